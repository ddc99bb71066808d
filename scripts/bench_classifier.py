"""BASELINE.json configs[3]: ViT-B/16 classification fine-tune step (random-init MAE-shaped weights, 2-class polyp head,
BCE on logit1-logit0 as train_classification.py:3347-3374, AdamW, full fine-tune) on synthetic 224x224 batches.

    python scripts/bench_classifier.py [--batch 128] [--steps 20] [--warmup 5] [--backbone mae|any]
    torchrun --nproc-per-node N scripts/bench_classifier.py --batch 128          (one process per GPU, batch per GPU)

Same timing rules as bench.py: CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks,
whole step (fwd + loss + bwd + bucketed all-reduce + AdamW) replayed from one CUDA graph, two alternating input batches.
F_step = 105.378 GFLOP/img.  Prints one JSON line.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

F_STEP_GFLOP = 105.378


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--backbone", default="mae", choices=["mae", "any"])
    args = ap.parse_args()
    from ssl4polyp_b200 import utils
    from ssl4polyp_b200.engine import ClassifierEngine
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    if args.backbone == "mae":
        model = utils.get_MAE_backbone(None, True, 2, False, None, out_token="cls")
    else:
        model = utils.get_ImageNet_or_random_ViT(True, 2, False, None, False, out_token="cls")
    model = model.to(dev).train()
    eng = ClassifierEngine(model, lr=1e-5, weight_decay=0.05)
    B = args.batch
    g = torch.Generator().manual_seed(1000 + rank)
    imgs = [torch.randn(B, 3, 224, 224, generator=g).to(dev) for _ in range(2)]
    tgt = [torch.randint(0, 2, (B,), generator=torch.Generator().manual_seed(2000 + rank + i)).to(dev) for i in range(2)]
    for i in range(max(args.warmup, 3)):
        eng.step(imgs[i & 1], tgt[i & 1], graph=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = eng.step(imgs[i & 1], tgt[i & 1], graph=True)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = ms.item()
    if rank == 0:
        ips = B * world / (ms / 1e3)
        print(json.dumps({"metric": "ViT-B/16 classification fine-tune img/s", "value": ips, "unit": "images/s", "n_gpus": world,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                          "dtype": "bf16", "data": "synthetic",
                          "config": {"workload": f"ViT-B/16 ({'ViT_from_MAE' if args.backbone == 'mae' else 'VisionTransformer_from_Any'}) "
                                                 f"full fine-tune step, 2-class BCE, batch {B}/GPU, 224x224 fp32 images", "global_batch": B * world},
                          "step_tflops_per_gpu": F_STEP_GFLOP * 1e9 * B / (ms / 1e3) / 1e12, "loss": float(loss.item())}))
    if world > 1:
        torch.cuda.synchronize()
        os._exit(0)          # see bench.py: ProcessGroupNCCL teardown with captured collectives


if __name__ == "__main__":
    main()
