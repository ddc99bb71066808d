"""TEST INFRASTRUCTURE ONLY — pins the CLASSIFIER half of the oracle against outputs of the reference itself.

Run in the build container (the only place `/root/reference` exists):

    python oracle/make_cls_golden.py        # writes tests/golden/cls_golden.pt, prints the agreement table

What it does
  1. installs `oracle/timm_shim.py` and imports, UNMODIFIED,
       /root/reference/src/ssl4polyp/models/models.py                       (`ViT_from_MAE` :143-222,
                                                                             `VisionTransformer_from_Any` :26-140)
       /root/reference/src/ssl4polyp/classification/train_classification.py (`_compute_supervised_loss` :3361-3374,
                                                                             `_binary_logits_from_multiclass` :3347-3358)
       /root/reference/src/ssl4polyp/models/mae/util/lr_sched.py            (`adjust_learning_rate` :9-21)
  2. constructs each reference classifier under `torch.manual_seed(seed)`, runs logits -> loss -> backward with the
     reference's own loss glue (`nn.BCEWithLogitsLoss(pos_weight=neg/pos)` as built at :6090-6101,
     `nn.CrossEntropyLoss(weight=...)` at :6103) and REQUIRES `oracle.mae_oracle.classifier_forward`,
     `binary_polyp_loss` on the same state dict to agree: logits 1e-5, loss 1e-6, every gradient 1e-3 (relative L2);
  3. sweeps the pretraining lr schedule and REQUIRES `oracle.mae_oracle.cosine_lr` == the reference function;
  4. commits compact vectors (seeds, logits, losses, sub-sampled gradients + gradient norms, the lr sweep) to
     `tests/golden/cls_golden.pt`.  Weights are not stored for the MAE-backed model (they are regenerated from the seed
     by `oracle.mae_oracle.init_state_dict`, which make_golden.py pins bit-for-bit); for `VisionTransformer_from_Any`
     (trunc_normal init through the shim) the CPU/GPU tests rebuild it with the same seeded draws and the fixture keeps
     a weight sample to prove the rebuild is the same model.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import mae_oracle as O          # noqa: E402
from oracle.make_golden import GOLDEN_DIR, REF_SRC, import_reference_models_mae, subsample, synth_imgs   # noqa: E402

CLS_GRAD_KEYS = ["lin_head.weight", "lin_head.bias", "patch_embed.proj.weight", "cls_token", "blocks.0.attn.qkv.weight",
                 "blocks.5.norm1.weight", "blocks.6.mlp.fc1.bias", "blocks.11.mlp.fc2.weight", "norm.weight", "pos_embed"]


def import_reference_classifier_modules():
    import_reference_models_mae()                                     # shim + stub packages + models_mae
    models = importlib.import_module("ssl4polyp.models.models")
    tc = importlib.import_module("ssl4polyp.classification.train_classification")
    lr_sched = importlib.import_module("ssl4polyp.models.mae.util.lr_sched")
    return models, tc, lr_sched


def rel(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def run_case(models, tc, kind, out_token, B, seed, pos_weight, num_classes=2):
    torch.manual_seed(seed)
    if kind == "mae":
        ref = models.ViT_from_MAE(None, True, num_classes, False, None, embed_dim=768, depth=12, num_heads=12, out_token=out_token)
    else:
        ref = models.VisionTransformer_from_Any(True, num_classes, False, None, 768, 12, 12, out_token)
    ref.train()
    ref_sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    # weights the tests can rebuild without the reference tree
    if kind == "mae":
        sd = O.init_state_dict(O.MAE_BASE, seed=seed, head_classes=num_classes)
    else:
        sd = O.vit_any_state_dict(seed, num_classes)
    assert set(sd) == set(ref_sd), set(sd) ^ set(ref_sd)
    for k in ref_sd:
        assert torch.equal(sd[k], ref_sd[k]), f"{kind}: rebuilt weights differ from the reference constructor at {k}"

    imgs = synth_imgs(B, seed)
    target = torch.randint(0, num_classes, (B,), generator=torch.Generator().manual_seed(2000 + seed))
    if num_classes == 2:
        loss_fn = nn.BCEWithLogitsLoss(pos_weight=torch.tensor(pos_weight, dtype=torch.float32))      # :6098-6100
        mode = "binary_bce"
        cw = None
    else:
        cw = torch.linspace(0.5, 1.5, num_classes)
        loss_fn = nn.CrossEntropyLoss(weight=cw)                                                       # :6103
        mode = "multiclass_ce"
    logits_r = ref(imgs)
    loss_r = tc._compute_supervised_loss(loss_fn, logits_r, target, mode=mode)
    loss_r.backward()
    grads_r = {k: p.grad.detach().clone() for k, p in ref.named_parameters() if p.grad is not None}

    frozen = ("pos_embed", "decoder_pos_embed") if kind == "mae" else ()
    sd_g = {k: v.clone().requires_grad_(k not in frozen) for k, v in sd.items()}
    logits_o = O.classifier_forward(sd_g, imgs, O.MAE_BASE, out_token, True)
    if num_classes == 2:
        loss_o = O.binary_polyp_loss(logits_o, target, pos_weight=torch.tensor(pos_weight))
    else:
        loss_o = O.multiclass_loss(logits_o, target, cw)
    loss_o.backward()
    e_logit = rel(logits_o.detach(), logits_r.detach())
    e_loss = abs(loss_o.item() - loss_r.item()) / abs(loss_r.item())
    assert set(grads_r) == {k for k, v in sd_g.items() if v.grad is not None}, "gradient key sets differ"
    e_grad = max(rel(sd_g[k].grad, grads_r[k]) for k in grads_r)
    assert e_logit < 1e-5 and e_loss < 1e-6 and e_grad < 1e-3, (e_logit, e_loss, e_grad)
    print(f"[{kind} out_token={out_token} B={B} seed={seed} ncls={num_classes} pos_weight={pos_weight}] loss={loss_r.item():.6f} "
          f"oracle-vs-ref: logits {e_logit:.2e} loss {e_loss:.2e} grad {e_grad:.2e} ({len(grads_r)} tensors)")
    keys = [k for k in CLS_GRAD_KEYS if k in grads_r]
    return {"kind": kind, "out_token": out_token, "B": B, "seed": seed, "num_classes": num_classes, "pos_weight": pos_weight,
            "class_weights": cw, "target": target, "logits": logits_r.detach().clone(), "loss": loss_r.detach().clone(),
            "grad_norm": {k: grads_r[k].norm().clone() for k in grads_r},
            "grad_sample": {k: subsample(grads_r[k], 256) for k in keys},
            "w_sample": subsample(ref_sd["blocks.0.attn.qkv.weight"], 16), "head_w": ref_sd["lin_head.weight"].clone()}


def lr_sweep(lr_sched):
    """util/lr_sched.py:9-21 over warm-up + decay, with and without `lr_scale` groups."""
    args = types.SimpleNamespace(lr=1.5e-4 * 2048 / 256, min_lr=1e-6, warmup_epochs=40, epochs=400)

    class Opt:
        def __init__(self):
            self.param_groups = [{"lr": 0.0}, {"lr": 0.0, "lr_scale": 0.65}]

    pts = [0.0, 0.5, 1.0, 13.37, 39.999, 40.0, 40.001, 100.25, 220.0, 399.0, 399.999]
    out = []
    for e in pts:
        opt = Opt()
        lr = lr_sched.adjust_learning_rate(opt, e, args)
        assert opt.param_groups[0]["lr"] == lr and opt.param_groups[1]["lr"] == lr * 0.65
        mine = O.cosine_lr(e, args.lr, args.min_lr, args.warmup_epochs, args.epochs)
        assert mine == lr, (e, mine, lr)
        out.append((e, lr))
    short = types.SimpleNamespace(lr=1e-3, min_lr=0.0, warmup_epochs=5, epochs=20)
    for e in [0.0, 2.5, 5.0, 12.5, 19.9]:
        lr = lr_sched.adjust_learning_rate(Opt(), e, short)
        assert O.cosine_lr(e, short.lr, short.min_lr, short.warmup_epochs, short.epochs) == lr
        out.append((e, lr))
    print(f"lr schedule: {len(out)} points, oracle == reference exactly")
    return {"args": dict(lr=args.lr, min_lr=args.min_lr, warmup_epochs=args.warmup_epochs, epochs=args.epochs),
            "short_args": dict(lr=short.lr, min_lr=short.min_lr, warmup_epochs=short.warmup_epochs, epochs=short.epochs),
            "points": out[:len(pts)], "short_points": out[len(pts):]}


def main():
    torch.set_num_threads(os.cpu_count() or 8)
    torch.use_deterministic_algorithms(True, warn_only=True)
    models, tc, lr_sched = import_reference_classifier_modules()
    # _binary_logits_from_multiclass (:3347-3358) on the shapes it accepts
    z = torch.randn(5, 2)
    assert torch.equal(tc._binary_logits_from_multiclass(z), z[:, 1] - z[:, 0])
    cases = [
        run_case(models, tc, "mae", "cls", 4, 11, 1.0),
        run_case(models, tc, "mae", "spatial", 4, 12, 2.5),
        run_case(models, tc, "any", "cls", 3, 13, 0.4),
        run_case(models, tc, "any", "spatial", 3, 14, 1.0),
        run_case(models, tc, "mae", "cls", 4, 15, None, num_classes=3),
    ]
    out = {"cases": cases, "lr": lr_sweep(lr_sched), "torch": str(torch.__version__), "generator": "oracle/make_cls_golden.py",
           "reference": "irconde/SSL4POLYP models/models.py + classification/train_classification.py:3347-3374 + "
                        "models/mae/util/lr_sched.py (all unmodified) behind the timm 0.4.12 shim"}
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, "cls_golden.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
