"""Time the attention kernels alone at the shapes of the MAE ViT-B step (CUDA events, L2 flushed between calls).
    python scripts/attn_bench.py [reps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4polyp_b200 import ops

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for (B, T, H, hd) in [(256, 197, 16, 32), (256, 50, 12, 64), (128, 197, 12, 64)]:
    qkv = torch.randn(B * T, 3 * H * hd, device=dev).bfloat16()
    do = torch.randn(B * T, H * hd, device=dev).bfloat16()
    for impl in ("tcgen05",):
        tf, tb = [], []
        for _ in range(reps):
            flush.zero_()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            out, lse = ops.attention_fwd(qkv, B, T, H, hd, hd ** -0.5)
            e[1].record()
            dqkv = ops.attention_bwd(qkv, out, do, lse, B, T, H, hd, hd ** -0.5)
            e[2].record()
            torch.cuda.synchronize()
            tf.append(e[0].elapsed_time(e[1]) * 1e3); tb.append(e[1].elapsed_time(e[2]) * 1e3)
        fl = 4.0 * B * H * T * T * hd
        byt_f = B * T * H * hd * 2 * 4
        print(f"B={B} T={T} H={H} hd={hd} {impl:8s} fwd {min(tf):8.1f} us ({fl / min(tf) / 1e6:6.1f} TF/s, {byt_f / min(tf) / 1e3:6.0f} GB/s)  "
              f"bwd {min(tb):8.1f} us ({2.5 * fl / min(tb) / 1e6:6.1f} TF/s, {byt_f * 2 / min(tb) / 1e3:6.0f} GB/s)")
