"""Sweep (splits, N tile) for every wgrad (TN) shape of the MAE ViT-B / ViT-L step: GEMM + split-K reduce, CUDA events,
L2 flushed between calls.  Prints the best pair per shape next to what ssl4p_gemm_pick_splits chooses.
    python scripts/wgrad_sweep.py [reps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4polyp_b200 import ops
from ssl4polyp_b200.ops import GEMM_TN, EPI_PARTIAL_F32

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
Me, Md = 256 * 50, 256 * 197
shapes = []
for tag, K, D in (("enc", Me, 768), ("dec", Md, 512), ("encL", Me, 1024), ("cls", 128 * 197, 768)):
    shapes += [(f"{tag} wfc1", 4 * D, D, K), (f"{tag} wfc2", D, 4 * D, K), (f"{tag} wqkv", 3 * D, D, K), (f"{tag} wproj", D, D, K)]
shapes += [("pe", 768, 768, 256 * 49), ("pred", 768, 512, Md), ("dembed", 512, 768, Me)]


def run(a, b, out, s, bn):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gemm(a, b, GEMM_TN, EPI_PARTIAL_F32, out=out, splits=s, force_bn=bn)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return min(ts)


for name, M, N, K in shapes:
    a = torch.randn(K, M, device=dev).bfloat16(); b = torch.randn(K, N, device=dev).bfloat16()
    out = torch.empty(M, N, device=dev)
    s0 = ops.pick_splits(M, N, K)
    t0 = run(a, b, out, s0, 0)
    res = []
    kb = (K + 63) // 64
    for bn in (128, 192, 256):
        for s in range(1, 17):
            if kb // s < 4:
                continue
            tiles = ((M + 127) // 128) * ((N + bn - 1) // bn) * s
            if tiles > 3 * 148 + 20:
                continue
            res.append((run(a, b, out, s, bn), s, bn, tiles))
    res.sort()
    fl = 2.0 * M * N * K
    print(f"{name:10s} M={M:5d} N={N:5d} K={K:6d} picked s={s0:2d}: {t0:7.1f} us ({fl / t0 / 1e6:6.0f} TF/s) | best " +
          "  ".join(f"s={s:2d} bn={bn} tiles={tl:3d}: {t:6.1f}us ({fl / t / 1e6:5.0f})" for t, s, bn, tl in res[:4]), flush=True)
