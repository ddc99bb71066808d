"""Quick on-GPU sanity of the tcgen05 GEMM (all layouts / tile widths); prints relative errors."""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4polyp_b200 import ops

d = torch.device("cuda:0")
torch.manual_seed(0)
def rel(a, b): return ((a.float() - b.float()).norm() / b.float().norm()).item()
for (M, N, K) in [(128, 64, 64), (128, 256, 128), (300, 512, 512)]:
    a = torch.randn(M, K, device=d).bfloat16(); w = (torch.randn(N, K, device=d) / math.sqrt(K)).bfloat16()
    for bn in (64, 128, 192, 256):
        out = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, force_bn=bn); torch.cuda.synchronize()
        print("NT", M, N, K, bn, rel(out, a.float() @ w.float().t()), flush=True)
    wn = (torch.randn(K, N, device=d) / math.sqrt(K)).bfloat16()
    for bn in (64, 128, 192, 256):
        out = ops.gemm(a, wn, ops.GEMM_NN, ops.EPI_BIAS_BF16, force_bn=bn); torch.cuda.synchronize()
        print("NN", M, N, K, bn, rel(out, a.float() @ wn.float()), flush=True)
    at = torch.randn(K, M, device=d).bfloat16()
    for bn in (64, 128, 192, 256):
        out = ops.gemm(at, wn, ops.GEMM_TN, ops.EPI_PARTIAL_F32, force_bn=bn); torch.cuda.synchronize()
        print("TN", M, N, K, bn, rel(out, at.float().t() @ wn.float()), flush=True)
print("done")
