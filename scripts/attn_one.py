"""One forward + backward of the attention kernels at a single shape (ncu captures).
    python scripts/attn_one.py [B T H hd]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4polyp_b200 import ops
B, T, H, hd = (int(x) for x in sys.argv[1:5]) if len(sys.argv) >= 5 else (256, 197, 16, 32)
dev = torch.device("cuda:0")
qkv = torch.randn(B * T, 3 * H * hd, device=dev).bfloat16()
do = torch.randn(B * T, H * hd, device=dev).bfloat16()
for _ in range(2):
    out, lse = ops.attention_fwd(qkv, B, T, H, hd, hd ** -0.5)
    dqkv = ops.attention_bwd(qkv, out, do, lse, B, T, H, hd, hd ** -0.5)
torch.cuda.synchronize()
print("ok")
