"""Timing experiments on the pipelined attention forward: SSL4P_ATTN_DEBUG bit mask disables one phase at a time
(1 no exp2, 2 one P.V k-step, 4 no TMA after the ring is primed, 8 no P stores, 16 no O stores)."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import torch
    from ssl4polyp_b200 import ops
    dev = torch.device("cuda:0")
    for (B, T, H, hd) in [(256, 197, 16, 32), (128, 197, 12, 64), (256, 50, 12, 64)]:
        qkv = torch.randn(B * T, 3 * H * hd, device=dev).bfloat16()
        ts = []
        for _ in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.attention_fwd(qkv, B, T, H, hd, hd ** -0.5); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        print(f"  dbg={os.environ.get('SSL4P_ATTN_DEBUG', '0'):>3s} T={T} hd={hd}: {min(ts):7.1f} us")
        if int(os.environ.get('SSL4P_ATTN_DEBUG', '0')) & 64:
            out, lse = ops.attention_fwd(qkv, B, T, H, hd, hd ** -0.5)
            torch.cuda.synchronize()
            pc = lse.view(-1)[:20].view(torch.int64).cpu().tolist()
            names = ["wait s_full", "tmem ld+free", "max", "bar", "exp/pack", "wait o_full", "epilogue", "P store+arrive", "loop"]
            n = max(1, pc[9])
            print("     clocks/tile:", ", ".join(f"{nm} {pc[i] / n:.0f}" for i, nm in enumerate(names)), f"| tiles {pc[9]} | sum {sum(pc[:9]) / n:.0f}")
else:
    for m in (64, 64 + 31):
        subprocess.run([sys.executable, __file__, "run"], env=dict(os.environ, SSL4P_ATTN_DEBUG=str(m)))
