"""HBM bandwidth by access mix (B200): fill (write only), copy (read+write), read-mostly reduction."""
import torch
dev = torch.device("cuda:0")
n = 512 << 20
a = torch.empty(n, dtype=torch.uint8, device=dev); b = torch.empty(n, dtype=torch.uint8, device=dev)
af = a.view(torch.float32); bf = b.view(torch.float32)
def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms = t(lambda: af.fill_(1.0)); print(f"fill  (write {n>>20} MiB): {ms*1e3:.1f} us  {n/ms/1e9:.2f} TB/s")
ms = t(lambda: bf.copy_(af)); print(f"copy  (read+write): {ms*1e3:.1f} us  {2*n/ms/1e9:.2f} TB/s")
ms = t(lambda: af.sum()); print(f"sum   (read only): {ms*1e3:.1f} us  {n/ms/1e9:.2f} TB/s")
h = a.view(torch.bfloat16); o = b.view(torch.bfloat16)
ms = t(lambda: torch.mul(h, 2, out=o)); print(f"scale bf16 (read+write): {ms*1e3:.1f} us  {2*n/ms/1e9:.2f} TB/s")
