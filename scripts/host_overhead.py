"""How long does the host take to ENQUEUE one step vs how long the GPU takes to run it?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4polyp_b200 import models_mae
from ssl4polyp_b200.engine import PretrainEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = models_mae.mae_vit_base_patch16(norm_pix_loss=True).to(dev)
eng = PretrainEngine(m, lr=1e-4)
imgs = torch.randn(B, 3, 224, 224, device=dev)
for _ in range(3):
    eng.step(imgs)
torch.cuda.synchronize()
for trial in range(3):
    t0 = time.perf_counter()
    for _ in range(5):
        eng.step(imgs)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"B={B}: enqueue {1e3*(t1-t0)/5:.2f} ms/step, total {1e3*(t2-t0)/5:.2f} ms/step")
