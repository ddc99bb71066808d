"""cProfile of TrainAugment.__call__ at 128 frames of 576x720 (host side only matters: where do the milliseconds go)."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from ssl4polyp_b200.input_pipeline import TrainAugment

rng = np.random.default_rng(0)
frames = [rng.integers(0, 256, (576, 720, 3), dtype=np.uint8) for _ in range(128)]
tf = TrainAugment(224)
out = torch.empty(128, 3, 224, 224, device="cuda")
for _ in range(3):
    tf(frames, out=out)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    tf(frames, out=out)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
