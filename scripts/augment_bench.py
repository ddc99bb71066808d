"""Throughput of the fine-tune TRAIN augmentations on the device (csrc/augment.cu + the bilinear resample) at the fine-tune
batch: 128 frames -> fp32 [128,3,224,224], per kernel and end to end from host frames, with the reference's torchvision
Compose on one core beside it.

    python scripts/augment_bench.py [--frame 576x720] [--batch 128] [--cpu-sample 16]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--frame", default="576x720")
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--cpu-sample", type=int, default=16)
args = ap.parse_args()
H, W = (int(v) for v in args.frame.split("x"))
from ssl4polyp_b200 import ops
from ssl4polyp_b200.input_pipeline import TrainAugment, IMAGENET_MEAN, IMAGENET_STD
rng = np.random.default_rng(0)
frames = [rng.integers(0, 256, (H, W, 3), dtype=np.uint8) for _ in range(args.batch)]
tf = TrainAugment(224)
torch.manual_seed(0)
aug = [tf.draw_augment() for _ in frames]
ints, taps = tf.pack_augment(aug)
ints, taps = ints.cuda(), taps.cuda()
img0 = torch.from_numpy(rng.integers(0, 256, (args.batch, 224, 224, 3), dtype=np.uint8)).cuda()
out = torch.empty(args.batch, 3, 224, 224, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn):
    ts = []
    for r in range(args.reps + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        if r >= 2:
            ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


jit = torch.empty_like(img0)
blur = torch.empty_like(img0)
per_kernel = {"color_jitter_ms": timed(lambda: ops.color_jitter_u8(img0, ints, out=jit)),
              "gaussian_blur_ms": timed(lambda: ops.gaussian_blur_u8(jit, taps, 25, out=blur)),
              "flip_rotate_normalize_ms": timed(lambda: ops.flip_rotate_normalize_u8(blur, ints, IMAGENET_MEAN, IMAGENET_STD, out=out))}
for r in range(3):
    tf(frames, augment=aug, out=out)
torch.cuda.synchronize()
t0 = time.perf_counter()
for r in range(20):
    tf(frames, augment=aug, out=out)
torch.cuda.synchronize()
e2e_ms = (time.perf_counter() - t0) / 20 * 1e3
for r in range(2):
    tf(frames, out=out)
torch.cuda.synchronize()
t0 = time.perf_counter()
for r in range(20):
    tf(frames, out=out)                                  # parameters drawn per call (9 torch RNG calls per image, torchvision's order)
torch.cuda.synchronize()
e2e_draw_ms = (time.perf_counter() - t0) / 20 * 1e3
from PIL import Image
import torchvision.transforms as T
comp = T.Compose([T.Resize((224, 224)), T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01), T.GaussianBlur((25, 25), sigma=(0.001, 2.0)),
                  T.RandomHorizontalFlip(), T.RandomVerticalFlip(), T.RandomRotation(180), T.ToTensor(), T.Normalize(IMAGENET_MEAN, IMAGENET_STD)])
torch.set_num_threads(1)
pil = [Image.fromarray(f) for f in frames[:args.cpu_sample]]
t0 = time.perf_counter()
for im in pil:
    comp(im)
cpu_ms = (time.perf_counter() - t0) / len(pil) * 1e3
dev_ms = sum(per_kernel.values())
print(json.dumps({"metric": "fine-tune train augmentation images/s", "frame": args.frame, "batch": args.batch, **per_kernel,
                  "device_ms_per_batch_after_resize": dev_ms, "value": args.batch / dev_ms * 1e3,
                  "e2e_ms_per_batch": e2e_ms, "e2e_images_per_s": args.batch / e2e_ms * 1e3,
                  "e2e_with_draws_ms_per_batch": e2e_draw_ms, "e2e_with_draws_images_per_s": args.batch / e2e_draw_ms * 1e3,
                  "pack_threads": tf._pack_threads, "h2d_bytes_per_batch": tf.h2d_bytes,
                  "cpu_baseline": {"kind": "reference (Pillow + torchvision Compose, stage 'train')", "cores": 1, "ms_per_image": cpu_ms,
                                   "images_per_s": 1e3 / cpu_ms, "sample": f"{len(pil)} frames"}}))
