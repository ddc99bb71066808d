"""Pins the fine-tune TRAIN augmentation restatement (oracle/input_oracle.py) against the INSTALLED Pillow + torchvision and
writes tests/golden/augment_golden.npz (run in the build container; the GPU box only reads the committed file):

    python oracle/make_augment_golden.py

Reference: /root/reference/src/ssl4polyp/classification/data/transforms.py:233-258 (stage 'train').  Checked here:
  * ColorJitter's four operations one by one and as drawn by get_params (bit for bit), both HSV conversions over all 2^24
    inputs (bit for bit), Image.rotate NEAREST over random and special angles (bit for bit), the RNG call order of the whole
    Compose (the same parameters under one seed);
  * GaussianBlur((25, 25)): the restatement sums in float64, torch's CPU conv2d in float32 in an unspecified order -> at
    most one grey level apart, on a fraction of the values that the fixture records;
  * the whole stage-'train' transform on seeded frames; the fixture holds torchvision's uint8 result (before ToTensor)."""
import os
import sys

import numpy as np
import torch
from PIL import Image, ImageEnhance
import torchvision.transforms as T
import torchvision.transforms.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import input_oracle as IO  # noqa: E402

rng = np.random.default_rng(0)
# ---- HSV conversions, exhaustive
g = np.arange(256, dtype=np.uint8)
allrgb = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(4096, 4096, 3)
assert np.array_equal(IO.rgb2hsv_u8(allrgb), np.asarray(Image.fromarray(allrgb).convert("HSV")))
assert np.array_equal(IO.hsv2rgb_u8(allrgb), np.asarray(Image.fromarray(allrgb, mode="HSV").convert("RGB")))
print("rgb2hsv / hsv2rgb: all 2^24 inputs bit-identical")
# ---- the four colour operations
for t in range(60):
    h, w = int(rng.integers(8, 80)), int(rng.integers(8, 80))
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    if t % 3 == 0:
        img = (img // 4 + 100).astype(np.uint8)
    pil = Image.fromarray(img)
    f = float(rng.uniform(0.5, 1.5))
    assert np.array_equal(IO.adjust_brightness(img, f), np.asarray(F.adjust_brightness(pil, f)))
    assert np.array_equal(IO.adjust_contrast(img, f), np.asarray(F.adjust_contrast(pil, f)))
    assert np.array_equal(IO.adjust_saturation(img, f), np.asarray(F.adjust_saturation(pil, f)))
    hf = float(rng.uniform(-0.5, 0.5)) if t % 2 else float(rng.uniform(-0.01, 0.01))
    assert np.array_equal(IO.adjust_hue(img, hf), np.asarray(F.adjust_hue(pil, hf)))
print("brightness / contrast / saturation / hue: bit-identical")
# ---- rotation
for t in range(200):
    h, w = (224, 224) if t % 2 == 0 else (int(rng.integers(5, 60)), int(rng.integers(5, 60)))
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    ang = float(rng.uniform(-180, 180)) if t % 10 else float(rng.choice([0, 90, -90, 180, -180, 45, 270, 1e-9]))
    assert np.array_equal(IO.rotate_nearest(img, ang), np.asarray(F.rotate(Image.fromarray(img), ang))), (t, ang)
print("rotation (NEAREST): bit-identical")
# ---- blur: at most one grey level apart
worst_frac = 0.0
for t in range(12):
    img = IO.synthetic_frame(100 + t, 224, 224) if t % 2 else rng.integers(0, 256, (224, 224, 3), dtype=np.uint8)
    sigma = [0.001, 0.3, 0.7, 1.0, 1.5, 2.0][t % 6]
    ref = np.asarray(F.gaussian_blur(Image.fromarray(img), [25, 25], [sigma, sigma]))
    got = IO.gaussian_blur_u8(img, IO.gaussian_kernel1d(25, sigma))
    d = np.abs(ref.astype(int) - got.astype(int))
    assert d.max() <= 1, (sigma, d.max())
    worst_frac = max(worst_frac, float((d > 0).mean()))
print(f"blur: <= 1 grey level apart, on at most {worst_frac:.2e} of the values")
assert worst_frac < 1e-3
# ---- the whole stage-'train' transform under one seed
tf_pre = T.Compose([T.Resize((224, 224)), T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01),
                    T.GaussianBlur((25, 25), sigma=(0.001, 2.0)), T.RandomHorizontalFlip(), T.RandomVerticalFlip(), T.RandomRotation(180)])
tf_post = T.Compose([T.ToTensor(), T.Normalize(mean=IO.IMAGENET_MEAN, std=IO.IMAGENET_STD)])
out = {}
CASES = [(0, 288, 352), (1, 576, 720), (4, 100, 130), (6, 37, 41), (8, 224, 224), (9, 480, 640)]
torch.manual_seed(321)
refs = []
for seed, H, W in CASES:
    pil = tf_pre(Image.fromarray(IO.synthetic_frame(seed, H, W)))
    refs.append((np.asarray(pil), tf_post(pil).numpy()))
torch.manual_seed(321)
tot_bad = 0
for (seed, H, W), (ref_u8, ref_t) in zip(CASES, refs):
    t, u8, p = IO.transform_finetune_train(IO.synthetic_frame(seed, H, W))
    d = np.abs(u8.astype(int) - ref_u8.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3, (seed, d.max(), (d > 0).mean())
    tot_bad += int((d > 0).sum())
    # ToTensor + Normalize of the SAME uint8 image is bit-identical
    assert np.array_equal(IO.to_tensor_normalize(ref_u8, False).view(np.uint32), ref_t.view(np.uint32))
    out[f"u8_{seed}"] = ref_u8
    out[f"frame_{seed}"] = np.array([H, W], np.int32)
    out[f"params_{seed}"] = np.array(p["order"] + [IO.hue_shift_u8(p["hue"]), int(p["hflip"]), int(p["vflip"])], np.int32)
    out[f"fparams_{seed}"] = np.array([p["brightness"], p["contrast"], p["saturation"], p["hue"], p["sigma"], p["angle"]], np.float64)
print(f"stage 'train' under torch.manual_seed(321): same draws; uint8 results differ from torchvision on {tot_bad} values (blur rounding), <= 1 level")
out["seed"] = np.array([321], np.int32)
out["blur_mismatch_values"] = np.array([tot_bad], np.int64)
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "augment_golden.npz")
np.savez_compressed(dst, **out)
print("wrote", dst, os.path.getsize(dst) // 1024, "KiB")
