"""TEST INFRASTRUCTURE ONLY (checker for tests/, smoke and bench's cpu_baseline; never imported by the product).

CPU restatement of the pretraining input transform, reference call site models/mae/main_pretrain.py:156-160:

    RandomResizedCrop(224, scale=(0.2, 1.0), interpolation=3)   # 3 = PIL bicubic, on a PIL image
    RandomHorizontalFlip()
    ToTensor()
    Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])

The arithmetic lives in third-party code that is not under /root/reference:
  * Pillow (pinned 10.4.0, requirements.txt:3) `Image.crop` + `Image.resize(size, BICUBIC)` -> libImaging/Resample.c
    `ImagingResample` for 8-bit channels: two separable passes (horizontal first), coefficients of the a = -0.5 cubic
    with the support stretched by the down-scale factor, normalised, rounded to 22-bit fixed point; each pass
    accumulates in int32 from 1 << 21, shifts by 22 and clips to uint8 (so the intermediate image is uint8).
  * torchvision (pinned 0.10.0, requirements.txt:7) `RandomResizedCrop.get_params` (10 attempts of area in
    scale*area, log-uniform ratio in (3/4, 4/3), `torch.randint` offsets, centre-crop fallback), `hflip`,
    `to_tensor` (uint8 -> float32 / 255) and `normalize` ((x - mean) / std in float32).
Pinned HERE against the installed Pillow 12.2 / torchvision 0.26 (the resample algorithm and get_params are unchanged
between the pinned and installed versions) by oracle/make_input_golden.py -> tests/golden/input_golden.npz; tests
require bit-exact uint8 after the resample and bit-exact float32 after normalisation.
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2          # Resample.c
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def bicubic_filter(x: float) -> float:
    """Resample.c bicubic_filter, a = -0.5, evaluated in double exactly as written there."""
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def bilinear_filter(x: float) -> float:
    """Resample.c bilinear_filter (support 1.0)."""
    if x < 0.0:
        x = -x
    return 1.0 - x if x < 1.0 else 0.0


def precompute_coeffs(in_size: int, in0: float, in1: float, out_size: int, filt: str = "bicubic"):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc -> (ksize, bounds[out,2] = (xmin, count), kk[out,ksize] int32)."""
    scale = filterscale = (in1 - in0) / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    fn, base = (bicubic_filter, 2.0) if filt == "bicubic" else (bilinear_filter, 1.0)
    support = base * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = in0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        k = [fn((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for w in k:
            ww += w
        for x in range(xmax):
            v = k[x] / ww if ww != 0.0 else k[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _pass(img: np.ndarray, bounds, kk, axis: int) -> np.ndarray:
    """One separable pass over uint8 [H,W,C]; axis 1 = horizontal, 0 = vertical."""
    src = img.astype(np.int64)
    n = bounds.shape[0]
    shape = list(img.shape)
    shape[axis] = n
    out = np.empty(shape, np.uint8)
    for o in range(n):
        lo, cnt = int(bounds[o, 0]), int(bounds[o, 1])
        w = kk[o, :cnt].astype(np.int64)
        if axis == 1:
            acc = (src[:, lo:lo + cnt, :] * w[None, :, None]).sum(axis=1)
        else:
            acc = (src[lo:lo + cnt, :, :] * w[:, None, None]).sum(axis=0)
        acc = (acc + (1 << (PRECISION_BITS - 1))) >> PRECISION_BITS
        val = np.clip(acc, 0, 255).astype(np.uint8)
        if axis == 1:
            out[:, o, :] = val
        else:
            out[o, :, :] = val
    return out


def crop_resize_u8(frame: np.ndarray, top: int, left: int, height: int, width: int, size: int = 224, filt: str = "bicubic") -> np.ndarray:
    """`F.resized_crop(img, top, left, height, width, (size, size), BICUBIC)` on a PIL RGB image: crop, then
    ImagingResample with the box = the whole crop; a pass is skipped when that axis already has the target size."""
    crop = np.ascontiguousarray(frame[top:top + height, left:left + width, :])
    if width != size:
        _, b, k = precompute_coeffs(width, 0.0, float(width), size, filt)
        crop = _pass(crop, b, k, 1)
    if height != size:
        _, b, k = precompute_coeffs(height, 0.0, float(height), size, filt)
        crop = _pass(crop, b, k, 0)
    return crop


def to_tensor_normalize(img_u8: np.ndarray, flip: bool, mean=IMAGENET_MEAN, std=IMAGENET_STD) -> np.ndarray:
    """hflip -> to_tensor (float32 / 255, CHW) -> normalize ((x - mean) / std, all float32)."""
    if flip:
        img_u8 = img_u8[:, ::-1, :]
    x = img_u8.transpose(2, 0, 1).astype(np.float32) / np.float32(255.0)
    m = np.asarray(mean, np.float32)[:, None, None]
    s = np.asarray(std, np.float32)[:, None, None]
    return ((x - m) / s).astype(np.float32)


def random_resized_crop_params(height: int, width: int, scale=(0.2, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)):
    """torchvision RandomResizedCrop.get_params, consuming the global torch RNG call for call."""
    import torch
    area = height * width
    log_ratio = torch.log(torch.tensor(ratio))
    for _ in range(10):
        target_area = area * torch.empty(1).uniform_(scale[0], scale[1]).item()
        aspect = torch.exp(torch.empty(1).uniform_(log_ratio[0], log_ratio[1])).item()
        w = int(round(math.sqrt(target_area * aspect)))
        h = int(round(math.sqrt(target_area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            i = torch.randint(0, height - h + 1, size=(1,)).item()
            j = torch.randint(0, width - w + 1, size=(1,)).item()
            return i, j, h, w
    in_ratio = float(width) / float(height)
    if in_ratio < min(ratio):
        w = width
        h = int(round(w / min(ratio)))
    elif in_ratio > max(ratio):
        h = height
        w = int(round(h * max(ratio)))
    else:
        w, h = width, height
    return (height - h) // 2, (width - w) // 2, h, w


def transform_train(frame: np.ndarray, size: int = 224):
    """The whole Compose on one uint8 HWC frame with the global torch RNG; returns (tensor [3,size,size], params)."""
    import torch
    i, j, h, w = random_resized_crop_params(frame.shape[0], frame.shape[1])
    flip = bool(torch.rand(1) < 0.5)
    return to_tensor_normalize(crop_resize_u8(frame, i, j, h, w, size), flip), (i, j, h, w, flip)


def synthetic_frame(seed: int, H: int, W: int) -> np.ndarray:
    """Seeded uint8 HWC test frame (8x8 blocks xor 5-bit noise: hard edges and texture); PCG64 streams are identical on
    every machine, so fixtures store only what PIL produced from the frame, not the frame."""
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (H // 8 + 2, W // 8 + 2, 3), dtype=np.uint8)
    return np.kron(base, np.ones((8, 8, 1), np.uint8))[:H, :W] ^ rng.integers(0, 32, (H, W, 3), dtype=np.uint8)


def transform_eval(frame: np.ndarray, size: int = 224) -> np.ndarray:
    """classification/data/transforms.py:233-258 without augmentations (val / test stages): T.Resize((size, size)) - PIL
    BILINEAR over the whole frame - then ToTensor and Normalize."""
    return to_tensor_normalize(crop_resize_u8(frame, 0, 0, frame.shape[0], frame.shape[1], size, "bilinear"), False)
