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


# =====================================================================================================================
# Fine-tune TRAIN augmentations (classification/data/transforms.py:233-246), applied to the PIL image AFTER
# T.Resize((S, S)):  ColorJitter(0.4, 0.5, 0.25, 0.01) -> GaussianBlur((25, 25), sigma=(0.001, 2.0)) -> RandomHorizontalFlip
# -> RandomVerticalFlip -> RandomRotation(180) -> ToTensor -> Normalize.
# The arithmetic is Pillow's (ImageEnhance = Image.blend with a degenerate image, libImaging/Blend.c, Convert.c rgb2l /
# rgb2hsv_row / hsv2rgb, Image.rotate -> Geometry.c affine_fixed for NEAREST) and torchvision's (ColorJitter.get_params
# order, adjust_hue's uint8 wrap-around shift, gaussian_blur = float32 conv2d of the uint8 image with a reflect pad, then
# round-half-even).  Pinned against the INSTALLED libraries by oracle/make_augment_golden.py: every stage bit for bit
# (the two HSV conversions over ALL 2^24 inputs) EXCEPT the blur, whose float32 summation order inside torch's CPU conv2d is
# not specified: the restatement sums in float64 and may differ from torch by one grey level where the exact value lies
# within ~1e-5 of a rounding boundary (a few values per image; the fixture records how many).
# =====================================================================================================================
def rgb_to_l(img: np.ndarray) -> np.ndarray:
    """Convert.c rgb2l (ITU-R 601-2, 16-bit fixed point): L = (19595 R + 38470 G + 7471 B + 0x8000) >> 16."""
    r, g, b = (img[..., i].astype(np.int64) for i in range(3))
    return ((r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16).astype(np.uint8)


def blend_u8(a: np.ndarray, b: np.ndarray, alpha: float) -> np.ndarray:
    """Blend.c ImagingBlend(a, b, (float)alpha): a + alpha (b - a) in float32 (a multiply, then an add), truncated;
    clipped to [0, 255] when alpha extrapolates."""
    al = np.float32(alpha)
    a32, b32 = a.astype(np.float32), b.astype(np.float32)
    t = (a32 + (al * (b32 - a32)).astype(np.float32)).astype(np.float32)
    if np.float32(0.0) <= al <= np.float32(1.0):                          # the C function takes alpha as a float
        return t.astype(np.uint8)
    return np.where(t <= 0, 0, np.where(t >= 255, 255, np.floor(t))).astype(np.uint8)


def adjust_brightness(img, f):      # ImageEnhance.Brightness: degenerate = black
    return blend_u8(np.zeros_like(img), img, f)


def adjust_contrast(img, f):        # ImageEnhance.Contrast: degenerate = int(mean(L) + 0.5) everywhere
    L = rgb_to_l(img)
    mean = int(int(L.astype(np.int64).sum()) / L.size + 0.5)
    return blend_u8(np.full_like(img, mean), img, f)


def adjust_saturation(img, f):      # ImageEnhance.Color: degenerate = L replicated to three bands
    return blend_u8(rgb_to_l(img)[..., None].repeat(3, axis=2), img, f)


def rgb2hsv_u8(rgb: np.ndarray) -> np.ndarray:
    """Convert.c rgb2hsv_row: float32 quotients, the hue assembled in double, stored to float32, scaled in double, truncated."""
    f32, f64 = np.float32, np.float64
    r, g, b = (rgb[..., i].astype(np.int32) for i in range(3))
    maxc, minc = np.maximum(r, np.maximum(g, b)), np.minimum(r, np.minimum(g, b))
    cr = (maxc - minc).astype(f32)
    with np.errstate(divide="ignore", invalid="ignore"):
        s = cr / maxc.astype(f32)
        rc, gc, bc = ((maxc - c).astype(f32) / cr for c in (r, g, b))
        h = np.where(r == maxc, (bc - gc).astype(f64),
                     np.where(g == maxc, (2.0 + rc.astype(f64)) - bc.astype(f64), (4.0 + gc.astype(f64)) - rc.astype(f64))).astype(f32)
        h = np.fmod(h.astype(f64) / 6.0 + 1.0, 1.0).astype(f32)
        uh = np.clip((h.astype(f64) * 255.0).astype(np.int64), 0, 255)
        us = np.clip((s.astype(f64) * 255.0).astype(np.int64), 0, 255)
    gray = maxc == minc
    return np.stack([np.where(gray, 0, uh), np.where(gray, 0, us), maxc], -1).astype(np.uint8)


def hsv2rgb_u8(hsv: np.ndarray) -> np.ndarray:
    """Convert.c hsv2rgb: sextant i = floor(h 6 / 255) and remainder f in double -> float32, p/q/t = round(v (1 - ...)) in double."""
    f32, f64 = np.float32, np.float64
    h6 = hsv[..., 0].astype(f64) * 6.0 / 255.0
    i = np.floor(h6).astype(np.int64)
    f = (h6 - i.astype(f64)).astype(f32).astype(f64)
    fs = (hsv[..., 1].astype(f64) / 255.0).astype(f32).astype(f64)
    v = hsv[..., 2].astype(f64)
    rnd = lambda x: np.floor(x + 0.5).astype(np.int64)                # C round() of a non-negative value
    p, q, t = (np.clip(rnd(v * (1.0 - e)), 0, 255) for e in (fs, fs * f, fs * (1.0 - f)))
    vi = hsv[..., 2].astype(np.int64)
    k = [i % 6 == n for n in range(6)]
    R = np.select(k, [vi, q, p, p, t, vi]); G = np.select(k, [t, vi, vi, q, p, p]); B = np.select(k, [p, p, t, vi, vi, q])
    gray = hsv[..., 1] == 0
    return np.stack([np.where(gray, vi, R), np.where(gray, vi, G), np.where(gray, vi, B)], -1).astype(np.uint8)


def hue_shift_u8(hue_factor: float) -> int:
    """torchvision _functional_pil.adjust_hue: np.int32(hue_factor * 255).astype(np.uint8) (truncation toward zero, wrap)."""
    return int(np.int32(hue_factor * 255).astype(np.uint8))


def adjust_hue(img, hue_factor):    # RGB -> HSV, h += shift (uint8 wrap-around), HSV -> RGB; lossy even for a zero shift
    hsv = rgb2hsv_u8(img)
    hsv[..., 0] = hsv[..., 0] + np.uint8(hue_shift_u8(hue_factor))
    return hsv2rgb_u8(hsv)


def color_jitter(img, order, b, c, s, h):
    """torchvision ColorJitter.forward with get_params' (fn_idx, brightness, contrast, saturation, hue)."""
    for fn in order:
        img = (adjust_brightness(img, b) if fn == 0 else adjust_contrast(img, c) if fn == 1 else
               adjust_saturation(img, s) if fn == 2 else adjust_hue(img, h))
    return img


def gaussian_kernel1d(ksize: int, sigma: float) -> np.ndarray:
    """torchvision _get_gaussian_kernel1d, the same float32 torch CPU ops call for call (the taps are host-side scalars)."""
    import torch
    half = (ksize - 1) * 0.5
    x = torch.linspace(-half, half, steps=ksize, dtype=torch.float32)
    pdf = torch.exp(-0.5 * (x / sigma).pow(2))
    return (pdf / pdf.sum()).numpy()


def gaussian_blur_u8(img: np.ndarray, k1d: np.ndarray) -> np.ndarray:
    """F.gaussian_blur on a uint8 image: reflect pad, the separable kernel (ky kx^T, the same sigma on both axes), float
    accumulate, round half to even, uint8.  Summed in float64 here (see the header of this section)."""
    r = len(k1d) // 2
    x = np.pad(img.astype(np.float64), ((r, r), (r, r), (0, 0)), mode="reflect")
    k = k1d.astype(np.float64)
    H, W = img.shape[:2]
    tmp = sum(k[j] * x[:, j:j + W, :] for j in range(len(k)))
    out = sum(k[i] * tmp[i:i + H, :, :] for i in range(len(k)))
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)


def rotate_coeffs(angle: float, w: int, h: int):
    """Image.rotate(angle, NEAREST, expand=False, center=None) -> (mode, a0..a5): mode 0 = Geometry.c affine_fixed with the
    16.16 fixed-point coefficients (x_in = (a2 + a0 x + a1 y) >> 16, y_in = (a5 + a3 x + a4 y) >> 16), 1 = identity,
    2 = rotate 180, 3 = rotate 90 (counter-clockwise), 4 = rotate 270; modes 3/4 only for square images."""
    angle = angle % 360.0
    if angle == 0:
        return 1, (0,) * 6
    if angle == 180:
        return 2, (0,) * 6
    if angle in (90, 270) and w == h:
        return (3 if angle == 90 else 4), (0,) * 6
    cx, cy = w / 2.0, h / 2.0
    a = -math.radians(angle)
    m = [round(math.cos(a), 15), round(math.sin(a), 15), 0.0, round(-math.sin(a), 15), round(math.cos(a), 15), 0.0]
    m[2] = m[0] * (-cx) + m[1] * (-cy) + m[2]
    m[5] = m[3] * (-cx) + m[4] * (-cy) + m[5]
    m[2] += cx
    m[5] += cy
    fix = lambda v: int(math.floor(v * 65536.0 + 0.5))
    return 0, (fix(m[0]), fix(m[1]), fix(m[2] + m[0] * 0.5 + m[1] * 0.5), fix(m[3]), fix(m[4]), fix(m[5] + m[3] * 0.5 + m[4] * 0.5))


def rotate_nearest(img: np.ndarray, angle: float) -> np.ndarray:
    h, w = img.shape[:2]
    mode, (a0, a1, a2, a3, a4, a5) = rotate_coeffs(angle, w, h)
    if mode == 1:
        return img.copy()
    if mode == 2:
        return img[::-1, ::-1].copy()
    if mode in (3, 4):
        return np.rot90(img, 1 if mode == 3 else 3).copy()
    ys, xs = np.mgrid[0:h, 0:w].astype(np.int64)
    xin, yin = (a2 + a0 * xs + a1 * ys) >> 16, (a5 + a3 * xs + a4 * ys) >> 16
    ok = (xin >= 0) & (xin < w) & (yin >= 0) & (yin < h)
    out = np.zeros_like(img)
    out[ok] = img[yin[ok], xin[ok]]
    return out


def draw_augment_params(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01, sigma=(0.001, 2.0), degrees=180.0):
    """The random draws of the augmentation Compose for ONE image, global torch RNG, call for call:
    ColorJitter.get_params (randperm(4), four uniforms), GaussianBlur.get_params, two flips (torch.rand(1) < 0.5),
    RandomRotation.get_params."""
    import torch
    order = [int(v) for v in torch.randperm(4)]
    b = float(torch.empty(1).uniform_(max(0.0, 1 - brightness), 1 + brightness))
    c = float(torch.empty(1).uniform_(max(0.0, 1 - contrast), 1 + contrast))
    s = float(torch.empty(1).uniform_(max(0.0, 1 - saturation), 1 + saturation))
    h = float(torch.empty(1).uniform_(-hue, hue))
    sg = torch.empty(1).uniform_(sigma[0], sigma[1]).item()
    hflip = bool(torch.rand(1) < 0.5)
    vflip = bool(torch.rand(1) < 0.5)
    angle = float(torch.empty(1).uniform_(-degrees, degrees).item())
    return dict(order=order, brightness=b, contrast=c, saturation=s, hue=h, sigma=sg, hflip=hflip, vflip=vflip, angle=angle)


def augment_u8(img: np.ndarray, p: dict, ksize: int = 25) -> np.ndarray:
    """ColorJitter -> GaussianBlur -> flips -> rotation on the resized uint8 image."""
    img = color_jitter(img, p["order"], p["brightness"], p["contrast"], p["saturation"], p["hue"])
    img = gaussian_blur_u8(img, gaussian_kernel1d(ksize, p["sigma"]))
    if p["hflip"]:
        img = img[:, ::-1]
    if p["vflip"]:
        img = img[::-1]
    return rotate_nearest(np.ascontiguousarray(img), p["angle"])


def transform_finetune_train(frame: np.ndarray, size: int = 224):
    """classification/data/transforms.py:233-258, stage 'train': Resize (PIL bilinear) -> augmentations -> ToTensor -> Normalize,
    global torch RNG.  Returns (tensor [3,size,size] float32, uint8 image before ToTensor, params)."""
    img = crop_resize_u8(frame, 0, 0, frame.shape[0], frame.shape[1], size, "bilinear")
    p = draw_augment_params()
    u8 = augment_u8(img, p)
    return to_tensor_normalize(u8, False), u8, p
