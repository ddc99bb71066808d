"""The pretraining input transform with the pixel work on the device.

Reference (models/mae/main_pretrain.py:156-160), per image on DataLoader workers:

    transforms.Compose([RandomResizedCrop(input_size, scale=(0.2, 1.0), interpolation=3),   # PIL bicubic
                        RandomHorizontalFlip(), ToTensor(), Normalize(IMAGENET_MEAN, IMAGENET_STD)])

Here the workers only decode to uint8 HWC; `TrainTransform` draws the crop box and flip of every image on the host
with the SAME torch RNG calls in the same order as the Compose (so a seeded run crops identically), copies only the
crop rectangle of each frame into one pinned staging buffer (one H2D per batch, 3 bytes per source pixel of the
crop - the CPU cost left on the host is a memcpy, not a bicubic resample), and `ops.crop_resize_normalize` produces the fp32 [N,3,224,224] batch - bit-identical to the Compose
(Pillow's fixed-point bicubic resample included; csrc/input_pipeline.cu).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import ops

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def random_resized_crop_params(height: int, width: int, scale=(0.2, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)):
    """(top, left, h, w) drawn as torchvision's RandomResizedCrop.get_params draws them: up to ten tries of
    area ~ U(scale) * H * W and aspect ~ exp(U(log ratio)), offsets by torch.randint; centre crop clamped to the
    ratio range when no try fits."""
    lo, hi = torch.log(torch.tensor(ratio)).tolist()
    full = height * width
    for _attempt in range(10):
        area = full * torch.empty(1).uniform_(scale[0], scale[1]).item()
        aspect = torch.exp(torch.empty(1).uniform_(lo, hi)).item()
        w, h = int(round(math.sqrt(area * aspect))), int(round(math.sqrt(area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            top = torch.randint(0, height - h + 1, size=(1,)).item()
            left = torch.randint(0, width - w + 1, size=(1,)).item()
            return top, left, h, w
    shape = float(width) / float(height)
    if shape < min(ratio):
        w, h = width, int(round(width / min(ratio)))
    elif shape > max(ratio):
        h, w = height, int(round(height * max(ratio)))
    else:
        w, h = width, height
    return (height - h) // 2, (width - w) // 2, h, w


class TrainTransform:
    """`batch = TrainTransform(224)(frames)`; frames: sequence of uint8 HWC RGB arrays / CPU tensors of any sizes.
    Returns fp32 [N,3,size,size] on `device`.  `last_params` holds the drawn (top, left, h, w, flip) per image."""

    def __init__(self, size: int = 224, scale=(0.2, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), flip_p: float = 0.5,
                 mean=IMAGENET_MEAN, std=IMAGENET_STD, device=None, pack_threads=None):
        if not torch.cuda.is_available():
            raise RuntimeError("ssl4polyp_b200.input_pipeline needs a CUDA device (sm_100a); there is no CPU fallback")
        self.size, self.scale, self.ratio, self.flip_p, self.mean, self.std = size, scale, ratio, flip_p, mean, std
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._stage = [None, None]          # two pinned staging buffers: batch i+1 is packed while batch i's copy drains
        self._events = [None, None]
        self._turn = 0
        self.filter = "bicubic"                          # RandomResizedCrop(interpolation=3)
        import os
        from concurrent.futures import ThreadPoolExecutor
        # The crop rectangles are copied into the staging buffer by ONE native call per batch (ssl4p_host_pack_rects, memcpy on
        # `pack_threads` host threads) issued from a helper thread: ctypes drops the interpreter lock for the whole call.
        nt = min(8, os.cpu_count() or 1) if pack_threads is None else int(pack_threads)
        self._pack_threads = max(1, nt)
        self._pack_exec = ThreadPoolExecutor(max_workers=1)
        self.last_params = []
        self.h2d_bytes = 0

    def draw(self, height: int, width: int):
        top, left, h, w = random_resized_crop_params(height, width, self.scale, self.ratio)
        flip = bool(torch.rand(1) < self.flip_p)          # RandomHorizontalFlip.forward
        return top, left, h, w, flip

    def _staging(self, nbytes: int):
        k = self._turn
        self._turn ^= 1
        if self._events[k] is not None:
            self._events[k].synchronize()                 # the copy that last used this buffer has left the host
        buf = self._stage[k]
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(max(nbytes, 1 << 22), dtype=torch.uint8).pin_memory()
            self._stage[k] = buf
        return k, buf

    def __call__(self, frames, params=None, out=None):
        dev, desc_dev, max_crop, n = self._upload(frames, params)
        if n == 0:
            return torch.empty(0, 3, self.size, self.size, dtype=torch.float32, device=self.device)
        return ops.crop_resize_normalize(dev, desc_dev, self.size, max_crop, self.mean, self.std, out=out, filter=self.filter)

    def _upload(self, frames, params=None, overlap=None):
        """Draw (or take) the crop parameters, pack the crop rectangles into a pinned buffer, one H2D copy.
        -> (device bytes, device descriptors int64 [N,8], largest crop extent, N).  `overlap`: a callable run on the calling
        thread while the native copy runs on its helper thread; its result is left in `self._overlap_result`."""
        frames = [f.numpy() if isinstance(f, torch.Tensor) else np.asarray(f) for f in frames]
        for f in frames:
            if f.dtype != np.uint8 or f.ndim != 3 or f.shape[2] != 3:
                raise RuntimeError(f"TrainTransform: frames must be uint8 [H, W, 3], got {f.dtype} {f.shape}")
        if params is None:
            params = [self.draw(f.shape[0], f.shape[1]) for f in frames]
        self.last_params = list(params)
        n = len(frames)
        rows = np.zeros((n, 8), np.int64)
        total = 0
        for i, (f, (top, left, h, w, flip)) in enumerate(zip(frames, params)):
            if not (0 <= top and 0 <= left and h > 0 and w > 0 and top + h <= f.shape[0] and left + w <= f.shape[1]):
                raise RuntimeError(f"TrainTransform: crop {(top, left, h, w)} outside a {f.shape[0]}x{f.shape[1]} frame")
            rows[i] = (total, w * 3, 0, 0, h, w, int(flip), 0)       # only the crop rectangle travels: it becomes the frame
            total += h * w * 3
        desc = torch.from_numpy(rows)
        k, stage = self._staging(max(total, 1))

        self._overlap_result = None
        if n > 0:
            frames = [f if (f.strides[2] == 1 and f.strides[1] == 3 and f.strides[0] >= 3 * f.shape[1]) else np.ascontiguousarray(f) for f in frames]
            src = np.array([f.__array_interface__["data"][0] + p_[0] * f.strides[0] + p_[1] * 3 for f, p_ in zip(frames, params)], np.uint64)
            pitch = np.array([f.strides[0] for f in frames], np.int64)
            nrows = np.array([p_[2] for p_ in params], np.int64)
            row_bytes = np.array([p_[3] * 3 for p_ in params], np.int64)
            offs = np.ascontiguousarray(rows[:, 0])
            lib = ops._lib.load()

            def copy_all():                              # the error text is thread-local: read it on the thread that made the call
                rc = lib.ssl4p_host_pack_rects(src.ctypes.data, pitch.ctypes.data, nrows.ctypes.data, row_bytes.ctypes.data,
                                               stage.data_ptr(), offs.ctypes.data, n, self._pack_threads)
                return rc, (lib.ssl4p_last_error().decode("utf-8", "replace") if rc else "")

            fut = self._pack_exec.submit(copy_all)
            try:
                if overlap is not None:                  # Python-bound work of the same batch runs beside the copies
                    self._overlap_result = overlap()
            finally:
                rc, msg = fut.result()
            if rc:
                raise RuntimeError(f"host_pack_rects failed (code {rc}): {msg}")
        elif overlap is not None:
            self._overlap_result = overlap()
        dev = torch.empty(max(total, 1), dtype=torch.uint8, device=self.device)
        dev.copy_(stage[:max(total, 1)], non_blocking=True)
        desc_dev = desc.pin_memory().to(self.device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._events[k] = ev
        self.h2d_bytes = total + desc.numel() * 8
        max_crop = max((max(p[2], p[3]) for p in params), default=1)
        return dev, desc_dev, max_crop, n


class EvalTransform(TrainTransform):
    """The fine-tune script's val / test transform (classification/data/transforms.py:233-258 with augmentations off):
    T.Resize((size, size)) - PIL BILINEAR over the whole frame - then ToTensor and Normalize; no randomness."""

    def __init__(self, size: int = 224, mean=IMAGENET_MEAN, std=IMAGENET_STD, device=None, pack_threads=None):
        super().__init__(size, mean=mean, std=std, device=device, pack_threads=pack_threads)
        self.filter = "bilinear"

    def draw(self, height: int, width: int):
        return 0, 0, height, width, False


_GAUSS_X: dict = {}


def _gaussian_kernel1d(ksize: int, sigma: float) -> torch.Tensor:
    """torchvision _get_gaussian_kernel1d, the same float32 CPU ops call for call (25 host-side scalars per image)."""
    x = _GAUSS_X.get(ksize)
    if x is None:
        half = (ksize - 1) * 0.5
        x = _GAUSS_X[ksize] = torch.linspace(-half, half, steps=ksize, dtype=torch.float32)
    pdf = torch.exp(-0.5 * (x / sigma).pow(2))
    return pdf / pdf.sum()


def _rotate_coeffs(angle: float, w: int, h: int):
    """PIL Image.rotate(angle, NEAREST, expand=False, center=None): the special angles PIL turns into transposes, else the
    affine matrix it builds in Python doubles (cos / sin rounded to 15 decimals, rotation about (w/2, h/2)) converted to the
    16.16 fixed point of Geometry.c affine_fixed.  -> (mode, a0..a5), see include/ssl4p.h."""
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


class TrainAugment(EvalTransform):
    """The fine-tune script's TRAIN transform (classification/data/transforms.py:233-258, stage 'train') with the pixel
    work on the device:

        T.Resize((S, S)) -> ColorJitter(0.4, 0.5, 0.25, 0.01) -> GaussianBlur((25, 25), sigma=(0.001, 2.0)) ->
        RandomHorizontalFlip -> RandomVerticalFlip -> RandomRotation(180) -> ToTensor -> Normalize

    `draw_augment()` makes the random draws of ONE image with the same torch RNG calls in the same order as that Compose
    (randperm(4) + four uniforms, the blur sigma, two flips, the angle), so a seeded run augments identically; the frames
    travel as uint8 (one H2D per batch) and five kernels do the rest (bilinear resample, colour chain, separable blur,
    flip + rotation gather with ToTensor / Normalize).  `last_augment` holds the drawn parameters per image."""

    def __init__(self, size: int = 224, brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01, kernel_size=25,
                 sigma=(0.001, 2.0), degrees=180.0, flip_p: float = 0.5, mean=IMAGENET_MEAN, std=IMAGENET_STD, device=None,
                 pack_threads=None):
        super().__init__(size, mean=mean, std=std, device=device, pack_threads=pack_threads)
        # Whole frames travel (1.2 MB each at 576x720).  Measured on the B200 host (16 cores, 128 frames): torch copy_ 4.6 ms,
        # 8 native threads 2.9 ms, beside which the parameter draws and their packing - Python-bound - run on the calling thread
        # (scripts/augment_host_probe.py; Python threads around numpy copies did not overlap with that work: the interpreter
        # lock changed hands once per frame).
        if kernel_size % 2 == 0 or kernel_size > ops.AUG_MAX_TAPS:
            raise RuntimeError(f"TrainAugment: kernel_size must be odd and <= {ops.AUG_MAX_TAPS}")
        self.brightness = (max(0.0, 1 - brightness), 1 + brightness)      # ColorJitter._check_input
        self.contrast = (max(0.0, 1 - contrast), 1 + contrast)
        self.saturation = (max(0.0, 1 - saturation), 1 + saturation)
        self.hue = (-hue, hue)
        self.kernel_size, self.sigma, self.degrees, self.flip_p = kernel_size, sigma, float(degrees), flip_p
        self.last_augment = []
        self._scratch: dict = {}

    def draw_augment(self) -> dict:
        # torchvision draws every scalar as torch.empty(1).uniform_(lo, hi) / torch.rand(1): a one-element tensor of the default
        # dtype filled from the global generator.  The same fill on one reused scratch tensor consumes the generator identically
        # and yields the same value (rand(1) is uniform_(0, 1)); it saves ~1400 tensor allocations per batch of 128.
        u = self._scratch.get(torch.get_default_dtype())
        if u is None:
            u = self._scratch[torch.get_default_dtype()] = torch.empty(1)
            self._flip_thr = float(torch.tensor(self.flip_p, dtype=u.dtype))      # `rand(1) < p` compares in the tensor's dtype
        order = torch.randperm(4).tolist()                                 # ColorJitter.get_params
        b = u.uniform_(self.brightness[0], self.brightness[1]).item()
        c = u.uniform_(self.contrast[0], self.contrast[1]).item()
        s = u.uniform_(self.saturation[0], self.saturation[1]).item()
        h = u.uniform_(self.hue[0], self.hue[1]).item()
        sg = u.uniform_(self.sigma[0], self.sigma[1]).item()               # GaussianBlur.get_params
        hflip = u.uniform_(0.0, 1.0).item() < self._flip_thr
        vflip = u.uniform_(0.0, 1.0).item() < self._flip_thr
        angle = u.uniform_(-self.degrees, self.degrees).item()             # RandomRotation.get_params
        return dict(order=order, brightness=b, contrast=c, saturation=s, hue=h, sigma=sg, hflip=hflip, vflip=vflip, angle=angle)

    def pack_augment(self, aug: list):
        """-> (int32 [N,32], float32 [N,32]) host tensors in the layout of include/ssl4p.h"""
        n = len(aug)
        ints = np.zeros((n, ops.AUG_PARAM_INTS), np.int32)
        taps = torch.zeros(n, ops.AUG_MAX_TAPS, dtype=torch.float32)
        if n == 0:
            return torch.from_numpy(ints), taps
        S = self.size
        ints[:, 0:4] = [a["order"] for a in aug]
        # adjust_hue: hue * 255 in doubles, truncation toward zero, uint8 wrap
        ints[:, 4] = (np.array([a["hue"] for a in aug], np.float64) * 255).astype(np.int32).astype(np.uint8)
        ints[:, 5] = [int(a["hflip"]) for a in aug]
        ints[:, 6] = [int(a["vflip"]) for a in aug]
        rot = [_rotate_coeffs(a["angle"], S, S) for a in aug]
        ints[:, 7] = [m for m, _ in rot]
        ints[:, 8:14] = [c for _, c in rot]
        ints[:, 16:19] = np.array([(a["brightness"], a["contrast"], a["saturation"]) for a in aug], np.float32).view(np.int32)
        # the blur taps stay one torchvision-identical float32 op sequence per image (a batched exp / sum could round differently)
        taps[:, :self.kernel_size] = torch.stack([_gaussian_kernel1d(self.kernel_size, a["sigma"]) for a in aug])
        return torch.from_numpy(ints), taps

    def __call__(self, frames, augment=None, out=None, want_u8=False):
        frames = list(frames)
        n_frames = len(frames)

        def draw_and_pack():                             # Python-bound: runs while the pool threads copy the frames
            aug = augment if augment is not None else [self.draw_augment() for _ in range(n_frames)]
            return list(aug), self.pack_augment(aug)

        dev, desc_dev, max_crop, n = self._upload(frames, overlap=draw_and_pack)
        if n == 0:
            return torch.empty(0, 3, self.size, self.size, dtype=torch.float32, device=self.device)
        self.last_augment, (ints, taps) = self._overlap_result
        ints_d = ints.pin_memory().to(self.device, non_blocking=True)
        taps_d = taps.pin_memory().to(self.device, non_blocking=True)
        self.h2d_bytes += ints.numel() * 4 + taps.numel() * 4
        img = ops.crop_resize_u8(dev, desc_dev, self.size, max_crop, filter=self.filter)
        ops.color_jitter_u8(img, ints_d, out=img)
        blurred = ops.gaussian_blur_u8(img, taps_d, self.kernel_size)
        return ops.flip_rotate_normalize_u8(blurred, ints_d, self.mean, self.std, out=out, want_u8=want_u8)
