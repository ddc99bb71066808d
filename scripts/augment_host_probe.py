"""Where the end-to-end time of TrainAugment goes at 128 frames of 576x720: host pack into pinned staging (torch / numpy,
1-8 threads, per-frame or chunked tasks), the H2D copy, the resample kernel.  One JSON line."""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from ssl4polyp_b200 import ops

H, W, N = 576, 720, 128
rng = np.random.default_rng(0)
frames = [rng.integers(0, 256, (H, W, 3), dtype=np.uint8) for _ in range(N)]
sz = H * W * 3
stage = torch.empty(N * sz, dtype=torch.uint8).pin_memory()
st_np = stage.numpy()


def pack_t(i):
    stage[i * sz:(i + 1) * sz].view(H, W, 3).copy_(torch.from_numpy(frames[i])[0:H, 0:W])


def pack_n(i):
    st_np[i * sz:(i + 1) * sz].reshape(H, W, 3)[...] = frames[i][0:H, 0:W]


def rng_of(fn):
    def g(r):
        for i in r:
            fn(i)
    return g


def t(fn, reps=6):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return round((time.perf_counter() - t0) / reps * 1e3, 3)


res = {"cores": os.cpu_count()}
res["torch_1"] = t(lambda: rng_of(pack_t)(range(N)))
res["numpy_1"] = t(lambda: rng_of(pack_n)(range(N)))
for nt in (2, 4, 8, 16):
    pool = ThreadPoolExecutor(nt)
    ch = [range(k * N // nt, (k + 1) * N // nt) for k in range(nt)]
    res[f"torch_{nt}_tasks"] = t(lambda: list(pool.map(pack_t, range(N))))
    res[f"torch_{nt}_chunked"] = t(lambda: list(pool.map(rng_of(pack_t), ch)))
    res[f"numpy_{nt}_chunked"] = t(lambda: list(pool.map(rng_of(pack_n), ch)))
    pool.shutdown()
dev = torch.empty(N * sz, dtype=torch.uint8, device="cuda")


def ev(fn, reps=6):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / reps, 3)


res["h2d_ms"] = ev(lambda: dev.copy_(stage, non_blocking=True))
rows = np.zeros((N, 8), np.int64)
for i in range(N):
    rows[i] = (i * sz, W * 3, 0, 0, H, W, 0, 0)
desc = torch.from_numpy(rows).cuda()
res["resize_ms"] = ev(lambda: ops.crop_resize_u8(dev, desc, 224, max(H, W), filter="bilinear"))
print(json.dumps(res))

# ---- second part: the Python-bound parameter work, alone and under the copy threads; the device side alone
from ssl4polyp_b200.input_pipeline import TrainAugment, IMAGENET_MEAN, IMAGENET_STD
tf = TrainAugment(224)
torch.manual_seed(0)
aug = [tf.draw_augment() for _ in range(N)]
res2 = {"draw_ms": t(lambda: [tf.draw_augment() for _ in range(N)]), "pack_augment_ms": t(lambda: tf.pack_augment(aug))}
pool = ThreadPoolExecutor(8)
ch = [range(k * N // 8, (k + 1) * N // 8) for k in range(8)]


def both():
    fs = [pool.submit(rng_of(pack_n), r) for r in ch]
    tf.pack_augment(aug)
    for f in fs:
        f.result()


res2["numpy_8_chunked_with_pack_augment_ms"] = t(both)
ints, taps = tf.pack_augment(aug)
ints_d, taps_d = ints.cuda(), taps.cuda()
out = torch.empty(N, 3, 224, 224, device="cuda")


def device_side():
    d = torch.empty(N * sz, dtype=torch.uint8, device="cuda")
    d.copy_(stage, non_blocking=True)
    img = ops.crop_resize_u8(d, desc, 224, max(H, W), filter="bilinear")
    ops.color_jitter_u8(img, ints_d, out=img)
    b = ops.gaussian_blur_u8(img, taps_d, 25)
    ops.flip_rotate_normalize_u8(b, ints_d, IMAGENET_MEAN, IMAGENET_STD, out=out)


res2["device_side_ms"] = ev(device_side)


def device_and_host():
    device_side()
    both()


res2["device_side_while_host_packs_ms"] = ev(device_and_host)
frames_l = frames
res2["call_predrawn_ms"] = ev(lambda: tf(frames_l, augment=aug, out=out))
res2["call_with_draws_ms"] = ev(lambda: tf(frames_l, out=out))
print(json.dumps(res2))
