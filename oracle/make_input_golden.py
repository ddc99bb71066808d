"""Generates tests/golden/input_golden.npz from the INSTALLED Pillow + torchvision (run in the build container; the GPU
box only reads the committed file):  python oracle/make_input_golden.py
Each case: a seeded uint8 frame, the crop box / flip drawn by torchvision's own RandomResizedCrop + RandomHorizontalFlip
under torch.manual_seed(seed), the uint8 image after PIL's bicubic resize and the sha256 of the float32 tensor after ToTensor+Normalize (the frame itself is
regenerated from its seed by input_oracle.synthetic_frame).
Also asserts the oracle restatement (oracle/input_oracle.py) reproduces all of it bit for bit."""
import hashlib
import os
import sys

import numpy as np
import torch
from PIL import Image
import torchvision.transforms as T
import torchvision.transforms.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import input_oracle as IO  # noqa: E402

CASES = [(0, 288, 352), (1, 576, 720), (2, 1072, 1920), (3, 224, 224), (4, 100, 130), (5, 224, 500), (6, 37, 41), (7, 1350, 1080)]
out, frames = {}, {}
for seed, H, W in CASES:
    frame = IO.synthetic_frame(seed, H, W)
    pil = Image.fromarray(frame)
    torch.manual_seed(seed)
    i, j, h, w = T.RandomResizedCrop.get_params(pil, scale=(0.2, 1.0), ratio=(3. / 4., 4. / 3.))
    flip = bool(torch.rand(1) < 0.5)
    if seed == 3:
        i, j, h, w = 0, 0, 224, 224                  # both passes skipped
    if seed == 5:
        i, j, h, w = 0, 100, 224, 300                # vertical pass skipped
    resized = F.resized_crop(pil, i, j, h, w, (224, 224), interpolation=T.InterpolationMode.BICUBIC)
    u8 = np.asarray(resized)
    t = F.normalize(F.to_tensor(F.hflip(resized) if flip else resized), IO.IMAGENET_MEAN, IO.IMAGENET_STD).numpy()
    # the oracle restatement must agree bit for bit, including the RNG-driven parameters
    torch.manual_seed(seed)
    pi, pj, ph, pw = IO.random_resized_crop_params(H, W)
    pflip = bool(torch.rand(1) < 0.5)
    if seed not in (3, 5):
        assert (pi, pj, ph, pw, pflip) == (i, j, h, w, flip), (seed, (pi, pj, ph, pw, pflip), (i, j, h, w, flip))
    o8 = IO.crop_resize_u8(frame, i, j, h, w)
    assert np.array_equal(o8, u8), (seed, np.abs(o8.astype(int) - u8.astype(int)).max())
    ot = IO.to_tensor_normalize(o8, flip)
    assert ot.dtype == np.float32 and np.array_equal(ot.view(np.uint32), t.view(np.uint32)), seed
    frames[seed] = frame
    out[f"params{seed}"] = np.array([H, W, i, j, h, w, int(flip)], np.int32)
    out[f"u8_{seed}"] = u8
    out[f"tensor_sha256_{seed}"] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(t).tobytes()).digest(), np.uint8)
    print(f"case {seed}: {H}x{W} crop {(i, j, h, w)} flip {flip}: oracle == PIL/torchvision bit for bit")
# the full Compose under one seed
torch.manual_seed(123)
tf = T.Compose([T.RandomResizedCrop(224, scale=(0.2, 1.0), interpolation=T.InterpolationMode.BICUBIC), T.RandomHorizontalFlip(), T.ToTensor(),
                T.Normalize(mean=IO.IMAGENET_MEAN, std=IO.IMAGENET_STD)])
ref = [tf(Image.fromarray(frames[s])).numpy() for s in (0, 1, 4)]
torch.manual_seed(123)
mine = [IO.transform_train(frames[s])[0] for s in (0, 1, 4)]
for a, b in zip(ref, mine):
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
print("Compose under one seed: bit-identical")
# the fine-tune eval transform: T.Resize((224, 224)) (PIL bilinear) -> ToTensor -> Normalize
ev = T.Compose([T.Resize((224, 224)), T.ToTensor(), T.Normalize(mean=IO.IMAGENET_MEAN, std=IO.IMAGENET_STD)])
for seed in (0, 1, 4, 6, 7):
    t = ev(Image.fromarray(frames[seed])).numpy()
    o = IO.transform_eval(frames[seed])
    assert np.array_equal(o.view(np.uint32), t.view(np.uint32)), seed
    out[f"eval_sha256_{seed}"] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(t).tobytes()).digest(), np.uint8)
print("eval transform (bilinear Resize): bit-identical")
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "input_golden.npz")
np.savez_compressed(dst, **out)
print("wrote", dst, os.path.getsize(dst) // 1024, "KiB")
