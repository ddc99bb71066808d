"""CPU-only tests (`-m "not gpu"`): the oracle against the reference-generated golden vectors, the C-ABI surface,
the host-side mirror of the reference interface, and the data-parallel bucket logic over gloo (world_size 2).
No CUDA compute is called here."""
import ctypes
import math
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import mae_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "mae_golden.pt")


@pytest.fixture(scope="module")
def golden():
    return torch.load(GOLDEN)


def subsample(t, n=64):
    f = t.detach().reshape(-1)
    step = max(1, f.numel() // n)
    return f[::step][:n].clone()


def synth_imgs(B, seed, size=224):
    return torch.randn(B, 3, size, size, generator=torch.Generator().manual_seed(1000 + seed))


# ------------------------------------------------------------------------------------------------ oracle vs golden
def test_oracle_masking_matches_reference_golden(golden):
    for c in golden["cases"]:
        ids_keep, mask, ids_restore = O.masking_from_noise(c["noise"], c["mask_ratio"])
        assert torch.equal(ids_restore, c["ids_restore"])
        assert torch.equal(mask, c["mask"])
        assert ids_restore.dtype == torch.int64 and mask.dtype == torch.float32
        L = c["noise"].shape[1]
        assert ids_keep.shape[1] == int(L * (1 - c["mask_ratio"]))
    t = golden["ties"]
    ids_keep, mask, ids_restore = O.masking_from_noise(t["noise"], t["mask_ratio"])
    assert torch.equal(ids_keep, t["ids_keep"]) and torch.equal(mask, t["mask"]) and torch.equal(ids_restore, t["ids_restore"])


def test_oracle_tiny_case_full_forward_backward_matches_reference_golden(golden):
    c = golden["cases"][4]
    cfg = O.MAEConfig(**c["cfg"])
    sd = O.init_state_dict(cfg, seed=c["seed"])
    assert torch.equal(subsample(sd["blocks.0.attn.qkv.weight"], 16), c["w_sample"])
    imgs = synth_imgs(c["B"], c["seed"], cfg.img_size)
    assert torch.equal(subsample(imgs, 16), c["imgs_sample"])
    sd_g = {k: v.clone().requires_grad_(k not in ("pos_embed", "decoder_pos_embed")) for k, v in sd.items()}
    loss, pred, mask = O.mae_forward(sd_g, imgs, c["noise"], c["mask_ratio"], cfg)
    loss.backward()
    assert abs(loss.item() - c["loss"].item()) < 1e-5 * abs(c["loss"].item())
    assert torch.allclose(subsample(pred, 256), c["pred_sample"], rtol=1e-4, atol=1e-5)
    for k, gn in c["grad_norm"].items():
        assert abs(sd_g[k].grad.norm().item() - gn.item()) <= 1e-3 * gn.item() + 1e-8, k
    for k, gs in c["grad_sample"].items():
        assert torch.allclose(subsample(sd_g[k].grad, 64), gs, rtol=1e-3, atol=1e-6), k


def test_oracle_base_forward_matches_reference_golden(golden):
    c = golden["cases"][0]
    cfg = O.MAEConfig(**c["cfg"])
    sd = O.init_state_dict(cfg, seed=c["seed"])
    imgs = synth_imgs(c["B"], c["seed"])
    with torch.no_grad():
        loss, pred, mask = O.mae_forward(sd, imgs, c["noise"], c["mask_ratio"], cfg)
    assert abs(loss.item() - c["loss"].item()) < 1e-5 * abs(c["loss"].item())
    assert torch.allclose(subsample(pred, 256), c["pred_sample"], rtol=1e-4, atol=1e-5)
    assert torch.equal(mask, c["mask"])


CLS_GOLDEN = os.path.join(ROOT, "tests", "golden", "cls_golden.pt")


@pytest.fixture(scope="module")
def cls_golden():
    return torch.load(CLS_GOLDEN)


def cls_case_state_dict(c):
    """Weights of a cls_golden case, rebuilt from its seed (pinned bit-for-bit to the reference constructors by
    oracle/make_cls_golden.py)."""
    if c["kind"] == "mae":
        return O.init_state_dict(O.MAE_BASE, seed=c["seed"], head_classes=c["num_classes"])
    return O.vit_any_state_dict(c["seed"], c["num_classes"])


@pytest.mark.parametrize("idx", [0, 3, 4])
def test_oracle_classifier_matches_reference_golden(cls_golden, idx):
    """a13/a14/a17: `O.classifier_forward` + the loss glue against vectors produced by the UNMODIFIED reference
    `models/models.py` (ViT_from_MAE :143-222, VisionTransformer_from_Any :26-140) and
    `train_classification._compute_supervised_loss` (:3361-3374).  Cases 1 and 2 are covered by the GPU tests."""
    c = cls_golden["cases"][idx]
    sd = cls_case_state_dict(c)
    assert torch.equal(subsample(sd["blocks.0.attn.qkv.weight"], 16), c["w_sample"])
    assert torch.equal(sd["lin_head.weight"], c["head_w"])
    frozen = ("pos_embed", "decoder_pos_embed") if c["kind"] == "mae" else ()
    sd_g = {k: v.clone().requires_grad_(k not in frozen) for k, v in sd.items()}
    logits = O.classifier_forward(sd_g, synth_imgs(c["B"], c["seed"]), O.MAE_BASE, c["out_token"], True)
    if c["num_classes"] == 2:
        loss = O.binary_polyp_loss(logits, c["target"], pos_weight=torch.tensor(c["pos_weight"]))
    else:
        loss = O.multiclass_loss(logits, c["target"], c["class_weights"])
    loss.backward()
    assert torch.allclose(logits.detach(), c["logits"], rtol=1e-5, atol=1e-6)
    assert abs(loss.item() - c["loss"].item()) < 1e-6 * abs(c["loss"].item())
    for k, gn in c["grad_norm"].items():
        assert abs(sd_g[k].grad.norm().item() - gn.item()) <= 1e-3 * gn.item() + 1e-10, k
    for k, gs in c["grad_sample"].items():
        assert torch.allclose(subsample(sd_g[k].grad, 256), gs, rtol=1e-3, atol=1e-7), k


def test_pretrain_lr_schedule_matches_reference_golden(cls_golden):
    """a19: `engine.cosine_lr` (and the oracle's copy) against values computed by the reference's own
    `util/lr_sched.py:9-21` over warm-up + decay (fixture written by oracle/make_cls_golden.py)."""
    from ssl4polyp_b200.engine import cosine_lr
    g = cls_golden["lr"]
    for args, pts in ((g["args"], g["points"]), (g["short_args"], g["short_points"])):
        for e, want in pts:
            assert cosine_lr(e, args["lr"], args["min_lr"], args["warmup_epochs"], args["epochs"]) == want, e
            assert O.cosine_lr(e, args["lr"], args["min_lr"], args["warmup_epochs"], args["epochs"]) == want, e
    a = g["args"]
    assert cosine_lr(0.0, **a) == 0.0 and cosine_lr(a["warmup_epochs"], **a) == a["lr"]
    assert abs(cosine_lr(a["epochs"], **a) - a["min_lr"]) < 1e-12


def test_oracle_patchify_roundtrip_and_len_keep():
    x = torch.randn(2, 3, 64, 64)
    assert torch.equal(O.unpatchify(O.patchify(x)), x)
    assert O.len_keep_of(196, 0.75) == 49 and O.len_keep_of(196, 0.6) == 78      # int(196*0.4) = int(78.4)


# ------------------------------------------------------------------------------------------------ C-ABI surface
def test_abi_header_symbols_exported_and_bound():
    from ssl4polyp_b200 import _lib
    with open(os.path.join(ROOT, "include", "ssl4p.h")) as f:
        header = f.read()
    declared = set(re.findall(r"\b(ssl4p_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 30
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} not exported by libssl4p.so"
    loaded = _lib.load()
    assert loaded.ssl4p_abi_version() == 1
    assert loaded.ssl4p_last_error() is not None


def test_library_is_sm100a_tcgen05_and_tma():
    """SASS evidence (B200_PROFILING.md): UTC*MMA = tcgen05.mma, UTMALDG = TMA loads, LDTM = tcgen05.ld."""
    from ssl4polyp_b200 import _lib
    out = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    sass = out.stdout
    assert "sm_100a" in sass
    assert re.search(r"UTC\w*MMA", sass), "no tcgen05.mma in the library"
    assert "UTMALDG" in sass and "LDTM" in sass


def test_product_never_imports_the_oracle_and_has_no_cpu_path():
    pkg = os.path.join(ROOT, "ssl4polyp_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith(".py"):
                src = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f"{fn} imports the oracle"
    from ssl4polyp_b200 import models_mae, ops
    m = models_mae.MaskedAutoencoderViT(img_size=32, embed_dim=64, depth=1, num_heads=2, decoder_embed_dim=64,
                                        decoder_depth=1, decoder_num_heads=2)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 3, 32, 32))
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.mask_from_noise(torch.rand(2, 16), 4)


# ------------------------------------------------------------------------------------------------ host-side mirror
def test_constructor_init_is_bit_identical_to_reference_init():
    from ssl4polyp_b200 import models_mae
    torch.manual_seed(0)
    m = models_mae.mae_vit_base_patch16(norm_pix_loss=True)
    sd = O.init_state_dict(O.MAE_BASE, seed=0)      # pinned bit-for-bit to the reference constructor by make_golden.py
    msd = m.state_dict()
    assert set(sd) == set(msd)
    for k in sd:
        assert msd[k].shape == sd[k].shape and torch.equal(msd[k], sd[k]), k
    assert not m.pos_embed.requires_grad and not m.decoder_pos_embed.requires_grad
    assert m.norm.eps == 1e-6 and m.norm_pix_loss is True
    assert m.patch_embed.patch_size == (16, 16) and m.patch_embed.num_patches == 196
    assert isinstance(m.blocks, torch.nn.ModuleList) and len(m.blocks) == 12 and len(m.decoder_blocks) == 8


def test_reference_factories_and_classifier_surface():
    from ssl4polyp_b200 import models, models_mae, utils
    assert models_mae.mae_vit_large_patch16 is models_mae.mae_vit_large_patch16_dec512d8b
    assert callable(models_mae.mae_vit_huge_patch14)
    m = utils.get_MAE_backbone(None, True, 2, False, None, out_token="cls")
    assert isinstance(m, models.ViT_from_MAE)
    keys = set(m.state_dict())
    assert {"lin_head.weight", "lin_head.bias", "cls_token", "pos_embed", "decoder_pos_embed", "norm.weight"} <= keys
    assert not any(k.startswith(("decoder_blocks", "decoder_embed", "decoder_pred", "decoder_norm", "mask_token")) for k in keys)
    assert m.lin_head.weight.shape == (2, 768) and m.head is True and m.frozen is False and m.out_token == "cls"
    # finetune.py:49-91 contract: blocks is a ModuleList, head module is `lin_head`
    for p in m.parameters():
        p.requires_grad_(False)
    for p in list(m.blocks)[-1].parameters():
        p.requires_grad_(True)
    assert sum(p.requires_grad for p in m.parameters()) == 12
    # --pretraining random / ImageNet_class backbone (models.py:26-140): timm 0.4.12 key set, learned pos_embed,
    # `blocks` an nn.Sequential (so finetune.py's head+k modes leave it frozen, as in the reference)
    v = utils.get_ImageNet_or_random_ViT(True, 2, False, None, False, out_token="spatial")
    assert isinstance(v, models.VisionTransformer_from_Any) and isinstance(v.blocks, torch.nn.Sequential)
    vk = list(v.state_dict())
    assert vk[:4] == ["cls_token", "pos_embed", "patch_embed.proj.weight", "patch_embed.proj.bias"] and len(vk) == 152
    assert v.pos_embed.requires_grad and v.pos_embed.shape == (1, 197, 768) and 0 < float(v.pos_embed.std()) < 0.03
    assert v.blocks[0].attn.qkv.bias is not None and v.head_bool is True and isinstance(v.head, torch.nn.Identity)
    with pytest.raises(RuntimeError):
        utils.get_ImageNet_or_random_ViT(True, 2, False, None, True)


def test_mae_to_classifier_checkpoint_handoff(tmp_path):
    from ssl4polyp_b200 import models, models_mae
    torch.manual_seed(1)
    mae = models_mae.mae_vit_base_patch16()
    path = tmp_path / "checkpoint-0.pth"
    torch.save({"model": mae.state_dict()}, path)                      # util/misc.py:312-318 layout
    clf = models.ViT_from_MAE(str(path), True, 2, False, None, 768, 12, 12, "cls")
    for k, v in clf.state_dict().items():
        if k in mae.state_dict():
            assert torch.equal(v, mae.state_dict()[k]), k


def test_sincos_table_matches_oracle():
    from ssl4polyp_b200.util.pos_embed import get_2d_sincos_pos_embed
    for dim in (768, 512, 1024):
        t = torch.from_numpy(get_2d_sincos_pos_embed(dim, 14, cls_token=True)).float()
        assert torch.equal(t, O.sincos_pos_embed(dim, 14))


# ------------------------------------------------------------------------------------------------ DP bucket logic
def test_bucket_plan_covers_trainable_range_in_backward_order():
    from ssl4polyp_b200.dp import plan_buckets
    offsets, off = {}, 0
    sizes = {"cls": 768, "pos": 197 * 768, "pe.w": 768 * 768, "b0.w": 3 * 768 * 768, "b0.b": 2304, "b1.w": 4 * 768 * 768, "head": 512 * 768}
    for k, n in sizes.items():
        offsets[k] = (off, n)
        off += (n + 255) // 256 * 256
    trainable = [k for k in sizes if k != "pos"]
    buckets = plan_buckets(offsets, trainable, limit_elems=2_000_000)
    assert buckets[0].end == off and "head" in buckets[0].names
    assert buckets[-1].start == offsets["cls"][0]
    for a, b in zip(buckets, buckets[1:]):
        assert b.end == a.start                       # contiguous, no gaps
    assert set().union(*[b.names for b in buckets]) == set(trainable)


_DP_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[3])
from ssl4polyp_b200.dp import plan_buckets, BucketReducer, broadcast_parameters
rank, world = int(sys.argv[1]), 2
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{sys.argv[2]}", rank=rank, world_size=world)
# ranks are seeded seed+rank before the model is built (main_pretrain.py:146-148): the arenas differ until rank 0's is broadcast
torch.manual_seed(7 + rank)
arena = torch.randn(5000)
want0 = torch.randn(5000, generator=torch.Generator().manual_seed(7))
assert torch.equal(arena, want0) == (rank == 0)
broadcast_parameters(arena)
assert torch.equal(arena, want0), "arena differs from rank 0's after the constructor broadcast"
offsets, off = {}, 0
for i in range(12):
    n = 1000 + 37 * i
    offsets[f"p{i}"] = (off, n); off += (n + 255) // 256 * 256
trainable = [f"p{i}" for i in range(12) if i != 3]           # p3 frozen
flat = torch.zeros(off)
for i in range(12):
    a, n = offsets[f"p{i}"]
    if i != 3:
        flat[a:a + n] = (rank + 1) * (i + 1)
buckets = plan_buckets(offsets, trainable, limit_elems=3000)
red = BucketReducer(flat, buckets)
assert red.world == 2
for i in reversed(range(12)):                                 # backward order; p5 never reports (unused parameter)
    if i not in (3, 5):
        red.ready(f"p{i}")
red.finish()
for i in range(12):
    a, n = offsets[f"p{i}"]
    want = 0.0 if i == 3 else 3.0 * (i + 1)
    assert torch.all(flat[a:a + n] == want), (i, flat[a].item(), want)
assert red.launch_order[0] == 0 and sorted(red.launch_order) == list(range(len(buckets)))
dist.destroy_process_group()
print("ok", rank)
"""


def test_bucket_reducer_world_size_2_gloo(tmp_path):
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    script = tmp_path / "dp_worker.py"
    script.write_text(_DP_WORKER)
    procs = [subprocess.Popen([sys.executable, str(script), str(r), str(port), ROOT], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
             for r in range(2)]
    for p in procs:
        out, err = p.communicate(timeout=180)
        assert p.returncode == 0, err[-2000:]
        assert "ok" in out


def test_synthetic_frames_shards_like_distributed_sampler():
    """Every rank must see the same number of equally-sized batches (DistributedSampler pads by wrapping around)."""
    from ssl4polyp_b200.train_classification import SyntheticFrames
    for n, b, w in [(40, 8, 2), (41, 8, 2), (7, 4, 3), (40, 8, 1), (9, 2, 4)]:
        loaders = [SyntheticFrames(n, b, 2, 8, 0, (r, w), shuffle=True) for r in range(w)]
        shapes = [[x[0].shape[0] for x in l] for l in loaders]
        assert all(s == shapes[0] for s in shapes), (n, b, w, shapes)
        assert all(len(l) == len(shapes[0]) for l in loaders)
        seen = torch.cat([l._rank_order() for l in loaders]).tolist()
        assert set(seen) == set(range(n)) and len(seen) == (n + w - 1) // w * w


def test_optimizer_state_layout_is_the_reference_adamw_state_dict():
    """engine.optimizer_state_dict() must be what `torch.optim.AdamW(add_weight_decay(model, wd)).state_dict()` looks like
    (timm 0.4.12 group order: no-decay first; main_pretrain.py:216-218), and for the fine-tune script the head / backbone
    groups of train_classification.py:5751-5768 - checked against a real torch optimizer built the reference's way."""
    from ssl4polyp_b200.engine import finetune_group_plan, timm_weight_decay_plan
    net = torch.nn.ModuleDict({"patch": torch.nn.Conv2d(3, 8, 4, 4), "norm": torch.nn.LayerNorm(8), "fc": torch.nn.Linear(8, 8),
                               "lin_head": torch.nn.Linear(8, 2)})
    net.pos = torch.nn.Parameter(torch.zeros(1, 5, 8), requires_grad=False)
    named = list(net.named_parameters())
    plan = timm_weight_decay_plan(named, 0.05)
    no_decay = [p for n, p in named if p.requires_grad and (p.ndim == 1 or n.endswith(".bias"))]
    decay = [p for n, p in named if p.requires_grad and not (p.ndim == 1 or n.endswith(".bias"))]
    opt = torch.optim.AdamW([{"params": no_decay, "weight_decay": 0.}, {"params": decay, "weight_decay": 0.05}], lr=1e-3, betas=(0.9, 0.95))
    by_id = {id(p): n for n, p in named}
    want = [[by_id[id(p)] for p in g["params"]] for g in opt.param_groups]
    assert [g for _, g in plan] == want and "pos" not in sum(want, [])
    assert [e["weight_decay"] for e, _ in plan] == [0.0, 0.05]
    fplan = finetune_group_plan(named, 1e-4)
    assert [e["name"] for e, _ in fplan] == ["head", "backbone"]
    assert fplan[0][1] == ["lin_head.weight", "lin_head.bias"] and "pos" in fplan[1][1] and len(fplan[1][1]) == len(named) - 2


def test_main_pretrain_parser_matches_the_reference_flag_for_flag():
    """Same flags and defaults as /root/reference/.../mae/main_pretrain.py:47-135 (parsed out of the reference source when
    the reference tree is present; the frozen expectation below otherwise)."""
    import argparse
    from pathlib import Path
    from ssl4polyp_b200 import main_pretrain as MP
    ours = vars(MP.get_args_parser().parse_args([]))
    ref_path = "/root/reference/src/ssl4polyp/models/mae/main_pretrain.py"
    if os.path.exists(ref_path):
        src = open(ref_path).read()
        ns = {"argparse": argparse, "Path": Path}
        exec(src[src.index("def get_args_parser"):src.index("def main(")], ns)
        extra = {"device_transform", "no_graph"}         # this implementation's own switches (input_pipeline.TrainTransform; eager launches)
        assert vars(ns["get_args_parser"]().parse_args([])) == {k: v for k, v in ours.items() if k not in extra}
        assert ours["device_transform"] is False
    assert ours["batch_size"] == 64 and ours["model"] == "mae_vit_large_patch16" and ours["mask_ratio"] == 0.75
    assert ours["blr"] == 1e-3 and ours["warmup_epochs"] == 40 and ours["weight_decay"] == 0.05 and ours["precision"] == "amp"
    a = MP.get_args_parser().parse_args(["--acum_iter", "4", "--no_train_dir", "--save-freq-mins", "30", "--keep-last", "2"])
    assert a.accum_iter == 4 and a.no_train_dir and a.save_freq_mins == 30 and a.keep_last == 2


def test_finetune_schedulers_match_torch():
    """train_classification.py:3943-3971 `create_scheduler`: cosine LambdaLR and ReduceLROnPlateau(mode="max") on floats."""
    from ssl4polyp_b200.engine import PlateauLR, finetune_cosine_factor
    w = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([w], lr=1e-3)
    sch = torch.optim.lr_scheduler.LambdaLR(opt, lambda e: finetune_cosine_factor(e, 20, 3))
    for e in range(20):
        assert abs(opt.param_groups[0]["lr"] - 1e-3 * finetune_cosine_factor(e, 20, 3)) < 1e-12
        opt.step(); sch.step()
    assert finetune_cosine_factor(0, 20, 3) == 1 / 3 and finetune_cosine_factor(3, 20, 3) == 1.0 and finetune_cosine_factor(25, 20, 3) == 0.0
    opt = torch.optim.AdamW([w], lr=1e-3)
    ref = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="max", factor=0.5, patience=2, min_lr=1e-6)
    ours = PlateauLR(1e-3, factor=0.5, patience=2, min_lr=1e-6)
    metrics = [0.5, 0.6, 0.6, 0.59, 0.6, 0.61, 0.6, 0.6, 0.6, 0.6, 0.6, 0.6, 0.62, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1]
    for m in metrics:
        ref.step(m)
        assert abs(ours.step(m) - opt.param_groups[0]["lr"]) < 1e-15, m


def test_classifier_engine_decays_every_parameter_like_the_reference():
    """train_classification.py:5751-5768 builds torch.optim.AdamW with one weight decay for all parameters; the pretraining
    script uses timm's rule (no decay for 1-D tensors and biases, main_pretrain.py:216-218)."""
    import inspect
    from ssl4polyp_b200 import engine
    src = inspect.getsource(engine.ClassifierEngine.__init__)
    assert "no_decay_rule" in src and "lambda name, p: False" in src
    assert inspect.signature(engine.ClassifierEngine.__init__).parameters["weight_decay"].default == 0.0


def test_train_classification_parser_matches_the_reference_flag_for_flag():
    """Every flag, dest and default of reference get_args (train_classification.py:7706-7982), checked by executing the
    reference's own parser-building lines when the tree is present; the frozen expectation below otherwise."""
    import argparse
    from pathlib import Path
    from ssl4polyp_b200 import train_classification as TC
    ours = vars(TC.get_args(["--pretraining", "random"]))
    ref_path = "/root/reference/src/ssl4polyp/classification/train_classification.py"
    if os.path.exists(ref_path):
        src = open(ref_path).read()
        body = src[src.index("def get_args():"):src.index("    seed_flag_present = any(")]
        ns = {"argparse": argparse, "Path": Path, "data_packs_root": lambda: Path("data_packs")}
        exec(body + "    return parser\n", ns)
        ref = vars(ns["get_args"]().parse_args(["--pretraining", "random"]))
        extra = {"synthetic_samples", "num_classes", "_seed_explicit"}
        assert set(ours) - extra == set(ref)
        assert {k: v for k, v in ours.items() if k not in extra} == ref
    assert ours["lr"] == 1e-4 and ours["batch_size"] == 16 and ours["epochs"] == 50 and ours["weight_decay"] == 0.05
    assert ours["finetune_mode"] == "full" and ours["scheduler"] == "none" and ours["early_stop_monitor"] == "val_loss"
    assert ours["seed"] == 42 and ours["pin_memory"] is True and ours["perturbation_splits"] == ["test"]
    a = TC.get_args(["--seeds", "7", "9"]);            assert TC._active_seed(a) == 7
    a = TC.get_args(["--seeds", "7", "--seed", "3"]);  assert TC._active_seed(a) == 3


def test_finetune_modes_select_the_reference_parameter_sets():
    """finetune.py:49-92: full / none / head+1 / head+2; nn.Sequential blocks are skipped by the isinstance test."""
    from ssl4polyp_b200 import train_classification as TC

    class Toy(torch.nn.Module):
        def __init__(self, seq):
            super().__init__()
            blocks = [torch.nn.Linear(2, 2) for _ in range(3)]
            self.blocks = torch.nn.Sequential(*blocks) if seq else torch.nn.ModuleList(blocks)
            self.norm = torch.nn.LayerNorm(2)
            self.lin_head = torch.nn.Linear(2, 2)
            self.head = True          # ViT_from_MAE.head is a bool, not a module (models.py:163)
            self.frozen = False

    def trainable(m):
        return sorted(n for n, p in m.named_parameters() if p.requires_grad)

    m = Toy(False)
    TC.configure_finetune_parameters(m, "none");   assert trainable(m) == ["lin_head.bias", "lin_head.weight"] and m.frozen
    TC.configure_finetune_parameters(m, "head+1"); assert trainable(m) == ["blocks.2.bias", "blocks.2.weight", "lin_head.bias", "lin_head.weight"]
    TC.configure_finetune_parameters(m, "HEAD+2"); assert len(trainable(m)) == 6 and "blocks.1.weight" in trainable(m) and not m.frozen
    TC.configure_finetune_parameters(m, "full");   assert len(trainable(m)) == len(list(m.parameters()))
    s = Toy(True)
    TC.configure_finetune_parameters(s, "head+2"); assert trainable(s) == ["lin_head.bias", "lin_head.weight"]
    with pytest.raises(ValueError):
        TC.normalise_finetune_mode("head+3")
    ref_path = "/root/reference/src/ssl4polyp/classification/finetune.py"
    if os.path.exists(ref_path):
        import importlib.util
        spec = importlib.util.spec_from_file_location("_ref_finetune", ref_path)
        ref = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref)
        for seq in (False, True):
            for mode in ("none", "full", "head+1", "head+2"):
                a, b = Toy(seq), Toy(seq)
                TC.configure_finetune_parameters(a, mode); ref.configure_finetune_parameters(b, mode)
                assert trainable(a) == trainable(b) and a.frozen == b.frozen, (seq, mode)


def test_monitor_early_stop_and_selection_tag_helpers():
    """train_classification.py:3297-3345, :3878-3915."""
    from ssl4polyp_b200 import train_classification as TC
    assert TC.resolve_monitor_mode("val_loss", "auto") == "min" and TC.resolve_monitor_mode("val_auroc", None) == "max"
    assert TC.resolve_monitor_mode("val_auroc", "MIN") == "min"
    assert TC.resolve_monitor_key("val_auroc") == "auroc" and TC.resolve_monitor_key(None) == "loss"
    assert TC.improved(0.4, None, mode="min", min_delta=0) and TC.improved(0.4, 0.5, mode="min", min_delta=0.05)
    assert not TC.improved(0.46, 0.5, mode="min", min_delta=0.05) and not TC.improved(float("nan"), 0.5, mode="max", min_delta=0)
    assert TC.should_stop(3, 3, 5, 0) and not TC.should_stop(3, 0, 5, 0) and not TC.should_stop(3, 3, 5, 6)
    assert [TC.selection_tag(m) for m in ("val_loss", "val_auroc", "loss", "", "val_f1_score")] == ["valLoss", "valAUROC", "loss", "best", "valF1Score"]
    ref_path = "/root/reference/src/ssl4polyp/classification/train_classification.py"
    if os.path.exists(ref_path):
        import re
        src = open(ref_path).read()
        ns = {"re": re, "Optional": __import__("typing").Optional}
        exec(src[src.index("def _format_selection_tag"):src.index("def _should_trigger_early_stop")], ns)
        for m in ("val_loss", "val_auroc", "loss", "", None, "val_f1_score", "AUC-test", "val__auprc"):
            assert TC.selection_tag(m) == ns["_format_selection_tag"](m), m


def test_metrics_match_sklearn_on_seeded_inputs_with_ties():
    """classification/metrics/performance.py + thresholds.py call scikit-learn; ssl4polyp_b200.metrics restates them."""
    sk = pytest.importorskip("sklearn.metrics")
    from ssl4polyp_b200 import metrics as M
    rng = np.random.default_rng(0)
    for n in (5, 64, 501):
        for ties in (False, True):
            y = rng.integers(0, 2, n); y[0], y[1] = 0, 1
            s = rng.random(n)
            s = np.round(s, 1) if ties else s
            p = np.stack([1 - s, s], 1)
            assert abs(M.auroc(p, y, 2) - sk.roc_auc_score(y, s)) < 1e-12
            assert abs(M.auprc(p, y, 2) - sk.average_precision_score(y, s)) < 1e-12
            assert abs(M.balanced_accuracy(p, y, 2) - sk.balanced_accuracy_score(y, p.argmax(1))) < 1e-12
            assert abs(M.mcc(p, y) - sk.matthews_corrcoef(y, p.argmax(1))) < 1e-12
            fpr, tpr, th = sk.roc_curve(y, s)
            tau = th[np.argmax(tpr - fpr)]
            tau = np.nextafter(s.max(), 1.0) if np.isinf(tau) else tau
            assert M.youden_j_threshold(p, y) == tau
            lab = (s >= tau).astype(int)
            assert abs(M.mean_f1(p, y, 2, tau) - sk.f1_score(y, lab, average="macro")) < 1e-6
            assert abs(M.mean_recall(p, y, 2, tau) - sk.recall_score(y, lab, average="macro")) < 1e-6
    y = rng.integers(0, 3, 300); p = rng.random((300, 3)); p /= p.sum(1, keepdims=True)
    assert abs(M.auroc(p, y, 3) - sk.roc_auc_score(y, p, multi_class="ovr", average="macro")) < 1e-12
    assert abs(M.auprc(p, y, 3) - sk.average_precision_score(np.eye(3)[y], p, average="macro")) < 1e-12
    assert math.isnan(M.auroc(np.array([[.5, .5]] * 3), np.zeros(3, int), 2))
    sep = np.array([0.1, 0.2, 0.8, 0.9]); yy = np.array([0, 0, 1, 1])           # perfectly separable: argmax lands on a real score
    assert M.youden_j_threshold(np.stack([1 - sep, sep], 1), yy) == 0.8


def test_synthetic_frames_shard_and_class_weights():
    from ssl4polyp_b200 import train_classification as TC
    full = TC.SyntheticFrames(37, 8, 2, image_size=16, seed=3)
    assert len(full) == 5 and [b[0].shape[0] for b in full] == [8, 8, 8, 8, 5]
    r0 = TC.SyntheticFrames(37, 8, 2, image_size=16, seed=3, shard=(0, 2)); r1 = TC.SyntheticFrames(37, 8, 2, image_size=16, seed=3, shard=(1, 2))
    assert len(r0) == 3 and len(r1) == 3             # DistributedSampler semantics: 19 samples per rank (one wrapped around)
    assert [b[0].shape[0] for b in r0] == [b[0].shape[0] for b in r1] == [8, 8, 3]
    seen = torch.cat([r0._rank_order(), r1._rank_order()])
    assert len(seen) == 38 and set(seen.tolist()) == set(range(37))
    a, b = next(iter(full)), next(iter(TC.SyntheticFrames(37, 8, 2, image_size=16, seed=3)))
    assert torch.equal(a[0], b[0]) and a[0].dtype == torch.float32
    args = TC.get_args([])
    w, c = TC._class_weights(args, [0, 0, 0, 1], 2)
    assert c == [3, 1] and w == [4 / 6, 2.0] and TC._pos_weight(w, c) == 3.0          # :5617-5622, :6089-6094
    args.class_weights = "1,5"
    assert TC._class_weights(args, [0, 0, 0, 1], 2)[0] == [1.0, 5.0]
    args.class_weights = "1,2,3"
    with pytest.raises(ValueError):
        TC._class_weights(args, [0, 1], 2)


def test_input_oracle_matches_pillow_torchvision_golden():
    """oracle/input_oracle.py against tests/golden/input_golden.npz (made by oracle/make_input_golden.py from the
    installed Pillow + torchvision): uint8 after the bicubic resample and fp32 after ToTensor+Normalize, bit for bit;
    the crop boxes drawn by the product's host code follow torchvision's RNG sequence."""
    import hashlib
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import input_pipeline as P
    g = np.load(os.path.join(ROOT, "tests", "golden", "input_golden.npz"))
    for case in range(8):
        H, W, i, j, h, w, flip = (int(v) for v in g[f"params{case}"])
        if H * W > 800 * 800:
            continue                                   # the two large frames take the numpy oracle ~10 s: GPU test covers them
        frame = IO.synthetic_frame(case, H, W)
        u8 = IO.crop_resize_u8(frame, i, j, h, w)
        assert np.array_equal(u8, g[f"u8_{case}"]), case
        t = IO.to_tensor_normalize(u8, bool(flip))
        assert hashlib.sha256(np.ascontiguousarray(t).tobytes()).digest() == g[f"tensor_sha256_{case}"].tobytes(), case
        if case not in (3, 5):                         # those two boxes were forced (skipped passes), not drawn
            torch.manual_seed(case)
            assert P.random_resized_crop_params(H, W) == (i, j, h, w)
            assert bool(torch.rand(1) < 0.5) == bool(flip)
    for s in range(30):
        torch.manual_seed(s); a = P.random_resized_crop_params(40 + 37 * s, 2000 - 61 * s)
        torch.manual_seed(s); b = IO.random_resized_crop_params(40 + 37 * s, 2000 - 61 * s)
        assert a == b, s
    for sd in (0, 4, 6):
        H, W = (int(v) for v in g[f"params{sd}"][:2])
        t = IO.transform_eval(IO.synthetic_frame(sd, H, W))
        assert hashlib.sha256(np.ascontiguousarray(t).tobytes()).digest() == g[f"eval_sha256_{sd}"].tobytes(), sd
    ks, b, k = IO.precompute_coeffs(1155, 0.0, 1155.0, 224)
    assert ks == 23 and int(b[:, 1].max()) <= ks and abs(int(k[100].sum()) - (1 << 22)) <= ks


def test_augment_oracle_matches_torchvision_golden():
    """Fine-tune TRAIN augmentations (classification/data/transforms.py:233-258): oracle/input_oracle.py against
    tests/golden/augment_golden.npz = what the INSTALLED Pillow + torchvision Compose produced under torch.manual_seed(321)
    (oracle/make_augment_golden.py, which also pins every stage separately and both HSV conversions exhaustively).  Same
    random draws; uint8 result identical except where the float32 conv2d of torchvision's blur rounds the other way
    (<= 1 grey level, on the number of values the fixture recorded)."""
    import numpy as np
    from oracle import input_oracle as IO
    g = np.load(os.path.join(ROOT, "tests", "golden", "augment_golden.npz"))
    torch.manual_seed(int(g["seed"][0]))
    bad = 0
    for sd in (0, 1, 4, 6, 8, 9):
        H, W = (int(v) for v in g[f"frame_{sd}"])
        t, u8, p = IO.transform_finetune_train(IO.synthetic_frame(sd, H, W))
        ip, fp = g[f"params_{sd}"], g[f"fparams_{sd}"]
        assert p["order"] == [int(v) for v in ip[:4]] and IO.hue_shift_u8(p["hue"]) == int(ip[4])
        assert (int(p["hflip"]), int(p["vflip"])) == (int(ip[5]), int(ip[6]))
        assert [p["brightness"], p["contrast"], p["saturation"], p["hue"], p["sigma"], p["angle"]] == [float(v) for v in fp]
        d = np.abs(u8.astype(int) - g[f"u8_{sd}"].astype(int))
        assert d.max() <= 1, (sd, d.max())
        bad += int((d > 0).sum())
        assert t.dtype == np.float32 and np.array_equal(t.view(np.uint32), IO.to_tensor_normalize(u8, False).view(np.uint32))
    assert bad == int(g["blur_mismatch_values"][0])
    # stage properties: a zero hue shift still goes through the lossy HSV round trip; identity / 180 rotations are exact
    img = IO.synthetic_frame(5, 32, 32)
    assert np.array_equal(IO.rotate_nearest(img, 0.0), img) and np.array_equal(IO.rotate_nearest(img, 180.0), img[::-1, ::-1])
    assert np.array_equal(IO.rotate_nearest(img, 90.0), np.rot90(img, 1))
    assert np.array_equal(IO.gaussian_blur_u8(img, IO.gaussian_kernel1d(25, 0.001)), img)          # delta kernel
    assert np.array_equal(IO.adjust_brightness(img, 1.0), img) and np.array_equal(IO.adjust_saturation(img, 1.0), img)


def test_host_pack_rects_matches_numpy_slices_on_every_thread_count():
    """ssl4p_host_pack_rects (the host side of the frame upload, no CUDA): ragged crop rectangles of ragged frames, offset
    views, one-row and one-column boxes, byte-balanced over 1..16 threads, against numpy slicing."""
    import numpy as np
    from ssl4polyp_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(5)
    frames = [rng.integers(0, 256, (int(rng.integers(200, 600)), int(rng.integers(200, 700)), 3), dtype=np.uint8) for _ in range(24)]
    frames[2] = frames[2][7:, :]                                   # a view that does not start at its buffer
    params = []
    for fr in frames:
        H, W = fr.shape[:2]
        h, w = int(rng.integers(1, H + 1)), int(rng.integers(1, W + 1))
        params.append((int(rng.integers(0, H - h + 1)), int(rng.integers(0, W - w + 1)), h, w))
    params[0] = (0, 0) + frames[0].shape[:2]                       # a whole frame: one memcpy
    params[1] = (3, 5, 1, frames[1].shape[1] - 5)                  # one row
    params[4] = (0, 9, frames[4].shape[0], 1)                      # one column
    sizes = [p[2] * p[3] * 3 for p in params]
    offs = np.cumsum([0] + sizes).astype(np.int64)
    total = int(offs[-1])
    assert total > (1 << 20)                                       # beyond the single-thread cut-off
    want = np.concatenate([fr[t:t + h, l:l + w].reshape(-1) for fr, (t, l, h, w) in zip(frames, params)])
    src = np.array([fr.__array_interface__["data"][0] + t * fr.strides[0] + l * 3 for fr, (t, l, h, w) in zip(frames, params)], np.uint64)
    pitch = np.array([fr.strides[0] for fr in frames], np.int64)
    nrows = np.array([p[2] for p in params], np.int64)
    row_bytes = np.array([p[3] * 3 for p in params], np.int64)
    o = np.ascontiguousarray(offs[:-1])
    for threads in (1, 2, 3, 7, 8, 16):
        dst = np.full(total + 64, 7, np.uint8)
        rc = lib.ssl4p_host_pack_rects(src.ctypes.data, pitch.ctypes.data, nrows.ctypes.data, row_bytes.ctypes.data, dst.ctypes.data,
                                       o.ctypes.data, len(frames), threads)
        assert rc == 0, lib.ssl4p_last_error()
        assert np.array_equal(dst[:total], want), threads
        assert (dst[total:] == 7).all()
    assert lib.ssl4p_host_pack_rects(src.ctypes.data, pitch.ctypes.data, nrows.ctypes.data, row_bytes.ctypes.data, dst.ctypes.data,
                                     o.ctypes.data, 0, 4) == 0     # empty batch
    assert lib.ssl4p_host_pack_rects(None, pitch.ctypes.data, nrows.ctypes.data, row_bytes.ctypes.data, dst.ctypes.data,
                                     o.ctypes.data, 3, 4) != 0     # null pointer: an error, not a crash


def test_train_augment_draws_follow_torchvision_call_for_call():
    """TrainAugment.draw_augment (one reused scratch tensor, .tolist()) against the installed torchvision transforms drawing
    in the Compose's order under the same seed (classification/data/transforms.py:236-244): identical values, identical
    generator consumption (the state after 500 images is the same)."""
    import types
    import torchvision.transforms as T
    from ssl4polyp_b200 import input_pipeline as ip
    me = types.SimpleNamespace(brightness=(0.6, 1.4), contrast=(0.5, 1.5), saturation=(0.75, 1.25), hue=(-0.01, 0.01),
                               sigma=(0.001, 2.0), degrees=180.0, flip_p=0.5, _scratch={})
    torch.manual_seed(77)
    mine = [ip.TrainAugment.draw_augment(me) for _ in range(500)]
    state_mine = torch.get_rng_state()
    cj = T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01)
    torch.manual_seed(77)
    for a in mine:
        order, b, c, s, h = T.ColorJitter.get_params(cj.brightness, cj.contrast, cj.saturation, cj.hue)
        sigma = T.GaussianBlur.get_params(0.001, 2.0)
        hflip = bool(torch.rand(1) < 0.5)                        # RandomHorizontalFlip.forward
        vflip = bool(torch.rand(1) < 0.5)                        # RandomVerticalFlip.forward
        angle = T.RandomRotation.get_params([-180.0, 180.0])
        assert a == dict(order=order.tolist(), brightness=b, contrast=c, saturation=s, hue=h, sigma=sigma, hflip=hflip, vflip=vflip,
                         angle=angle)
    assert torch.equal(torch.get_rng_state(), state_mine)
