"""GPU parity of the drop-in model (ssl4polyp_b200.models_mae / models) against
  (1) the committed golden vectors produced by the UNMODIFIED reference (tests/golden/mae_golden.pt), and
  (2) the CPU oracle (oracle/mae_oracle.py) run on the same weights, images and noise.
Tolerances are BASELINE.json's: masking indices / ids_restore / mask bit-exact; loss rel 1e-2; gradient cosine > 0.999.
"""
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import mae_oracle as O  # noqa: E402  (checker only)

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "mae_golden.pt")
LOSS_RTOL = 1e-2
GRAD_COS = 0.999


def dev():
    return torch.device("cuda:0")


def synth_imgs(B, seed, size=224):
    g = torch.Generator().manual_seed(1000 + seed)
    return torch.randn(B, 3, size, size, generator=g)


def subsample(t, n=64):
    f = t.detach().reshape(-1)
    step = max(1, f.numel() // n)
    return f[::step][:n].clone()


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm()).clamp_min(1e-300)).item()


def build_base(seed, norm_pix):
    from ssl4polyp_b200 import models_mae
    torch.manual_seed(seed)
    m = models_mae.mae_vit_base_patch16(norm_pix_loss=norm_pix)
    return m.to(dev()).train()


@pytest.fixture(scope="module")
def golden():
    return torch.load(GOLDEN)


@pytest.mark.parametrize("idx", [0, 1, 2, 3])
def test_mae_base_against_reference_golden(golden, idx):
    c = golden["cases"][idx]
    assert c["cfg_name"] == "base"
    B, seed, mr = c["B"], c["seed"], c["mask_ratio"]
    m = build_base(seed, c["cfg"]["norm_pix_loss"])
    # same init as the reference constructor under the same seed (bit-exact on CPU, checked in test_host.py)
    assert torch.equal(subsample(m.state_dict()["blocks.0.attn.qkv.weight"].cpu(), 16), c["w_sample"])
    imgs = synth_imgs(B, seed)
    assert torch.equal(subsample(imgs, 16), c["imgs_sample"])
    loss, pred, mask, ids_restore = m.forward_with_noise(imgs.to(dev()), c["noise"].to(dev()), mr)
    assert ids_restore.dtype == torch.int64 and mask.dtype == torch.float32
    assert torch.equal(ids_restore.cpu(), c["ids_restore"]), "ids_restore must be bit-exact"
    assert torch.equal(mask.cpu(), c["mask"]), "mask must be bit-exact"
    assert pred.shape == (B, 196, 768)
    rel = abs(loss.item() - c["loss"].item()) / abs(c["loss"].item())
    assert rel < LOSS_RTOL, (loss.item(), c["loss"].item())
    ps = subsample(pred.cpu(), 256)
    assert cos(ps, c["pred_sample"]) > 0.995
    assert abs(pred.float().norm().item() - c["pred_norm"].item()) / c["pred_norm"].item() < 2e-2
    loss.backward()
    grads = {k: p.grad for k, p in m.named_parameters() if p.grad is not None}
    assert set(grads) == set(c["grad_norm"]), set(grads) ^ set(c["grad_norm"])
    for k, gs in c["grad_sample"].items():
        assert cos(subsample(grads[k].cpu(), 64), gs) > 0.99, k
    bad = []
    for k, gn in c["grad_norm"].items():
        r = abs(grads[k].norm().item() - gn.item()) / max(gn.item(), 1e-12)
        if r > 5e-2:
            bad.append((k, r))
    assert not bad, bad[:8]


def _oracle_run(cfg, sd, imgs, noise, mr):
    sd_g = {k: v.clone().requires_grad_(k not in ("pos_embed", "decoder_pos_embed")) for k, v in sd.items()}
    loss, pred, mask = O.mae_forward(sd_g, imgs, noise, mr, cfg)
    loss.backward()
    _, _, ids_restore = O.masking_from_noise(noise, mr)
    return loss.detach(), pred.detach(), mask, ids_restore, {k: v.grad for k, v in sd_g.items() if v.grad is not None}


def _check_against_oracle(m, cfg, B, seed, mr, size):
    m.zero_grad(set_to_none=True)
    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    imgs = synth_imgs(B, seed, size)
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        noise = torch.rand(B, cfg.num_patches)
    loss_o, pred_o, mask_o, ids_o, grads_o = _oracle_run(cfg, sd, imgs, noise, mr)
    loss, pred, mask, ids_restore = m.forward_with_noise(imgs.to(dev()), noise.to(dev()), mr)
    assert torch.equal(ids_restore.cpu(), ids_o) and torch.equal(mask.cpu(), mask_o)
    assert abs(loss.item() - loss_o.item()) / abs(loss_o.item()) < LOSS_RTOL
    assert cos(pred.cpu(), pred_o) > 0.999
    loss.backward()
    worst = 1.0
    for k, g in grads_o.items():
        p = dict(m.named_parameters())[k]
        assert p.grad is not None, k
        c = cos(p.grad.cpu(), g)
        worst = min(worst, c)
        assert c > GRAD_COS, (k, c)
    return worst


def test_mae_base_all_grads_against_oracle():
    m = build_base(5, True)
    _check_against_oracle(m, O.MAEConfig(norm_pix_loss=True), B=2, seed=5, mr=0.75, size=224)


def test_mae_small_config_against_oracle():
    """64x64 images, 16 patches, D=64 (2 heads of 32), decoder D=64: ragged tiles everywhere (M = 15, 51 rows)."""
    from functools import partial
    from ssl4polyp_b200 import models_mae
    torch.manual_seed(11)
    kw = dict(img_size=64, patch_size=16, in_chans=3, embed_dim=64, depth=2, num_heads=2, decoder_embed_dim=64,
              decoder_depth=2, decoder_num_heads=2, mlp_ratio=4.)
    m = models_mae.MaskedAutoencoderViT(norm_layer=partial(torch.nn.LayerNorm, eps=1e-6), norm_pix_loss=True, **kw).to(dev())
    with torch.no_grad():                      # non-trivial LayerNorm / bias / token values
        for p in m.parameters():
            if p.requires_grad and p.ndim == 1:
                p.add_(0.1 * torch.randn_like(p))
    cfg = O.MAEConfig(img_size=64, embed_dim=64, depth=2, num_heads=2, decoder_embed_dim=64, decoder_depth=2,
                      decoder_num_heads=2, norm_pix_loss=True)
    _check_against_oracle(m, cfg, B=3, seed=4, mr=0.75, size=64)
    _check_against_oracle(m, cfg, B=5, seed=6, mr=0.5, size=64)
    # edges of random_masking (models_mae.py:123-148): a single image; nothing masked away from the encoder except one patch
    # (len_keep = L - 1); almost everything masked (len_keep = 1: the encoder sees the cls token and one patch)
    _check_against_oracle(m, cfg, B=1, seed=8, mr=0.75, size=64)
    _check_against_oracle(m, cfg, B=2, seed=9, mr=0.05, size=64)
    _check_against_oracle(m, cfg, B=4, seed=10, mr=0.9, size=64)


def test_forward_draws_the_reference_noise():
    """`forward` must consume exactly one torch.rand(N, L) from the device generator (models_mae.py:132)."""
    m = build_base(0, True)
    imgs = synth_imgs(2, 0).to(dev())
    torch.manual_seed(123)
    noise = torch.rand(2, 196, device=dev())
    after = torch.rand(4, device=dev())
    torch.manual_seed(123)
    with torch.no_grad():
        loss, pred, mask = m(imgs, mask_ratio=0.75)
    assert torch.equal(torch.rand(4, device=dev()), after)
    ids_keep, mask_o, _ = O.masking_from_noise(noise.cpu(), 0.75)
    assert torch.equal(mask.cpu(), mask_o)
    assert mask.sum().item() == 2 * 147 and loss.dim() == 0 and not loss.requires_grad


def test_piecewise_api_matches_fused_forward():
    m = build_base(1, True)
    imgs = synth_imgs(2, 1).to(dev())
    torch.manual_seed(7)
    loss, pred, mask = m(imgs, 0.75)
    loss.backward()
    g_fused = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
    m.zero_grad(set_to_none=True)
    torch.manual_seed(7)
    latent, mask2, ids_restore = m.forward_encoder(imgs, 0.75)
    assert latent.shape == (2, 50, 768) and latent.dtype == torch.float32
    pred2 = m.forward_decoder(latent, ids_restore)
    loss2 = m.forward_loss(imgs, pred2, mask2)
    assert torch.equal(mask, mask2)
    assert abs(loss.item() - loss2.item()) / abs(loss.item()) < 1e-3
    assert cos(pred, pred2) > 0.9999
    loss2.backward()
    for k, g in g_fused.items():
        g2 = dict(m.named_parameters())[k].grad
        assert g2 is not None and cos(g, g2) > 0.999, k
    # patchify / unpatchify / random_masking
    x = m.patchify(imgs)
    assert torch.equal(x.cpu(), O.patchify(imgs.cpu()))
    assert torch.equal(m.unpatchify(x), imgs)
    torch.manual_seed(3)
    xm, mk, ir = m.random_masking(x, 0.75)
    torch.manual_seed(3)
    nz = torch.rand(2, 196, device=dev())
    ik, mk_o, ir_o = O.masking_from_noise(nz.cpu(), 0.75)
    assert torch.equal(ir.cpu(), ir_o) and torch.equal(mk.cpu(), mk_o)
    assert torch.equal(xm.cpu(), torch.gather(x.cpu(), 1, ik.unsqueeze(-1).expand(-1, -1, 768)))


def test_classifier_against_oracle():
    from ssl4polyp_b200 import models
    torch.manual_seed(2)
    m = models.ViT_from_MAE(None, True, 2, False, None, embed_dim=768, depth=12, num_heads=12, out_token="cls").to(dev()).train()
    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    B = 4
    imgs = synth_imgs(B, 9)
    target = torch.tensor([0, 1, 1, 0])
    sd_g = {k: v.clone().requires_grad_(k not in ("pos_embed", "decoder_pos_embed")) for k, v in sd.items()}
    logits_o = O.classifier_forward(sd_g, imgs, O.MAE_BASE, "cls", True)
    loss_o = O.binary_polyp_loss(logits_o, target)
    loss_o.backward()
    logits = m(imgs.to(dev()))
    assert logits.shape == (B, 2) and logits.dtype == torch.float32
    assert (logits.cpu() - logits_o.detach()).abs().max().item() < 3e-2 * max(1.0, logits_o.abs().max().item())
    loss = O.binary_polyp_loss(logits, target.to(dev()))          # 4 scalars of glue, stays in torch
    assert abs(loss.item() - loss_o.item()) / abs(loss_o.item()) < LOSS_RTOL
    loss.backward()
    for k, v in sd_g.items():
        if v.grad is None:
            continue
        g = dict(m.named_parameters())[k].grad
        assert g is not None, k
        assert cos(g.cpu(), v.grad) > 0.995, (k, cos(g.cpu(), v.grad))


def test_vit_from_any_against_oracle():
    """timm-shaped ViT with a LEARNED pos_embed (reference models.py:26-140): logits, BCE loss and every gradient,
    including d pos_embed (sum over images) and d cls_token, against the CPU oracle on the same weights."""
    from ssl4polyp_b200 import models, utils
    torch.manual_seed(5)
    m = utils.get_ImageNet_or_random_ViT(True, 2, False, None, False, out_token="spatial").to(dev()).train()
    assert isinstance(m, models.VisionTransformer_from_Any) and isinstance(m.blocks, torch.nn.Sequential)
    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    B = 3
    imgs = synth_imgs(B, 11)
    target = torch.tensor([1, 0, 1])
    sd_g = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    logits_o = O.classifier_forward(sd_g, imgs, O.MAE_BASE, "spatial", True)
    loss_o = O.binary_polyp_loss(logits_o, target)
    loss_o.backward()
    logits = m(imgs.to(dev()))
    assert logits.shape == (B, 2) and logits.dtype == torch.float32
    assert (logits.cpu() - logits_o.detach()).abs().max().item() < 3e-2 * max(1.0, logits_o.abs().max().item())
    loss = O.binary_polyp_loss(logits, target.to(dev()))
    assert abs(loss.item() - loss_o.item()) / abs(loss_o.item()) < LOSS_RTOL
    loss.backward()
    params = dict(m.named_parameters())
    assert params["pos_embed"].grad is not None
    for k, v in sd_g.items():
        g = params[k].grad
        assert g is not None, k
        assert cos(g.cpu(), v.grad) > 0.995, (k, cos(g.cpu(), v.grad))
    # the fused fine-tune step (ClassifierEngine) accepts it too
    from ssl4polyp_b200.engine import ClassifierEngine
    eng = ClassifierEngine(m, lr=5e-5, weight_decay=0.05)
    l0 = eng.step(imgs.to(dev()), target.to(dev())).item()
    for _ in range(6):
        l1 = eng.step(imgs.to(dev()), target.to(dev())).item()
    assert l1 < l0, (l0, l1)


def test_classifier_finetune_modes_prune_backward():
    from ssl4polyp_b200 import models
    torch.manual_seed(3)
    m = models.ViT_from_MAE(None, True, 2, False, None, embed_dim=768, depth=12, num_heads=12, out_token="spatial").to(dev()).train()
    for p in m.parameters():
        p.requires_grad_(False)
    for p in m.lin_head.parameters():
        p.requires_grad_(True)
    for p in m.blocks[-1].parameters():
        p.requires_grad_(True)
    imgs = synth_imgs(2, 3).to(dev())
    m(imgs).sum().backward()
    with_grad = {k for k, p in m.named_parameters() if p.grad is not None}
    assert with_grad == {k for k, p in m.named_parameters() if p.requires_grad}
    assert all(k.startswith(("lin_head.", "blocks.11.")) for k in with_grad)
    with torch.no_grad():
        out = m.eval()(imgs)
    assert not out.requires_grad


def _small_mae(seed=0):
    from functools import partial
    from ssl4polyp_b200 import models_mae
    torch.manual_seed(seed)
    kw = dict(img_size=64, patch_size=16, in_chans=3, embed_dim=64, depth=2, num_heads=2, decoder_embed_dim=64,
              decoder_depth=1, decoder_num_heads=2, mlp_ratio=4.)
    return models_mae.MaskedAutoencoderViT(norm_layer=partial(torch.nn.LayerNorm, eps=1e-6), norm_pix_loss=True, **kw).to(dev()).train()


def test_engine_graph_replay_matches_eager_steps():
    """The captured step (noise draw, fwd, bwd, AdamW with device-resident lr/step) must walk the same trajectory as
    the eager step on the same images and noise: losses and weights agree to fp32 round-off.  (With self-drawn noise
    the captured torch.rand consumes the Philox stream at a different offset, so only the loss trend is checked.)"""
    from ssl4polyp_b200.engine import PretrainEngine
    imgs = [synth_imgs(4, s, 64).to(dev()) for s in range(2)]
    out = {}
    for mode in (False, True):
        m = _small_mae(0)
        eng = PretrainEngine(m, lr=1e-3, weight_decay=0.05)
        gen = torch.Generator().manual_seed(123)
        losses = []
        for i in range(6):
            lr = 1e-3 * (1.0 - 0.1 * i)                      # the schedule writes a new lr every iteration
            noise = torch.rand(4, 16, generator=gen).to(dev())
            losses.append(float(eng.step(imgs[i % 2], lr=lr, noise=noise, graph=mode).item()))
        out[mode] = (losses, eng.store.flat.detach().clone(), eng.t)
        assert eng.nonfinite_count() == 0
    (le, we, te), (lg, wg, tg) = out[False], out[True]
    assert te == tg == 6
    assert lg[-1] < lg[0]
    for a, b in zip(le, lg):
        assert abs(a - b) <= 1e-5 * max(1.0, abs(a)), (le, lg)
    assert torch.allclose(we, wg, rtol=1e-5, atol=1e-6)
    m = _small_mae(0)
    eng = PretrainEngine(m, lr=1e-3, weight_decay=0.05)
    own = [float(eng.step(imgs[0], graph=True).item()) for _ in range(8)]     # noise drawn inside the captured step
    assert own[-1] < own[0] and all(x == x for x in own)
    assert len({round(x, 6) for x in own[3:]}) > 1           # replays do advance the noise stream


def test_engine_step_matches_oracle_adamw_step():
    """One engine step == the reference's step restated on the CPU oracle (engine_pretrain.py:50-72 with fp32,
    torch.optim.AdamW over timm's weight-decay groups): updated weights agree (bf16 operand noise only)."""
    from ssl4polyp_b200.engine import PretrainEngine
    m = _small_mae(1)
    cfg = O.MAEConfig(img_size=64, embed_dim=64, depth=2, num_heads=2, decoder_embed_dim=64, decoder_depth=1,
                      decoder_num_heads=2, norm_pix_loss=True)
    sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    params = {k: v.clone().requires_grad_(k not in ("pos_embed", "decoder_pos_embed")) for k, v in sd.items()}
    opt = torch.optim.AdamW(O.weight_decay_groups(params.items(), 0.05), lr=1e-3, betas=(0.9, 0.95))
    imgs = synth_imgs(4, 7, 64)
    noise = torch.rand(4, 16, generator=torch.Generator().manual_seed(5))
    eng = PretrainEngine(m, lr=1e-3, weight_decay=0.05, betas=(0.9, 0.95))
    for _ in range(3):
        loss_o, _, _ = O.mae_forward(params, imgs, noise, 0.75, cfg)
        opt.zero_grad(set_to_none=True)
        loss_o.backward()
        opt.step()
        loss = eng.step(imgs.to(dev()), noise=noise.to(dev()))
        assert abs(loss.item() - loss_o.item()) <= LOSS_RTOL * abs(loss_o.item())
    for k, p in m.named_parameters():
        delta_o = params[k].detach() - sd[k]
        delta = p.detach().cpu() - sd[k]
        if k.endswith("attn.qkv.bias"):
            # the k-bias gradient is analytically zero (softmax is invariant to a shift of all keys), so Adam turns
            # pure round-off into +-lr steps there on both sides: compare the q and v thirds only
            n = delta.numel() // 3
            delta, delta_o = torch.cat([delta[:n], delta[2 * n:]]), torch.cat([delta_o[:n], delta_o[2 * n:]])
        if delta_o.abs().max() > 0:
            assert cos(delta, delta_o) > 0.98, k


def test_main_pretrain_cli_synthetic_run_and_resume(tmp_path):
    """The reference's command line (main_pretrain.py:47-135) end to end on synthetic batches: 2 epochs, checkpoint
    dict layout of util/misc.py:306-318, `last.pth`, log.txt, then --resume continues from the saved epoch."""
    import json
    from ssl4polyp_b200 import main_pretrain as MP
    out = tmp_path / "run"
    argv = ["--model", "mae_vit_base_patch16", "--norm_pix_loss", "--batch_size", "4", "--epochs", "2", "--warmup_epochs", "1",
            "--blr", "1.5e-4", "--data_path", "synthetic:3", "--output_dir", str(out), "--save-freq-epochs", "1"]
    args = MP.get_args_parser().parse_args(argv)
    out.mkdir(parents=True)
    MP.main(args)
    ck = torch.load(out / "ckpts" / "checkpoint-1.pth", map_location="cpu", weights_only=False)
    assert set(ck) == {"model", "optimizer", "epoch", "scaler", "args"} and ck["epoch"] == 1
    assert "decoder_blocks.7.mlp.fc2.weight" in ck["model"] and ck["model"]["patch_embed.proj.weight"].shape == (768, 3, 16, 16)
    assert (out / "ckpts" / "last.pth").is_symlink()
    logs = [json.loads(l) for l in open(out / "log.txt")]
    assert [l["epoch"] for l in logs] == [0, 1] and all(math.isfinite(l["train_loss"]) for l in logs)
    args2 = MP.get_args_parser().parse_args(argv[:argv.index("--epochs")] + ["--epochs", "3"] + argv[argv.index("--epochs") + 2:] +
                                            ["--resume", str(out / "ckpts" / "checkpoint-1.pth")])
    MP.main(args2)
    assert args2.start_epoch == 2 and (out / "ckpts" / "checkpoint-2.pth").exists()


def test_train_classification_cli_synthetic_run(tmp_path):
    """The fine-tune command line (train_classification.py:7706-7982) end to end on synthetic frames: MAE checkpoint ->
    ViT_from_MAE, head+1 regime, BCE on logit1-logit0, cosine schedule, Youden tau on val, best/last checkpoints with the
    reference's payload keys and pointer files; frozen tensors stay bit-identical, trainable ones move."""
    import json
    from ssl4polyp_b200 import models_mae, utils
    from ssl4polyp_b200 import train_classification as TC
    torch.manual_seed(0)
    ck = tmp_path / "mae.pth"
    torch.save({"model": models_mae.mae_vit_base_patch16(norm_pix_loss=True).state_dict()}, ck)
    out = tmp_path / "cls"
    args = TC.get_args(["--pretraining", "Hyperkvasir", "--ss-framework", "mae", "--checkpoint", str(ck), "--finetune-mode", "head+1",
                        "--dataset", "synthetic", "--synthetic-samples", "48", "--batch-size", "16", "--epochs", "3", "--learning-rate", "1e-3",
                        "--scheduler", "cosine", "--early-stop-monitor", "val_auroc", "--log-interval", "1", "--output-dir", str(out), "--seed", "5"])
    lines = []
    summary = TC.run(args, log=lines.append)
    assert summary["epochs_completed"] == 3 and summary["global_step"] == 9 and summary["monitor_metric"] == "val_auroc"
    assert 0.0 <= summary["val"]["auroc"] <= 1.0 and math.isfinite(summary["train_loss"]) and summary["val"]["tau"] is not None
    stem = out / "Hyperkvasir_synthetic_headp1_s5"
    best = torch.load(str(stem) + ".pth", map_location="cpu", weights_only=False)
    need = {"epoch", "model_state_dict", "optimizer_state_dict", "scaler_state_dict", "loss", "val_perf", "val_loss", "val_auroc", "test_perf",
            "test_loss", "test_auroc", "monitor_value", "monitor_metric", "test_monitor_value", "py_state", "np_state", "torch_state"}
    assert need <= set(best) and best["monitor_metric"] == "val_auroc" and best["epoch"] == summary["best_epoch"]
    assert os.path.islink(str(stem) + ".pth") and os.path.exists(str(stem) + "_last.pth")
    last = torch.load(str(stem) + "_last.pth", map_location="cpu", weights_only=False)
    assert last["epoch"] == 3
    assert json.loads(open(str(stem) + ".metrics.json").read())["best_epoch"] == summary["best_epoch"]
    assert any("Train Epoch: 1" in l for l in lines) and os.path.getsize(str(stem) + ".log") > 0
    # head+1: lin_head and blocks.11 moved, everything else is the checkpoint bit for bit
    ref = utils.get_MAE_backbone(str(ck), True, 2, False, None).state_dict()
    sd = last["model_state_dict"]
    moved = {k for k in sd if not torch.equal(sd[k].cpu(), ref[k].cpu())}
    assert moved and all(k.startswith("lin_head.") or k.startswith("blocks.11.") for k in moved), sorted(moved)[:5]
    assert "blocks.11.mlp.fc1.weight" in moved and "blocks.10.mlp.fc1.weight" not in moved and "patch_embed.proj.weight" not in moved
    # the saved weights reproduce the logged validation loss through the module API (eval mode, no_grad)
    model = utils.get_MAE_backbone(str(ck), True, 2, False, None)
    model.load_state_dict(sd)
    model = model.cuda()
    val = TC.evaluate(model, TC.SyntheticFrames(12, 16, 2, 224, 6), torch.device("cuda"), 2,
                      pos_weight=torch.tensor(TC._pos_weight(*TC._class_weights(args, TC.SyntheticFrames(48, 16, 2, 224, 5).labels_list, 2)), device="cuda"))
    assert abs(val["loss"] - last["val_loss"]) <= 2e-3 * max(1.0, abs(last["val_loss"]))
    assert abs(val["auroc"] - last["val_auroc"]) < 1e-6


def test_main_pretrain_image_folder_with_device_transform(tmp_path):
    """`--device_transform`: ImageFolder workers decode to uint8 only, the crop / bicubic resize / flip / normalise of
    main_pretrain.py:156-160 runs on the GPU inside the loop; one epoch over PNG frames of mixed sizes, checkpoint written."""
    import numpy as np
    from PIL import Image
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import main_pretrain as MP
    root = tmp_path / "data"
    for cls in ("a", "b"):
        (root / cls).mkdir(parents=True)
        for k in range(4):
            Image.fromarray(IO.synthetic_frame(k + (10 if cls == "b" else 0), 240 + 40 * k, 320 + 30 * k)).save(root / cls / f"{k}.png")
    out = tmp_path / "run"
    out.mkdir()
    args = MP.get_args_parser().parse_args(["--model", "mae_vit_base_patch16", "--norm_pix_loss", "--batch_size", "4", "--epochs", "1",
                                            "--warmup_epochs", "0", "--blr", "1.5e-4", "--data_path", str(root), "--no_train_dir",
                                            "--num_workers", "2", "--device_transform", "--output_dir", str(out), "--save-freq-epochs", "1"])
    MP.main(args)
    ck = torch.load(out / "ckpts" / "checkpoint-0.pth", map_location="cpu", weights_only=False)
    assert ck["epoch"] == 0 and ck["args"].device_transform is True
    log = [l for l in open(out / "log.txt").read().splitlines() if l.strip()]
    import json
    assert math.isfinite(json.loads(log[-1])["train_loss"])


def test_full_size_batch_256_properties():
    """BASELINE configs[1] (ViT-B/16, B=256, mask 0.75, norm_pix_loss) is too large for the CPU oracle in a test, so the
    full size is checked through size-independent properties: masking is a permutation ranking (bit-exact against
    argsort(argsort(noise)) with ties broken by index), 147 masked patches per image, the loss equals an fp32 torch
    restatement of models_mae.py:198-214 applied to the kernel's own pred/mask (rel 1e-4), rows of the batch are
    independent (the first 8 images alone give the same pred), the run is bit-reproducible, and the decoder_pred.bias
    gradient equals the column sums of d loss / d pred from torch autograd (cosine > 0.9999)."""
    torch.manual_seed(0)
    m = build_base(0, True)
    B = 256
    imgs = torch.randn(B, 3, 224, 224, generator=torch.Generator().manual_seed(1000)).to(dev())
    noise = torch.rand(B, 196, generator=torch.Generator().manual_seed(7)).to(dev())
    noise[3, 10] = noise[3, 150]; noise[5, :] = 0.25                                   # ties: one pair, and a whole row
    for p in m.parameters():
        p.grad = None
    loss, pred, mask, ids_restore = m.forward_with_noise(imgs, noise, 0.75)
    loss.backward()
    torch.cuda.synchronize()
    # (1) masking: stable double argsort, permutation, counts
    ids_shuffle = torch.argsort(noise, dim=1, stable=True)
    want_restore = torch.argsort(ids_shuffle, dim=1, stable=True)
    assert ids_restore.dtype == torch.int64 and torch.equal(ids_restore, want_restore)
    assert torch.equal(ids_restore.sort(dim=1).values, torch.arange(196, device=dev()).expand(B, 196))
    assert torch.equal(mask, (ids_restore >= 49).float()) and torch.equal(mask.sum(1), torch.full((B,), 147.0, device=dev()))
    assert torch.equal(ids_restore[5], torch.arange(196, device=dev()))              # all-equal row keeps index order
    # (2) loss from the kernel's own pred / mask, fp32 torch restatement
    pred_f = pred.detach().float().requires_grad_(True)
    x = imgs.reshape(B, 3, 14, 16, 14, 16)
    target = torch.einsum('nchpwq->nhwpqc', x).reshape(B, 196, 768)
    mean, var = target.mean(dim=-1, keepdim=True), target.var(dim=-1, keepdim=True)
    target = (target - mean) / (var + 1.e-6) ** .5
    l = (((pred_f - target) ** 2).mean(dim=-1) * mask).sum() / mask.sum()
    assert abs(loss.item() - l.item()) <= 1e-4 * abs(l.item()), (loss.item(), l.item())
    # (3) d loss / d pred -> decoder_pred.bias gradient (column sums over every token row)
    l.backward()
    want_db = pred_f.grad.sum(dim=(0, 1)).double()
    got_db = m.decoder_pred.bias.grad.double()
    cos = float(got_db @ want_db / (got_db.norm() * want_db.norm()))
    assert cos > 0.9999 and abs(float(got_db.norm() / want_db.norm()) - 1.0) < 2e-2, (cos, float(got_db.norm()), float(want_db.norm()))
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters() if p.requires_grad)
    # (4) rows are independent: the first 8 images alone
    with torch.no_grad():
        l8, pred8, mask8, ids8 = m.forward_with_noise(imgs[:8].contiguous(), noise[:8].contiguous(), 0.75)
        assert torch.equal(ids8, ids_restore[:8]) and torch.equal(mask8, mask[:8])
        d = (pred8.float() - pred[:8].float()).abs().max().item()
        assert d <= 2e-2 * pred[:8].float().abs().max().item(), d
        # (5) bit-reproducible
        l2, pred2, _, _ = m.forward_with_noise(imgs, noise, 0.75)
        assert torch.equal(pred2, pred) and l2.item() == loss.item()


def test_full_size_classifier_batch_128_properties():
    """BASELINE configs[3] (ViT-B/16 classifier, B=128, T=197) through properties: rows independent (first 4 images alone),
    bit-reproducible, the engine's fused step == the module's autograd path (same loss, same lin_head gradient direction:
    `lin_head.bias.grad` equals the batch sum of d loss / d logits from torch), all gradients finite."""
    from ssl4polyp_b200 import models
    torch.manual_seed(2)
    m = models.ViT_from_MAE(None, True, 2, False, None, embed_dim=768, depth=12, num_heads=12, out_token="cls").to(dev()).train()
    B = 128
    imgs = torch.randn(B, 3, 224, 224, generator=torch.Generator().manual_seed(1000)).to(dev())
    target = torch.randint(0, 2, (B,), generator=torch.Generator().manual_seed(2000)).to(dev())
    logits = m(imgs)
    assert logits.shape == (B, 2) and torch.isfinite(logits).all()
    lg = logits.detach().clone().requires_grad_(True)
    O.binary_polyp_loss(lg, target).backward()
    loss = O.binary_polyp_loss(logits, target)
    loss.backward()
    got, want = m.lin_head.bias.grad.double(), lg.grad.sum(0).double()
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-7), (got, want)
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters() if p.requires_grad)
    with torch.no_grad():
        again = m(imgs)
        assert torch.equal(again, logits)
        first = m(imgs[:4].contiguous())
        assert (first - logits[:4]).abs().max().item() <= 2e-2 * max(1.0, logits.abs().max().item())


def test_evaluate_accepts_uint8_frames_through_the_device_eval_transform():
    """`evaluate` on batches of decoded uint8 frames == `evaluate` on the oracle-transformed fp32 tensors (the reference's
    val / test transform, classification/data/transforms.py:233-258)."""
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import models
    from ssl4polyp_b200 import train_classification as TC
    torch.manual_seed(3)
    m = models.ViT_from_MAE(None, True, 2, False, None, embed_dim=768, depth=12, num_heads=12, out_token="cls").to(dev())
    frames = [IO.synthetic_frame(60 + k, 200 + 31 * k, 260 + 17 * k) for k in range(6)]
    labels = torch.tensor([0, 1, 0, 1, 1, 0])
    as_frames = [(frames[:4], labels[:4]), (frames[4:], labels[4:])]
    tensors = torch.from_numpy(np.stack([IO.transform_eval(f) for f in frames]))
    as_tensors = [(tensors[:4], labels[:4]), (tensors[4:], labels[4:])]
    a = TC.evaluate(m, as_frames, dev(), 2)
    b = TC.evaluate(m, as_tensors, dev(), 2)
    assert a["count"] == 6 and a["loss"] == b["loss"] and a["auroc"] == b["auroc"]
    assert torch.equal(a["_probabilities"], b["_probabilities"])
