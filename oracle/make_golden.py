"""TEST INFRASTRUCTURE ONLY — pins the oracle against OUTPUTS OF THE REFERENCE ITSELF.

Run in the build container (the only place `/root/reference` exists):

    python oracle/make_golden.py            # writes tests/golden/*.pt, prints the agreement table

What it does
  1. installs `oracle/timm_shim.py` (restated timm 0.4.12, the un-vendored dependency) and imports
     `/root/reference/src/ssl4polyp/models/mae/models_mae.py` UNMODIFIED as a package module so its
     relative import of `util/pos_embed.py` resolves;
  2. constructs the reference `mae_vit_base_patch16(norm_pix_loss=…)` under `torch.manual_seed(seed)` and
     REQUIRES `oracle.mae_oracle.init_state_dict(seed)` to reproduce its state dict bit-for-bit;
  3. runs the reference forward/backward (noise drawn by the reference's own `torch.rand`, reproduced by
     seeding immediately before the call) and REQUIRES the functional oracle to agree: masks and
     ids_restore exactly, loss/pred/grads to fp32 round-off;
  4. cross-checks both against the independent `transformers` ViT-MAE implementation with an explicit
     noise tensor (substitute for the missing timm pin);
  5. commits compact golden vectors (seeds + integer outputs + losses + sub-sampled pred/grads) under
     `tests/golden/`.  Weights and images are NOT stored: they are regenerated from the seeds with the
     same torch CPU generator on the GPU box (same image, same torch build).
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import mae_oracle as O          # noqa: E402
from oracle import timm_shim                # noqa: E402

REF_SRC = "/root/reference/src"
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

GRAD_KEYS = [
    "patch_embed.proj.weight", "blocks.0.attn.qkv.weight", "blocks.11.mlp.fc2.weight",
    "decoder_blocks.7.attn.proj.weight", "mask_token", "cls_token", "decoder_pred.bias",
    "blocks.5.norm1.weight", "decoder_embed.weight", "decoder_blocks.0.mlp.fc1.bias",
]


def import_reference_models_mae():
    """Import the reference file as `ssl4polyp.models.mae.models_mae` without executing the (heavy,
    partly broken on this stack) package `__init__`s: stub packages with the right `__path__`."""
    timm_shim.install()
    for name, path in [("ssl4polyp", f"{REF_SRC}/ssl4polyp"),
                       ("ssl4polyp.models", f"{REF_SRC}/ssl4polyp/models"),
                       ("ssl4polyp.models.mae", f"{REF_SRC}/ssl4polyp/models/mae"),
                       ("ssl4polyp.models.mae.util", f"{REF_SRC}/ssl4polyp/models/mae/util")]:
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = [path]
            sys.modules[name] = m
    sys.dont_write_bytecode = True
    return importlib.import_module("ssl4polyp.models.mae.models_mae")


def subsample(t: torch.Tensor, n: int = 64) -> torch.Tensor:
    """Deterministic strided sub-sample of a flattened tensor (keeps fixtures small)."""
    f = t.detach().reshape(-1)
    step = max(1, f.numel() // n)
    return f[::step][:n].clone()


def synth_imgs(B: int, seed: int, size: int = 224) -> torch.Tensor:
    g = torch.Generator().manual_seed(1000 + seed)
    return torch.randn(B, 3, size, size, generator=g)


def noise_for(B: int, L: int, seed: int) -> torch.Tensor:
    """What the reference's `torch.rand(N, L)` (models_mae.py:132) returns when the global CPU
    generator is seeded with `seed` immediately before `forward`."""
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        return torch.rand(B, L)


def run_case(ref_mod, cfg_name, cfg: O.MAEConfig, ctor_kwargs, B, seed, mask_ratio, hf_check):
    # --- reference model, reference init --------------------------------------------------------
    torch.manual_seed(seed)
    if cfg_name == "base":
        ref = ref_mod.mae_vit_base_patch16(norm_pix_loss=cfg.norm_pix_loss)
    else:
        from functools import partial
        ref = ref_mod.MaskedAutoencoderViT(norm_layer=partial(torch.nn.LayerNorm, eps=cfg.ln_eps),
                                           norm_pix_loss=cfg.norm_pix_loss, **ctor_kwargs)
    ref.train()
    ref_sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}

    # --- oracle init must reproduce it bit-for-bit ----------------------------------------------
    sd = O.init_state_dict(cfg, seed=seed)
    assert set(sd) == set(ref_sd), (set(sd) ^ set(ref_sd))
    for k in ref_sd:
        assert sd[k].shape == ref_sd[k].shape, k
        assert torch.equal(sd[k], ref_sd[k]), f"init mismatch at {k}"

    imgs = synth_imgs(B, seed, cfg.img_size)
    noise = noise_for(B, cfg.num_patches, seed)

    # --- reference forward/backward --------------------------------------------------------------
    torch.manual_seed(seed)
    loss_r, pred_r, mask_r = ref(imgs, mask_ratio=mask_ratio)
    loss_r.backward()
    grads_r = {k: p.grad.detach().clone() for k, p in ref.named_parameters() if p.grad is not None}
    torch.manual_seed(seed)
    with torch.no_grad():
        latent_r, mask_r2, ids_restore_r = ref.forward_encoder(imgs, mask_ratio)
    assert torch.equal(mask_r, mask_r2)

    # --- oracle forward/backward -----------------------------------------------------------------
    sd_g = {k: v.clone().requires_grad_(k not in ("pos_embed", "decoder_pos_embed")) for k, v in sd.items()}
    loss_o, pred_o, mask_o = O.mae_forward(sd_g, imgs, noise, mask_ratio, cfg)
    loss_o.backward()
    _, _, ids_restore_o = O.masking_from_noise(noise, mask_ratio)
    assert torch.equal(mask_o, mask_r), "mask mismatch oracle vs reference"
    assert torch.equal(ids_restore_o, ids_restore_r), "ids_restore mismatch oracle vs reference"
    rel = lambda a, b: ((a - b).norm() / b.norm().clamp_min(1e-30)).item()
    e_loss = abs(loss_o.item() - loss_r.item()) / abs(loss_r.item())
    e_pred = rel(pred_o, pred_r)
    e_grad = max(rel(sd_g[k].grad, grads_r[k]) for k in grads_r)
    assert e_loss < 1e-5 and e_pred < 1e-4 and e_grad < 1e-3, (e_loss, e_pred, e_grad)

    # --- independent cross-check: transformers ViT-MAE with explicit noise -----------------------
    e_hf = None
    if hf_check:
        e_hf = hf_crosscheck(sd, cfg, imgs, noise, mask_ratio, loss_r, pred_r, mask_r, ids_restore_r)

    print(f"[{cfg_name} B={B} seed={seed} mr={mask_ratio} normpix={cfg.norm_pix_loss}] loss={loss_r.item():.6f} "
          f"oracle-vs-ref: loss {e_loss:.2e} pred {e_pred:.2e} grad {e_grad:.2e}  hf: {e_hf}")

    keys = [k for k in GRAD_KEYS if k in grads_r]
    return {
        "cfg_name": cfg_name, "cfg": dict(cfg.__dict__), "B": B, "seed": seed, "mask_ratio": mask_ratio,
        "noise": noise, "ids_restore": ids_restore_r.clone(), "mask": mask_r.clone(),
        "loss": loss_r.detach().clone(), "pred_sample": subsample(pred_r, 256),
        "pred_norm": pred_r.detach().norm().clone(), "latent_sample": subsample(latent_r, 256),
        "grad_norm": {k: grads_r[k].norm().clone() for k in grads_r},
        "grad_sample": {k: subsample(grads_r[k], 64) for k in keys},
        "imgs_sample": subsample(imgs, 16), "w_sample": subsample(ref_sd["blocks.0.attn.qkv.weight"], 16),
    }


def hf_crosscheck(sd, cfg, imgs, noise, mask_ratio, loss_r, pred_r, mask_r, ids_restore_r):
    from transformers import ViTMAEConfig, ViTMAEForPreTraining
    hc = ViTMAEConfig(hidden_size=cfg.embed_dim, num_hidden_layers=cfg.depth, num_attention_heads=cfg.num_heads,
                      intermediate_size=int(cfg.embed_dim * cfg.mlp_ratio), hidden_act="gelu",
                      hidden_dropout_prob=0., attention_probs_dropout_prob=0., layer_norm_eps=cfg.ln_eps,
                      image_size=cfg.img_size, patch_size=cfg.patch_size, num_channels=cfg.in_chans, qkv_bias=True,
                      decoder_num_attention_heads=cfg.decoder_num_heads, decoder_hidden_size=cfg.decoder_embed_dim,
                      decoder_num_hidden_layers=cfg.decoder_depth,
                      decoder_intermediate_size=int(cfg.decoder_embed_dim * cfg.mlp_ratio),
                      mask_ratio=mask_ratio, norm_pix_loss=cfg.norm_pix_loss, attn_implementation="eager")
    hf = ViTMAEForPreTraining(hc).eval()
    hsd = hf.state_dict()
    D, Dd = cfg.embed_dim, cfg.decoder_embed_dim

    def put(k, v):
        assert hsd[k].shape == v.shape, (k, hsd[k].shape, v.shape)
        hsd[k] = v.clone()

    put("vit.embeddings.cls_token", sd["cls_token"])
    put("vit.embeddings.position_embeddings", sd["pos_embed"])
    put("vit.embeddings.patch_embeddings.projection.weight", sd["patch_embed.proj.weight"])
    put("vit.embeddings.patch_embeddings.projection.bias", sd["patch_embed.proj.bias"])

    def put_block(dst, src, dim):
        qw, qb = sd[src + "attn.qkv.weight"], sd[src + "attn.qkv.bias"]
        for i, n in enumerate(("query", "key", "value")):
            put(dst + f"attention.attention.{n}.weight", qw[i * dim:(i + 1) * dim])
            put(dst + f"attention.attention.{n}.bias", qb[i * dim:(i + 1) * dim])
        put(dst + "attention.output.dense.weight", sd[src + "attn.proj.weight"])
        put(dst + "attention.output.dense.bias", sd[src + "attn.proj.bias"])
        put(dst + "layernorm_before.weight", sd[src + "norm1.weight"]); put(dst + "layernorm_before.bias", sd[src + "norm1.bias"])
        put(dst + "layernorm_after.weight", sd[src + "norm2.weight"]); put(dst + "layernorm_after.bias", sd[src + "norm2.bias"])
        put(dst + "intermediate.dense.weight", sd[src + "mlp.fc1.weight"]); put(dst + "intermediate.dense.bias", sd[src + "mlp.fc1.bias"])
        put(dst + "output.dense.weight", sd[src + "mlp.fc2.weight"]); put(dst + "output.dense.bias", sd[src + "mlp.fc2.bias"])

    for i in range(cfg.depth):
        put_block(f"vit.encoder.layer.{i}.", f"blocks.{i}.", D)
    put("vit.layernorm.weight", sd["norm.weight"]); put("vit.layernorm.bias", sd["norm.bias"])
    put("decoder.mask_token", sd["mask_token"])
    put("decoder.decoder_pos_embed", sd["decoder_pos_embed"])
    put("decoder.decoder_embed.weight", sd["decoder_embed.weight"]); put("decoder.decoder_embed.bias", sd["decoder_embed.bias"])
    for i in range(cfg.decoder_depth):
        put_block(f"decoder.decoder_layers.{i}.", f"decoder_blocks.{i}.", Dd)
    put("decoder.decoder_norm.weight", sd["decoder_norm.weight"]); put("decoder.decoder_norm.bias", sd["decoder_norm.bias"])
    put("decoder.decoder_pred.weight", sd["decoder_pred.weight"]); put("decoder.decoder_pred.bias", sd["decoder_pred.bias"])
    hf.load_state_dict(hsd)
    with torch.no_grad():
        out = hf(pixel_values=imgs, noise=noise)
    assert torch.equal(out.mask, mask_r), "HF mask mismatch"
    assert torch.equal(out.ids_restore, ids_restore_r), "HF ids_restore mismatch"
    e_loss = abs(out.loss.item() - loss_r.item()) / abs(loss_r.item())
    e_pred = ((out.logits - pred_r).norm() / pred_r.norm()).item()
    assert e_loss < 1e-5 and e_pred < 1e-4, (e_loss, e_pred)
    return f"loss {e_loss:.2e} pred {e_pred:.2e}"


def crafted_ties_case():
    """Noise rows with duplicated values: the documented tie rule is stable ascending (lower index first)."""
    L = 196
    g = torch.Generator().manual_seed(7)
    noise = torch.rand(4, L, generator=g)
    noise[0, 10] = noise[0, 3]; noise[0, 150] = noise[0, 3]
    noise[1, :] = 0.5                                         # all equal → identity permutation
    noise[2, ::2] = noise[2, 1::2]                            # pairwise ties
    noise[3, 195] = noise[3, 0]
    ids_keep, mask, ids_restore = O.masking_from_noise(noise, 0.75)
    assert torch.equal(ids_restore[1], torch.arange(L))
    return {"noise": noise, "ids_keep": ids_keep, "mask": mask, "ids_restore": ids_restore, "mask_ratio": 0.75}


def main():
    torch.set_num_threads(os.cpu_count() or 8)
    torch.use_deterministic_algorithms(True, warn_only=True)
    from transformers import ViTMAEConfig, ViTMAEForPreTraining   # noqa: F401  (before the timm shim: HF probes find_spec("timm"))
    ref_mod = import_reference_models_mae()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    cases = []
    base_np = O.MAEConfig(norm_pix_loss=True)
    base_raw = O.MAEConfig(norm_pix_loss=False)
    cases.append(run_case(ref_mod, "base", base_np, {}, B=2, seed=0, mask_ratio=0.75, hf_check=True))
    cases.append(run_case(ref_mod, "base", base_np, {}, B=8, seed=1, mask_ratio=0.75, hf_check=False))
    cases.append(run_case(ref_mod, "base", base_raw, {}, B=2, seed=2, mask_ratio=0.75, hf_check=True))
    cases.append(run_case(ref_mod, "base", base_np, {}, B=2, seed=3, mask_ratio=0.6, hf_check=False))
    tiny_kwargs = dict(img_size=64, patch_size=16, in_chans=3, embed_dim=64, depth=2, num_heads=2,
                       decoder_embed_dim=32, decoder_depth=1, decoder_num_heads=2, mlp_ratio=4.)
    tiny = O.MAEConfig(img_size=64, embed_dim=64, depth=2, num_heads=2, decoder_embed_dim=32, decoder_depth=1,
                       decoder_num_heads=2, norm_pix_loss=True)
    cases.append(run_case(ref_mod, "tiny", tiny, tiny_kwargs, B=3, seed=4, mask_ratio=0.75, hf_check=True))
    out = {"cases": cases, "ties": crafted_ties_case(), "torch": str(torch.__version__),
           "generator": "oracle/make_golden.py", "reference": "irconde/SSL4POLYP models_mae.py (unmodified) + timm 0.4.12 shim"}
    path = os.path.join(GOLDEN_DIR, "mae_golden.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
