"""TEST INFRASTRUCTURE ONLY — CPU/fp32 restatement of the reference's MAE ViT hot path.

This is the *checker*, never the product: only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s `cpu_baseline` / `--impl reference` legs may import it.  Nothing under
`ssl4polyp_b200/` imports it, and the product path raises if its CUDA library is missing.

Pinning status: the reference holds no golden vector, known-answer test or fixture for
this path (SURVEY.md §4, §8c — "parity unpinned" by the reference's own tests).  The
restatement is therefore pinned against OUTPUTS OF THE REFERENCE ITSELF: `make_golden.py`
imports `/root/reference/src/ssl4polyp/models/mae/models_mae.py` unmodified (behind
`timm_shim.install()`), requires this file to agree with it to fp32 round-off (exactly on the
integer outputs), cross-checks both against the independent `transformers` ViT-MAE
implementation, and commits the resulting vectors under `tests/golden/`.

The restatement is *functional*: weights come in as a state dict using the reference's key
names (`models_mae.py:33-58`), and the masking noise is an explicit argument (the reference
draws it inside `random_masking`, `models_mae.py:132`).  All math is plain fp32 torch so
autograd through it yields the oracle gradients.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch
import torch.nn.functional as F


@dataclass(frozen=True)
class MAEConfig:
    img_size: int = 224
    patch_size: int = 16
    in_chans: int = 3
    embed_dim: int = 768
    depth: int = 12
    num_heads: int = 12
    decoder_embed_dim: int = 512
    decoder_depth: int = 8
    decoder_num_heads: int = 16
    mlp_ratio: float = 4.0
    ln_eps: float = 1e-6           # constructors pass partial(LayerNorm, eps=1e-6): models_mae.py:227,235
    norm_pix_loss: bool = False

    @property
    def grid(self):
        return self.img_size // self.patch_size

    @property
    def num_patches(self):
        return self.grid * self.grid


MAE_BASE = MAEConfig()                                            # models_mae.py:223-228
MAE_LARGE = MAEConfig(embed_dim=1024, depth=24, num_heads=16)     # models_mae.py:231-236


# ----------------------------------------------------------------------------------------------
# a2 — sin-cos table (util/pos_embed.py:20-67)
# ----------------------------------------------------------------------------------------------
def sincos_1d(dim: int, pos: np.ndarray) -> np.ndarray:
    """util/pos_embed.py:48-67: omega_i = 10000^(-i/(dim/2)) in fp64; [sin | cos] halves."""
    assert dim % 2 == 0
    omega = np.arange(dim // 2, dtype=np.float64) / (dim / 2.0)
    omega = 1.0 / 10000 ** omega
    ang = np.einsum("m,d->md", pos.reshape(-1), omega)
    return np.concatenate([np.sin(ang), np.cos(ang)], axis=1)


def sincos_pos_embed(dim: int, grid: int, cls_token: bool = True) -> torch.Tensor:
    """util/pos_embed.py:20-45: meshgrid with *w first*; the first half of the channels encodes
    grid[0] (the w coordinate), the second half grid[1] (h); a zero row is prepended for cls.
    Returned as fp32 [1(+)grid², dim] exactly like `torch.from_numpy(...).float()` at models_mae.py:68."""
    gh = np.arange(grid, dtype=np.float32)
    gw = np.arange(grid, dtype=np.float32)
    g = np.stack(np.meshgrid(gw, gh), axis=0).reshape(2, 1, grid, grid)
    emb = np.concatenate([sincos_1d(dim // 2, g[0]), sincos_1d(dim // 2, g[1])], axis=1)
    if cls_token:
        emb = np.concatenate([np.zeros([1, dim]), emb], axis=0)
    return torch.from_numpy(emb).float()


# ----------------------------------------------------------------------------------------------
# a3 / a4 — patchify, unpatchify (models_mae.py:95-121)
# ----------------------------------------------------------------------------------------------
def patchify(imgs: torch.Tensor, p: int = 16) -> torch.Tensor:
    """x[n, h*W+w, (pi*p+qi)*3+c] = imgs[n, c, h*p+pi, w*p+qi]  (models_mae.py:100-107)."""
    n, c, H, W = imgs.shape
    assert H == W and H % p == 0
    h = w = H // p
    x = imgs.reshape(n, c, h, p, w, p).permute(0, 2, 4, 3, 5, 1)
    return x.reshape(n, h * w, p * p * c)


def unpatchify(x: torch.Tensor, p: int = 16, c: int = 3) -> torch.Tensor:
    """Inverse of `patchify` (models_mae.py:114-121)."""
    n, L, _ = x.shape
    h = w = int(L ** .5)
    assert h * w == L
    x = x.reshape(n, h, w, p, p, c).permute(0, 5, 1, 3, 2, 4)
    return x.reshape(n, c, h * p, w * p)


# ----------------------------------------------------------------------------------------------
# a5 — random masking as a function of the noise (models_mae.py:123-148)
# ----------------------------------------------------------------------------------------------
def len_keep_of(L: int, mask_ratio: float) -> int:
    """models_mae.py:130 — evaluated with Python floats, e.g. int(196*(1-0.75)) = 49."""
    return int(L * (1 - mask_ratio))


def masking_from_noise(noise: torch.Tensor, mask_ratio: float):
    """Returns (ids_keep int64 [N,keep], mask fp32 [N,L], ids_restore int64 [N,L]).

    Tie rule: stable ascending sort, lower index first — identical to the
    reference (`torch.argsort`, models_mae.py:135-136) whenever the noise row has no duplicates,
    and what torch's CPU sort does in practice when it has."""
    N, L = noise.shape
    keep = len_keep_of(L, mask_ratio)
    ids_shuffle = torch.argsort(noise, dim=1, stable=True)
    ids_restore = torch.argsort(ids_shuffle, dim=1, stable=True)
    ids_keep = ids_shuffle[:, :keep]
    mask = torch.ones(N, L, dtype=torch.float32, device=noise.device)
    mask[:, :keep] = 0
    mask = torch.gather(mask, 1, ids_restore)
    return ids_keep, mask, ids_restore


# ----------------------------------------------------------------------------------------------
# a11 / a12 — timm 0.4.12 PatchEmbed and Block, functional form
# ----------------------------------------------------------------------------------------------
def patch_embed(sd, imgs, p=16, prefix="patch_embed."):
    """Conv2d(k=s=p) + flatten(2).transpose(1,2) (timm PatchEmbed) == patch-vector GEMM with the
    conv weight viewed [D, c*p*p] (column order c, pi, qi)."""
    w, b = sd[prefix + "proj.weight"], sd[prefix + "proj.bias"]
    x = F.conv2d(imgs, w, b, stride=p)
    return x.flatten(2).transpose(1, 2)


def attention(sd, x, num_heads, prefix):
    B, T, C = x.shape
    hd = C // num_heads
    qkv = F.linear(x, sd[prefix + "qkv.weight"], sd[prefix + "qkv.bias"])
    qkv = qkv.reshape(B, T, 3, num_heads, hd).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    a = (q @ k.transpose(-2, -1)) * (hd ** -0.5)
    a = a.softmax(dim=-1)
    o = (a @ v).transpose(1, 2).reshape(B, T, C)
    return F.linear(o, sd[prefix + "proj.weight"], sd[prefix + "proj.bias"])


def block(sd, x, num_heads, prefix, eps):
    C = x.shape[-1]
    h = F.layer_norm(x, (C,), sd[prefix + "norm1.weight"], sd[prefix + "norm1.bias"], eps)
    x = x + attention(sd, h, num_heads, prefix + "attn.")
    h = F.layer_norm(x, (C,), sd[prefix + "norm2.weight"], sd[prefix + "norm2.bias"], eps)
    h = F.linear(h, sd[prefix + "mlp.fc1.weight"], sd[prefix + "mlp.fc1.bias"])
    h = F.gelu(h)                                  # exact erf GELU (nn.GELU default)
    h = F.linear(h, sd[prefix + "mlp.fc2.weight"], sd[prefix + "mlp.fc2.bias"])
    return x + h


# ----------------------------------------------------------------------------------------------
# a6 / a7 / a8 / a9 — encoder, decoder, loss, forward (models_mae.py:150-220)
# ----------------------------------------------------------------------------------------------
def encoder(sd, imgs, noise, mask_ratio, cfg: MAEConfig):
    x = patch_embed(sd, imgs, cfg.patch_size)
    x = x + sd["pos_embed"][:, 1:, :]
    ids_keep, mask, ids_restore = masking_from_noise(noise, mask_ratio)
    x = torch.gather(x, 1, ids_keep.unsqueeze(-1).expand(-1, -1, x.shape[-1]))
    cls = (sd["cls_token"] + sd["pos_embed"][:, :1, :]).expand(x.shape[0], -1, -1)
    x = torch.cat([cls, x], dim=1)
    for i in range(cfg.depth):
        x = block(sd, x, cfg.num_heads, f"blocks.{i}.", cfg.ln_eps)
    x = F.layer_norm(x, (x.shape[-1],), sd["norm.weight"], sd["norm.bias"], cfg.ln_eps)
    return x, mask, ids_restore


def decoder(sd, latent, ids_restore, cfg: MAEConfig):
    x = F.linear(latent, sd["decoder_embed.weight"], sd["decoder_embed.bias"])
    B, Te, Dd = x.shape
    L = ids_restore.shape[1]
    fill = sd["mask_token"].expand(B, L + 1 - Te, Dd)
    body = torch.cat([x[:, 1:, :], fill], dim=1)
    body = torch.gather(body, 1, ids_restore.unsqueeze(-1).expand(-1, -1, Dd))
    x = torch.cat([x[:, :1, :], body], dim=1) + sd["decoder_pos_embed"]
    for i in range(cfg.decoder_depth):
        x = block(sd, x, cfg.decoder_num_heads, f"decoder_blocks.{i}.", cfg.ln_eps)
    x = F.layer_norm(x, (Dd,), sd["decoder_norm.weight"], sd["decoder_norm.bias"], cfg.ln_eps)
    x = F.linear(x, sd["decoder_pred.weight"], sd["decoder_pred.bias"])
    return x[:, 1:, :]


def masked_mse(imgs, pred, mask, norm_pix_loss: bool, p: int = 16):
    """models_mae.py:198-214; `var` is the unbiased one (÷(P−1)), eps 1e-6 inside the sqrt."""
    target = patchify(imgs, p)
    if norm_pix_loss:
        mean = target.mean(dim=-1, keepdim=True)
        var = target.var(dim=-1, keepdim=True)
        target = (target - mean) / (var + 1.e-6) ** .5
    per_patch = ((pred - target) ** 2).mean(dim=-1)
    return (per_patch * mask).sum() / mask.sum()


def mae_forward(sd, imgs, noise, mask_ratio=0.75, cfg: MAEConfig = MAE_BASE):
    latent, mask, ids_restore = encoder(sd, imgs, noise, mask_ratio, cfg)
    pred = decoder(sd, latent, ids_restore, cfg)
    loss = masked_mse(imgs, pred, mask, cfg.norm_pix_loss, cfg.patch_size)
    return loss, pred, mask


# ----------------------------------------------------------------------------------------------
# a13 — ViT_from_MAE classifier forward (models.py:196-222) and a17 — the 2-class loss
# ----------------------------------------------------------------------------------------------
def classifier_forward(sd, imgs, cfg: MAEConfig = MAE_BASE, out_token: str = "cls", head: bool = True):
    x = patch_embed(sd, imgs, cfg.patch_size)
    x = x + sd["pos_embed"][:, 1:, :]
    cls = (sd["cls_token"] + sd["pos_embed"][:, :1, :]).expand(x.shape[0], -1, -1)
    x = torch.cat([cls, x], dim=1)
    for i in range(cfg.depth):
        x = block(sd, x, cfg.num_heads, f"blocks.{i}.", cfg.ln_eps)
    x = F.layer_norm(x, (x.shape[-1],), sd["norm.weight"], sd["norm.bias"], cfg.ln_eps)
    if out_token == "cls":
        x = x[:, 0]
    elif out_token == "spatial":
        x = x[:, 1:].mean(1)
    if head:
        x = F.linear(x, sd["lin_head.weight"], sd["lin_head.bias"])
    return x


def binary_polyp_loss(logits, target, pos_weight=None):
    """train_classification.py:3347-3374: 2-class logits → logit₁−logit₀ → BCE-with-logits."""
    z = logits[:, 1] - logits[:, 0]
    return F.binary_cross_entropy_with_logits(z, target.float(), pos_weight=pos_weight)


def multiclass_loss(logits, target, class_weights=None):
    """train_classification.py:6103 + :3374: `nn.CrossEntropyLoss(weight=class_weights)(logits, targets)`."""
    return F.cross_entropy(logits, target, weight=class_weights)


# ----------------------------------------------------------------------------------------------
# a1 — initialisation, restated (models_mae.py:65-93) so the oracle can build weights without
# the reference tree (the GPU box has no /root/reference).
# ----------------------------------------------------------------------------------------------
def init_state_dict(cfg: MAEConfig = MAE_BASE, seed: int = 0, head_classes: int | None = None) -> dict:
    """head_classes: also draw `lin_head = nn.Linear(D, head_classes)` AFTER the MAE constructor, which is where
    `ViT_from_MAE.__init__` (models.py:178-180) draws it; the decoder tensors that constructor deletes (:171-175) are
    dropped (`decoder_pos_embed` stays, as in the reference).  make_cls_golden.py pins this bit-for-bit.

    Draw weights with the same RNG call ORDER and distributions as constructing
    `MaskedAutoencoderViT` under `torch.manual_seed(seed)`: module construction first (default
    torch inits consume the generator), then `initialize_weights`.  make_golden.py checks that
    this reproduces the reference constructor bit-for-bit."""
    import torch.nn as nn
    g = torch.random.fork_rng(devices=[])
    with g:
        torch.manual_seed(seed)
        D, Dd, p, c = cfg.embed_dim, cfg.decoder_embed_dim, cfg.patch_size, cfg.in_chans
        hid, hidd = int(D * cfg.mlp_ratio), int(Dd * cfg.mlp_ratio)
        mods = {}                                   # construction order == models_mae.py:33-58
        mods["patch_embed.proj"] = nn.Conv2d(c, D, p, p)
        for i in range(cfg.depth):
            mods[f"blocks.{i}.attn.qkv"] = nn.Linear(D, 3 * D)
            mods[f"blocks.{i}.attn.proj"] = nn.Linear(D, D)
            mods[f"blocks.{i}.mlp.fc1"] = nn.Linear(D, hid)
            mods[f"blocks.{i}.mlp.fc2"] = nn.Linear(hid, D)
        mods["decoder_embed"] = nn.Linear(D, Dd)
        for i in range(cfg.decoder_depth):
            mods[f"decoder_blocks.{i}.attn.qkv"] = nn.Linear(Dd, 3 * Dd)
            mods[f"decoder_blocks.{i}.attn.proj"] = nn.Linear(Dd, Dd)
            mods[f"decoder_blocks.{i}.mlp.fc1"] = nn.Linear(Dd, hidd)
            mods[f"decoder_blocks.{i}.mlp.fc2"] = nn.Linear(hidd, Dd)
        mods["decoder_pred"] = nn.Linear(Dd, p * p * c)
        sd = {}
        # initialize_weights(): patch proj xavier on the [D, c*p*p] view, then the two tokens,
        # then self.apply(_init_weights) in module-tree order (children before self).
        w = mods["patch_embed.proj"].weight.data
        nn.init.xavier_uniform_(w.view(w.shape[0], -1))
        cls = torch.zeros(1, 1, D); mask_tok = torch.zeros(1, 1, Dd)
        nn.init.normal_(cls, std=.02); nn.init.normal_(mask_tok, std=.02)
        order = []                                   # nn.Module.apply order == registration order
        for i in range(cfg.depth):
            order += [f"blocks.{i}.attn.qkv", f"blocks.{i}.attn.proj", f"blocks.{i}.mlp.fc1", f"blocks.{i}.mlp.fc2"]
        order += ["decoder_embed"]
        for i in range(cfg.decoder_depth):
            order += [f"decoder_blocks.{i}.attn.qkv", f"decoder_blocks.{i}.attn.proj",
                      f"decoder_blocks.{i}.mlp.fc1", f"decoder_blocks.{i}.mlp.fc2"]
        order += ["decoder_pred"]
        for name in order:
            nn.init.xavier_uniform_(mods[name].weight)
            nn.init.zeros_(mods[name].bias)
        for name, m in mods.items():
            sd[name + ".weight"] = m.weight.data
            sd[name + ".bias"] = m.bias.data
        for pre, n, dim in ([("blocks", cfg.depth, D)] + [("decoder_blocks", cfg.decoder_depth, Dd)]):
            for i in range(n):
                for ln in ("norm1", "norm2"):
                    sd[f"{pre}.{i}.{ln}.weight"] = torch.ones(dim)
                    sd[f"{pre}.{i}.{ln}.bias"] = torch.zeros(dim)
        sd["norm.weight"], sd["norm.bias"] = torch.ones(D), torch.zeros(D)
        sd["decoder_norm.weight"], sd["decoder_norm.bias"] = torch.ones(Dd), torch.zeros(Dd)
        sd["cls_token"], sd["mask_token"] = cls, mask_tok
        sd["pos_embed"] = sincos_pos_embed(D, cfg.grid).unsqueeze(0)
        sd["decoder_pos_embed"] = sincos_pos_embed(Dd, cfg.grid).unsqueeze(0)
        if head_classes is not None:
            head = nn.Linear(D, head_classes)
            gone = ("mask_token", "decoder_embed.", "decoder_blocks.", "decoder_norm.", "decoder_pred.")
            sd = {k: v for k, v in sd.items() if not k.startswith(gone)}
            sd["lin_head.weight"], sd["lin_head.bias"] = head.weight.data, head.bias.data
    return sd


def vit_any_state_dict(seed: int, num_classes: int = 2, embed_dim: int = 768, depth: int = 12, num_heads: int = 12):
    """Weights of `VisionTransformer_from_Any(head=True, ...)` under `torch.manual_seed(seed)`, drawn through the timm
    shim (construction order and trunc_normal init of timm 0.4.12) WITHOUT the reference tree, so the GPU box can
    rebuild them; make_cls_golden.py checks this against the reference constructor (models.py:47-62) bit-for-bit."""
    import torch.nn as nn
    from oracle import timm_shim
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        vt = timm_shim.VisionTransformer(patch_size=16, embed_dim=embed_dim, depth=depth, num_heads=num_heads)
        lin_head = nn.Linear(embed_dim, num_classes)
    sd = {k: v.detach().clone() for k, v in vt.state_dict().items() if not k.startswith("head.")}
    sd["lin_head.weight"], sd["lin_head.bias"] = lin_head.weight.detach().clone(), lin_head.bias.detach().clone()
    return sd


# ----------------------------------------------------------------------------------------------
# a16 / a18 / a19 — one fp32 pretraining step (engine_pretrain.py:47-72 without the scaler)
# ----------------------------------------------------------------------------------------------
def weight_decay_groups(named_params, weight_decay=0.05):
    """timm 0.4.12 `optim_factory.add_weight_decay` as called at main_pretrain.py:217: tensors with
    ndim==1 or whose name ends in `.bias` get wd 0, everything else (cls/mask tokens too) gets wd."""
    decay, no_decay = [], []
    for name, p in named_params:
        if not p.requires_grad:
            continue
        (no_decay if (p.ndim == 1 or name.endswith(".bias")) else decay).append(p)
    return [{"params": no_decay, "weight_decay": 0.}, {"params": decay, "weight_decay": weight_decay}]


def cosine_lr(epoch_f, lr, min_lr, warmup_epochs, epochs):
    """util/lr_sched.py:9-21."""
    if epoch_f < warmup_epochs:
        return lr * epoch_f / warmup_epochs
    return min_lr + (lr - min_lr) * 0.5 * (1. + math.cos(math.pi * (epoch_f - warmup_epochs) / (epochs - warmup_epochs)))
