"""TEST INFRASTRUCTURE ONLY — restated `timm==0.4.12` layer library for the oracle.

The reference's hot-path arithmetic lives in an un-vendored third-party package:
`timm==0.4.12` (pinned at /root/reference `pyproject.toml:25`, `requirements.txt:5`,
asserted at `src/ssl4polyp/models/mae/main_pretrain.py:37`).  It is imported at
`src/ssl4polyp/models/mae/models_mae.py:17` (`PatchEmbed`, `Block`) and
`src/ssl4polyp/models/models.py:14` (`VisionTransformer`).  timm is neither installed
nor installable in this image, so this file restates the *published* timm 0.4.12
modules the reference instantiates (`timm/models/vision_transformer.py` and
`timm/models/layers/{patch_embed,mlp}.py` at tag v0.4.12), keeping the sub-module
names that fix the state-dict keys (`proj`, `norm1`, `attn.qkv`, `attn.proj`,
`norm2`, `mlp.fc1`, `mlp.fc2`).

`install()` registers the shim under `sys.modules["timm…"]` so the reference's
`models_mae.py` can be imported UNMODIFIED by `oracle/make_golden.py`.

Nothing under `ssl4polyp_b200/` may import this file (see oracle/README.md).
"""
from __future__ import annotations

import sys
import types
from functools import partial

import torch
import torch.nn as nn


def _pair(v):
    return v if isinstance(v, tuple) else (v, v)


class PatchEmbed(nn.Module):
    """timm 0.4.12 `layers/patch_embed.py::PatchEmbed` — Conv2d(k=s=patch) then flatten/transpose."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768, norm_layer=None, flatten=True):
        super().__init__()
        img_size = _pair(img_size)
        patch_size = _pair(patch_size)
        self.img_size = img_size
        self.patch_size = patch_size
        self.grid_size = (img_size[0] // patch_size[0], img_size[1] // patch_size[1])
        self.num_patches = self.grid_size[0] * self.grid_size[1]
        self.flatten = flatten
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_size)
        self.norm = norm_layer(embed_dim) if norm_layer else nn.Identity()

    def forward(self, x):
        B, C, H, W = x.shape
        assert H == self.img_size[0] and W == self.img_size[1], \
            f"Input image size ({H}*{W}) doesn't match model ({self.img_size[0]}*{self.img_size[1]})."
        x = self.proj(x)
        if self.flatten:
            x = x.flatten(2).transpose(1, 2)  # BCHW -> BNC
        x = self.norm(x)
        return x


class Mlp(nn.Module):
    """timm 0.4.12 `layers/mlp.py::Mlp` — fc1 → GELU(erf) → drop → fc2 → drop."""

    def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.):
        super().__init__()
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_features, out_features)
        self.drop = nn.Dropout(drop)

    def forward(self, x):
        x = self.fc1(x)
        x = self.act(x)
        x = self.drop(x)
        x = self.fc2(x)
        x = self.drop(x)
        return x


class Attention(nn.Module):
    """timm 0.4.12 `vision_transformer.py::Attention`."""

    def __init__(self, dim, num_heads=8, qkv_bias=False, attn_drop=0., proj_drop=0.):
        super().__init__()
        self.num_heads = num_heads
        head_dim = dim // num_heads
        self.scale = head_dim ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=qkv_bias)
        self.attn_drop = nn.Dropout(attn_drop)
        self.proj = nn.Linear(dim, dim)
        self.proj_drop = nn.Dropout(proj_drop)

    def forward(self, x):
        B, N, C = x.shape
        qkv = self.qkv(x).reshape(B, N, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        attn = (q @ k.transpose(-2, -1)) * self.scale
        attn = attn.softmax(dim=-1)
        attn = self.attn_drop(attn)
        x = (attn @ v).transpose(1, 2).reshape(B, N, C)
        x = self.proj(x)
        x = self.proj_drop(x)
        return x


class Block(nn.Module):
    """timm 0.4.12 `vision_transformer.py::Block` — pre-LN; DropPath is Identity at rate 0."""

    def __init__(self, dim, num_heads, mlp_ratio=4., qkv_bias=False, drop=0., attn_drop=0.,
                 drop_path=0., act_layer=nn.GELU, norm_layer=nn.LayerNorm):
        super().__init__()
        assert drop_path == 0., "oracle shim only restates the drop_path=0 configuration the reference uses"
        self.norm1 = norm_layer(dim)
        self.attn = Attention(dim, num_heads=num_heads, qkv_bias=qkv_bias, attn_drop=attn_drop, proj_drop=drop)
        self.drop_path = nn.Identity()
        self.norm2 = norm_layer(dim)
        self.mlp = Mlp(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer, drop=drop)

    def forward(self, x):
        x = x + self.drop_path(self.attn(self.norm1(x)))
        x = x + self.drop_path(self.mlp(self.norm2(x)))
        return x


def _trunc_normal_(t, std=.02):
    return nn.init.trunc_normal_(t, mean=0., std=std, a=-2., b=2.)


class VisionTransformer(nn.Module):
    """The part of timm 0.4.12 `VisionTransformer` that `models.py:26-140` relies on.

    Defaults as in 0.4.12: img 224, in_chans 3, num_classes 1000, mlp_ratio 4, qkv_bias True,
    norm = LayerNorm(eps=1e-6), act = GELU, learned `pos_embed` trunc_normal(.02), `cls_token`
    trunc_normal(.02), `blocks` held in `nn.Sequential`, Linear init trunc_normal(.02)/zero bias,
    LayerNorm (1, 0).  `dist_token`, `pre_logits` and the `head` are the no-distillation,
    no-representation defaults.
    """

    def __init__(self, img_size=224, patch_size=16, in_chans=3, num_classes=1000, embed_dim=768, depth=12,
                 num_heads=12, mlp_ratio=4., qkv_bias=True, norm_layer=None, act_layer=None):
        super().__init__()
        self.num_classes = num_classes
        self.num_features = self.embed_dim = embed_dim
        self.num_tokens = 1
        norm_layer = norm_layer or partial(nn.LayerNorm, eps=1e-6)
        act_layer = act_layer or nn.GELU
        self.patch_embed = PatchEmbed(img_size=img_size, patch_size=patch_size, in_chans=in_chans, embed_dim=embed_dim)
        num_patches = self.patch_embed.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.dist_token = None
        self.pos_embed = nn.Parameter(torch.zeros(1, num_patches + self.num_tokens, embed_dim))
        self.pos_drop = nn.Dropout(p=0.)
        self.blocks = nn.Sequential(*[
            Block(dim=embed_dim, num_heads=num_heads, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias,
                  norm_layer=norm_layer, act_layer=act_layer)
            for _ in range(depth)])
        self.norm = norm_layer(embed_dim)
        self.pre_logits = nn.Identity()
        self.head = nn.Linear(self.num_features, num_classes) if num_classes > 0 else nn.Identity()
        _trunc_normal_(self.pos_embed, std=.02)
        _trunc_normal_(self.cls_token, std=.02)
        self.apply(self._init_vit_weights)

    @staticmethod
    def _init_vit_weights(m):
        if isinstance(m, nn.Linear):
            _trunc_normal_(m.weight, std=.02)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.LayerNorm):
            nn.init.zeros_(m.bias)
            nn.init.ones_(m.weight)


def install():
    """Register the shim as `timm`, `timm.models`, `timm.models.vision_transformer` (+ layers)."""
    import numpy as np
    if not hasattr(np, "float"):          # util/pos_embed.py:56 uses the removed alias np.float
        np.float = float                   # type: ignore[attr-defined]
    if "timm" in sys.modules and not getattr(sys.modules["timm"], "_ssl4p_oracle_shim", False):
        return sys.modules["timm"]         # a real timm is present: use it
    timm = types.ModuleType("timm")
    timm.__version__ = "0.4.12"
    timm._ssl4p_oracle_shim = True
    models = types.ModuleType("timm.models")
    vt = types.ModuleType("timm.models.vision_transformer")
    layers = types.ModuleType("timm.models.layers")
    for m in (vt, layers):
        m.PatchEmbed, m.Mlp = PatchEmbed, Mlp
    vt.Attention, vt.Block, vt.VisionTransformer = Attention, Block, VisionTransformer
    timm.models, models.vision_transformer, models.layers = models, vt, layers
    sys.modules.update({"timm": timm, "timm.models": models,
                        "timm.models.vision_transformer": vt, "timm.models.layers": layers})
    return timm
