"""Parameter containers with the module tree (and therefore the state-dict keys) of timm 0.4.12's
`PatchEmbed` / `Block{Attention, Mlp}` as the reference instantiates them
(/root/reference/src/ssl4polyp/models/mae/models_mae.py:17,33,39-41,53-55):

    patch_embed.proj.{weight,bias}
    blocks.i.norm1.{weight,bias}  blocks.i.attn.qkv.{weight,bias}  blocks.i.attn.proj.{weight,bias}
    blocks.i.norm2.{weight,bias}  blocks.i.mlp.fc1.{weight,bias}   blocks.i.mlp.fc2.{weight,bias}

They hold parameters and hyper-parameters only.  The arithmetic of the hot path is scheduled over whole models by
`vit_core` on our sm_100a kernels; a `Block` called on its own runs the same kernels through `vit_core.block_fwd/bwd`.
"""
from __future__ import annotations

import torch
import torch.nn as nn


def _pair(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


class PatchEmbed(nn.Module):
    """Image -> patch tokens: Conv2d(in_chans, embed_dim, kernel=stride=patch) viewed as a GEMM over patch vectors."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768, norm_layer=None, flatten=True):
        super().__init__()
        self.img_size = _pair(img_size)
        self.patch_size = _pair(patch_size)
        self.grid_size = (self.img_size[0] // self.patch_size[0], self.img_size[1] // self.patch_size[1])
        self.num_patches = self.grid_size[0] * self.grid_size[1]
        self.flatten = flatten
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=self.patch_size, stride=self.patch_size)
        self.norm = norm_layer(embed_dim) if norm_layer else nn.Identity()

    def check_input(self, x):
        H, W = x.shape[-2:]
        assert H == self.img_size[0] and W == self.img_size[1], \
            f"Input image size ({H}*{W}) doesn't match model ({self.img_size[0]}*{self.img_size[1]})."

    def forward(self, x):
        """[N,C,H,W] fp32 -> [N, num_patches, embed_dim] fp32 (im2col kernel + tcgen05 GEMM, differentiable)."""
        from .functional import patch_embed_all
        self.check_input(x)
        return patch_embed_all(x, self.proj.weight, self.proj.bias, self.patch_size[0])


class Mlp(nn.Module):
    def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.):
        super().__init__()
        if drop:
            raise NotImplementedError("dropout > 0 is not on the MAE hot path (the reference always uses 0)")
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_features, out_features)


class Attention(nn.Module):
    def __init__(self, dim, num_heads=8, qkv_bias=False, attn_drop=0., proj_drop=0.):
        super().__init__()
        if attn_drop or proj_drop:
            raise NotImplementedError("dropout > 0 is not on the MAE hot path (the reference always uses 0)")
        if not qkv_bias:
            raise NotImplementedError("qkv_bias=False is not on the MAE hot path (the reference always passes True)")
        self.num_heads = num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.proj = nn.Linear(dim, dim)


class Block(nn.Module):
    """Pre-LN transformer block: x + attn(norm1(x)); x + mlp(norm2(x))."""

    def __init__(self, dim, num_heads, mlp_ratio=4., qkv_bias=False, drop=0., attn_drop=0., drop_path=0.,
                 act_layer=nn.GELU, norm_layer=nn.LayerNorm):
        super().__init__()
        if drop or attn_drop or drop_path:
            raise NotImplementedError("dropout / drop-path > 0 is not on the MAE hot path")
        if act_layer is not nn.GELU:
            raise NotImplementedError("only exact-erf nn.GELU is implemented (what the reference uses)")
        self.dim = dim
        self.num_heads = num_heads
        self.norm1 = norm_layer(dim)
        self.attn = Attention(dim, num_heads=num_heads, qkv_bias=qkv_bias)
        self.norm2 = norm_layer(dim)
        self.mlp = Mlp(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer)

    def forward(self, x):
        """[B,T,D] fp32 -> [B,T,D] fp32 on our kernels (differentiable)."""
        from .functional import block_apply
        return block_apply(self, x)
