"""Fixed 2-D sin-cos position tables, computed on the host in fp64 exactly as the reference does
(/root/reference/src/ssl4polyp/models/mae/util/pos_embed.py:20-67), so that the fp32 table is bit-identical.

Layout facts that matter for parity: the meshgrid is built *w first*; the first half of the channels encodes the
w coordinate, the second half the h coordinate; inside each half the first quarter is sin, the second cos; the
cls row (if requested) is all zeros and comes first.
"""
from __future__ import annotations

import numpy as np
import torch


def get_1d_sincos_pos_embed_from_grid(embed_dim: int, pos) -> np.ndarray:
    """pos: any-shaped array of positions -> [pos.size, embed_dim] = [sin(pos*w_i) | cos(pos*w_i)], w_i = 10000^(-2i/dim)."""
    if embed_dim % 2 != 0:
        raise AssertionError("embed_dim must be even")
    freq = np.arange(embed_dim // 2, dtype=np.float64)
    freq /= embed_dim / 2.0
    freq = 1.0 / 10000 ** freq
    angle = np.outer(np.asarray(pos).reshape(-1), freq)
    return np.concatenate([np.sin(angle), np.cos(angle)], axis=1)


def get_2d_sincos_pos_embed_from_grid(embed_dim: int, grid) -> np.ndarray:
    if embed_dim % 2 != 0:
        raise AssertionError("embed_dim must be even")
    first = get_1d_sincos_pos_embed_from_grid(embed_dim // 2, grid[0])
    second = get_1d_sincos_pos_embed_from_grid(embed_dim // 2, grid[1])
    return np.concatenate([first, second], axis=1)


def get_2d_sincos_pos_embed(embed_dim: int, grid_size: int, cls_token: bool = False) -> np.ndarray:
    """-> [grid_size**2 (+1), embed_dim] float64 (callers cast with torch.from_numpy(...).float())."""
    coords_h = np.arange(grid_size, dtype=np.float32)
    coords_w = np.arange(grid_size, dtype=np.float32)
    grid = np.stack(np.meshgrid(coords_w, coords_h), axis=0).reshape([2, 1, grid_size, grid_size])
    table = get_2d_sincos_pos_embed_from_grid(embed_dim, grid)
    if cls_token:
        table = np.concatenate([np.zeros([1, embed_dim]), table], axis=0)
    return table


def interpolate_pos_embed(model, checkpoint_model) -> None:
    """Bicubic resize of a checkpoint's `pos_embed` to the model's grid (reference util/pos_embed.py:75-96)."""
    if "pos_embed" not in checkpoint_model:
        return
    ckpt = checkpoint_model["pos_embed"]
    dim = ckpt.shape[-1]
    n_patches = model.patch_embed.num_patches
    n_extra = model.pos_embed.shape[-2] - n_patches
    old = int((ckpt.shape[-2] - n_extra) ** 0.5)
    new = int(n_patches ** 0.5)
    if old == new:
        return
    print("Position interpolate from %dx%d to %dx%d" % (old, old, new, new))
    extra = ckpt[:, :n_extra]
    grid = ckpt[:, n_extra:].reshape(-1, old, old, dim).permute(0, 3, 1, 2)
    grid = torch.nn.functional.interpolate(grid, size=(new, new), mode="bicubic", align_corners=False)
    grid = grid.permute(0, 2, 3, 1).flatten(1, 2)
    checkpoint_model["pos_embed"] = torch.cat((extra, grid), dim=1)
