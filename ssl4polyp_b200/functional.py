"""Stand-alone differentiable pieces (a `Block` or `PatchEmbed` called on its own, as `blk(x)` /
`self.patch_embed(x)` in the reference: models_mae.py:152,166-167).  They run the same kernels and the same
hand-written backward as the whole-model path; bf16 operand copies of the weights are made per call, so the
whole-model entry points (`MaskedAutoencoderViT.forward`, `ViT_from_MAE.forward`) are the fast path.
"""
from __future__ import annotations

import torch

from . import ops, vit_core
from .ops import EPI_BIAS_F32, GEMM_NT


class _DictWeights(vit_core.Weights):
    def __init__(self, named):
        self.p = dict(named)
        self._b = {}

    def f(self, name):
        return self.p[name].data

    def b(self, name):
        t = self._b.get(name)
        if t is None:
            w = self.p[name].data
            t = ops.cast_bf16(w.contiguous()).view(w.shape[0], -1)
            self._b[name] = t
        return t

    def need(self, name):
        return name in self.p and self.p[name].requires_grad

    def has(self, name):
        return name in self.p

    def names(self):
        return self.p.keys()


def _cuda_f32(x):
    if not x.is_cuda:
        raise RuntimeError("ssl4polyp_b200: input must be a CUDA tensor (no CPU path)")
    return x.float().contiguous()


class _BlockFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, blk, x, save, names, *params):
        W = _DictWeights(zip(names, params))
        B, T, D = x.shape
        eps = float(blk.norm1.eps)
        y, sv = vit_core.block_fwd(W, "", x.view(B * T, D), B, T, blk.num_heads, eps, save)
        ctx.W, ctx.sv, ctx.names, ctx.shape, ctx.H = W, sv, names, (B, T, D), blk.num_heads
        return y.view(B, T, D)

    @staticmethod
    def backward(ctx, dy):
        B, T, D = ctx.shape
        sink = vit_core.FreshGrads()
        dx = dy.detach().float().contiguous().view(B * T, D).clone()
        dxb = ops.cast_bf16(dx)
        dx, _ = vit_core.block_bwd(ctx.W, sink, "", ctx.sv, dx, dxb, B, T, ctx.H, want_bf16=False)
        ctx.sv = None
        grads = tuple(sink.grads.get(n) if ctx.W.need(n) else None for n in ctx.names)
        return (None, dx.view(B, T, D), None, None) + grads


def block_apply(blk, x):
    x = _cuda_f32(x)
    named = list(blk.named_parameters())
    names = tuple(n for n, _ in named)
    params = [p for _, p in named]
    save = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params))
    return _BlockFn.apply(blk, x, save, names, *params)


class _PatchEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, imgs, weight, bias, p):
        N = imgs.shape[0]
        cols = ops.im2col_patches(imgs, None, p)
        wb = ops.cast_bf16(weight.detach().contiguous()).view(weight.shape[0], -1)
        tok = ops.gemm(cols, wb, GEMM_NT, EPI_BIAS_F32, bias=bias.detach())
        ctx.cols = cols
        ctx.wshape = weight.shape
        return tok.view(N, -1, weight.shape[0])

    @staticmethod
    def backward(ctx, dtok):
        D = ctx.wshape[0]
        d = ops.cast_bf16(dtok.detach().float().contiguous().view(-1, D))
        W = _DictWeights([])
        dW = ops.gemm(d, ctx.cols, ops.GEMM_TN, ops.EPI_PARTIAL_F32, splits=ops.pick_splits(D, ctx.cols.shape[1], d.shape[0]))
        db = ops.colsum(d)
        return None, dW.view(ctx.wshape), db, None


def patch_embed_all(imgs, weight, bias, p):
    """timm PatchEmbed.forward: [N,C,H,W] -> [N, L, D] (no gradient w.r.t. the image)."""
    return _PatchEmbedFn.apply(_cuda_f32(imgs), weight, bias, p)
