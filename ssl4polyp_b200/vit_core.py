"""Functional forward/backward of the MAE ViT hot path over the sm_100a kernels (include/ssl4p.h).

This file is the host-side schedule only: which kernel runs on which buffer, in which order.  Every tensor op
below is one of our CUDA kernels reached through the C-ABI (ssl4polyp_b200/ops.py); torch allocates the buffers.
The backward is written by hand (no autograd inside), so the same code serves
  * the autograd.Function wrappers of `models_mae.MaskedAutoencoderViT` / `models.ViT_from_MAE` (drop-in API), and
  * the fused data-parallel step engine (`engine.py`: flat gradient arena, bucketed NCCL all-reduce overlapped
    with this backward, fused AdamW, CUDA-graph capture).

Reference semantics restated here (file:line under /root/reference/src/ssl4polyp/models):
  mae/models_mae.py:150-170 forward_encoder, :172-196 forward_decoder, :198-214 forward_loss, :123-148
  random_masking; timm 0.4.12 Block/Attention/Mlp (pre-LN, exact-erf GELU, scale = head_dim**-0.5);
  models.py:196-222 ViT_from_MAE.forward_encoder/forward.

Numerics: residual stream fp32; LayerNorm statistics, softmax, loss in fp32; GEMM/attention operands bf16 with
fp32 accumulation; weight gradients fp32.
"""
from __future__ import annotations

import contextlib
from dataclasses import dataclass, field

import torch

from . import ops
from .ops import (EPI_BIAS_BF16, EPI_BIAS_F32, EPI_BIAS_GELU_BF16, EPI_BIAS_RESIDUAL_F32, EPI_DGELU_BF16,
                  EPI_PARTIAL_F32, GEMM_NN, GEMM_NT, GEMM_TN)


@dataclass
class ViTDims:
    img_size: int
    patch_size: int
    in_chans: int
    embed_dim: int
    depth: int
    num_heads: int
    decoder_embed_dim: int = 0
    decoder_depth: int = 0
    decoder_num_heads: int = 0
    ln_eps: float = 1e-6
    block_prefix: str = "blocks."

    @property
    def grid(self) -> int:
        return self.img_size // self.patch_size

    @property
    def num_patches(self) -> int:
        return self.grid * self.grid

    @property
    def patch_dim(self) -> int:
        return self.patch_size * self.patch_size * self.in_chans


class Weights:
    """What the schedule needs from the parameter store: fp32 tensors by state-dict name, bf16 shadows of the
    GEMM weights (2-D, [out, in]), and which names want a gradient."""

    def f(self, name: str) -> torch.Tensor: raise NotImplementedError
    def b(self, name: str) -> torch.Tensor: raise NotImplementedError
    def need(self, name: str) -> bool: raise NotImplementedError
    def has(self, name: str) -> bool: raise NotImplementedError
    def names(self): raise NotImplementedError


class GradSink:
    """Where weight gradients go.  `buf(name, shape)` -> (fp32 tensor to write, accumulate?);
    `ready(name)` is called once the kernels producing that gradient have been enqueued.

    Weight gradients are OFF the critical path of the backward (the dgrad chain never reads them), so a sink may run them
    on a second stream: `side(*operands)` is the context in which `_wgrad` enqueues its kernels (default: the current
    stream), `block_boundary()` is called between transformer blocks (a sink with a side stream joins lagging side work
    there and releases the operands it kept alive for it)."""

    def buf(self, name: str, like: torch.Tensor): raise NotImplementedError
    def ready(self, name: str) -> None: pass
    def side(self, *operands): return contextlib.nullcontext()
    def block_boundary(self) -> None: pass


class FreshGrads(GradSink):
    """Allocates a new fp32 tensor per gradient (the autograd.Function path hands these to autograd)."""

    def __init__(self):
        self.grads: dict[str, torch.Tensor] = {}

    def buf(self, name, like):
        t = torch.empty(like.shape, dtype=torch.float32, device=like.device)
        self.grads[name] = t
        return t, False


_BLOCK_PARAMS = ("norm1.weight", "norm1.bias", "attn.qkv.weight", "attn.qkv.bias", "attn.proj.weight", "attn.proj.bias",
                 "norm2.weight", "norm2.bias", "mlp.fc1.weight", "mlp.fc1.bias", "mlp.fc2.weight", "mlp.fc2.bias")


def _block_needs(W: Weights, pre: str) -> bool:
    return any(W.need(pre + s) for s in _BLOCK_PARAMS)


# ----------------------------------------------------------------------------------------------------------
# linear layer backward helpers
# ----------------------------------------------------------------------------------------------------------
def _wgrad(W: Weights, sink: GradSink, wname: str, bname: str | None, dy: torch.Tensor, x: torch.Tensor):
    """dW[out,in] = dy^T x (TN GEMM, deterministic split-K), db = column sums of dy (bname None: the bias gradient
    was already produced by the LayerNorm backward that wrote dy, see _ln_bwd)."""
    if not (W.need(wname) or (bname is not None and W.need(bname))):
        return
    with sink.side(dy, x):
        if W.need(wname):
            w = W.f(wname)
            buf, acc = sink.buf(wname, w)
            buf2 = buf.view(w.shape[0], -1)
            M, N, K = dy.shape[1], x.shape[1], dy.shape[0]
            s = ops.pick_splits(M, N, K)
            if acc:
                ops.gemm(dy, x, GEMM_TN, EPI_PARTIAL_F32, accumulate_into=buf2, splits=s)
            else:
                ops.gemm(dy, x, GEMM_TN, EPI_PARTIAL_F32, out=buf2, splits=s)
            sink.ready(wname)
        if bname is not None and W.need(bname):
            buf, acc = sink.buf(bname, W.f(bname))
            ops.colsum(dy, out=buf, accumulate=acc)
            sink.ready(bname)


def _ln_bwd(W: Weights, sink: GradSink, pre: str, dy, x, mean, rstd, residual_grad, want_bf16=True, upstream_bias=None):
    """LayerNorm backward; gamma/beta gradients go to the sink (or to a scratch pair when frozen).
    upstream_bias: name of the Linear bias whose output (+ skip) this norm read; its gradient is the column sum of
    the dx produced here and comes out of the same kernel (no separate pass over dx)."""
    gname, bname = pre + "weight", pre + "bias"
    gamma = W.f(gname)
    if W.need(gname) or W.need(bname):
        dg, acc_g = sink.buf(gname, gamma)
        db, acc_b = sink.buf(bname, gamma)
        assert acc_g == acc_b
        acc = acc_g
    else:
        dg = torch.empty_like(gamma, dtype=torch.float32)
        db = torch.empty_like(gamma, dtype=torch.float32)
        acc = False
    cs, cs_acc = None, False
    if upstream_bias is not None and W.need(upstream_bias):
        cs, cs_acc = sink.buf(upstream_bias, W.f(upstream_bias))
    out = ops.layernorm_bwd(dy, x, mean, rstd, gamma, residual_grad, dg, db, want_bf16=want_bf16, accumulate=acc,
                            dx_colsum=cs, accumulate_colsum=cs_acc)
    sink.ready(gname)
    sink.ready(bname)
    if cs is not None:
        sink.ready(upstream_bias)
    return out


# ----------------------------------------------------------------------------------------------------------
# transformer block (timm 0.4.12 Block): x + proj(attn(norm1(x)));  x + fc2(gelu(fc1(norm2(x))))
# ----------------------------------------------------------------------------------------------------------
def block_fwd(W: Weights, pre: str, x: torch.Tensor, B: int, T: int, H: int, eps: float, save: bool):
    D = x.shape[1]
    hd = D // H
    h, mean1, rstd1 = ops.layernorm_fwd(x, W.f(pre + "norm1.weight"), W.f(pre + "norm1.bias"), eps)
    qkv = ops.gemm(h, W.b(pre + "attn.qkv.weight"), GEMM_NT, EPI_BIAS_BF16, bias=W.f(pre + "attn.qkv.bias"))
    o, lse = ops.attention_fwd(qkv, B, T, H, hd, hd ** -0.5)
    x1 = ops.gemm(o, W.b(pre + "attn.proj.weight"), GEMM_NT, EPI_BIAS_RESIDUAL_F32, bias=W.f(pre + "attn.proj.bias"), aux=x)
    h2, mean2, rstd2 = ops.layernorm_fwd(x1, W.f(pre + "norm2.weight"), W.f(pre + "norm2.bias"), eps)
    dact, act = ops.gemm(h2, W.b(pre + "mlp.fc1.weight"), GEMM_NT, EPI_BIAS_GELU_BF16, bias=W.f(pre + "mlp.fc1.bias"))
    x2 = ops.gemm(act, W.b(pre + "mlp.fc2.weight"), GEMM_NT, EPI_BIAS_RESIDUAL_F32, bias=W.f(pre + "mlp.fc2.bias"), aux=x1)
    saved = (x, mean1, rstd1, h, qkv, o, lse, x1, mean2, rstd2, h2, dact, act) if save else None
    return x2, saved


def block_bwd(W: Weights, sink: GradSink, pre: str, saved, dx: torch.Tensor, dxb: torch.Tensor, B: int, T: int, H: int,
              want_bf16: bool = True, prev_fc2_bias: str | None = None, fc2_bias_done: bool = False):
    """dx (fp32, updated in place) / dxb (its bf16 copy) = gradient w.r.t. the block output.  fc2_bias_done: the
    LayerNorm backward that produced them has already written this block's mlp.fc2.bias gradient.
    prev_fc2_bias: bias of the previous block's fc2 (its gradient = column sums of the dx returned here).
    Returns (dx, dxb) w.r.t. the block input."""
    x, mean1, rstd1, h, qkv, o, lse, x1, mean2, rstd2, h2, dact, act = saved
    D = x.shape[1]
    hd = D // H
    sink.block_boundary()
    # ---- MLP branch
    _wgrad(W, sink, pre + "mlp.fc2.weight", None if fc2_bias_done else pre + "mlp.fc2.bias", dxb, act)
    b1 = pre + "mlp.fc1.bias"
    if W.need(b1):                               # fc1 bias gradient = column sums of dpre, out of the same GEMM epilogue
        cs, cs_acc = sink.buf(b1, W.f(b1))
        dpre = ops.gemm(dxb, W.b(pre + "mlp.fc2.weight"), GEMM_NN, EPI_DGELU_BF16, aux=dact, colsum_out=cs, colsum_accumulate=cs_acc)
        sink.ready(b1)
    else:
        dpre = ops.gemm(dxb, W.b(pre + "mlp.fc2.weight"), GEMM_NN, EPI_DGELU_BF16, aux=dact)
    _wgrad(W, sink, pre + "mlp.fc1.weight", None, dpre, h2)
    dh2 = ops.gemm(dpre, W.b(pre + "mlp.fc1.weight"), GEMM_NN, EPI_BIAS_BF16)
    dx, dxb = _ln_bwd(W, sink, pre + "norm2.", dh2, x1, mean2, rstd2, dx, upstream_bias=pre + "attn.proj.bias")
    # ---- attention branch
    _wgrad(W, sink, pre + "attn.proj.weight", None, dxb, o)
    do = ops.gemm(dxb, W.b(pre + "attn.proj.weight"), GEMM_NN, EPI_BIAS_BF16)
    dqkv = ops.attention_bwd(qkv, o, do, lse, B, T, H, hd, hd ** -0.5)
    _wgrad(W, sink, pre + "attn.qkv.weight", pre + "attn.qkv.bias", dqkv, h)
    dh = ops.gemm(dqkv, W.b(pre + "attn.qkv.weight"), GEMM_NN, EPI_BIAS_BF16)
    dx, dxb = _ln_bwd(W, sink, pre + "norm1.", dh, x, mean1, rstd1, dx, want_bf16=want_bf16, upstream_bias=prev_fc2_bias)
    return dx, dxb


# ----------------------------------------------------------------------------------------------------------
# encoder: patch-embed (kept patches only) -> +pos, cls -> blocks -> norm
# ----------------------------------------------------------------------------------------------------------
@dataclass
class EncoderSaved:
    cols: torch.Tensor = None            # bf16 [N*keep, C*p*p]  (patch vectors of the kept patches)
    blocks: list = field(default_factory=list)
    x_final: torch.Tensor = None         # fp32 [N*Te, D] input of the final LayerNorm
    mean: torch.Tensor = None
    rstd: torch.Tensor = None
    N: int = 0
    keep: int = 0
    all_patches: bool = False            # no masking: token r of every image is patch r


def encoder_fwd(W: Weights, dims: ViTDims, imgs: torch.Tensor, ids_keep, keep: int, save: bool):
    """ids_keep None -> all patches in order (classifier path, models.py:196-209).
    Returns (latent bf16 [N*Te, D], EncoderSaved | None)."""
    N = imgs.shape[0]
    D, p = dims.embed_dim, dims.patch_size
    Te = keep + 1
    cols = ops.im2col_patches(imgs, ids_keep, p)
    w_pe = W.b("patch_embed.proj.weight")
    tok = ops.gemm(cols, w_pe, GEMM_NT, EPI_BIAS_F32, bias=W.f("patch_embed.proj.bias"))
    x = ops.assemble_tokens(tok, ids_keep, W.f("pos_embed"), W.f("cls_token"), N, keep, D).view(N * Te, D)
    sv = EncoderSaved(cols=cols, N=N, keep=keep, all_patches=ids_keep is None) if save else None
    for i in range(dims.depth):
        x, s = block_fwd(W, f"{dims.block_prefix}{i}.", x, N, Te, dims.num_heads, dims.ln_eps, save)
        if save:
            sv.blocks.append(s)
    latent, mean, rstd = ops.layernorm_fwd(x, W.f("norm.weight"), W.f("norm.bias"), dims.ln_eps)
    if save:
        sv.x_final, sv.mean, sv.rstd = x, mean, rstd
    return latent, sv


def _first_needed_block(W: Weights, dims: ViTDims, prefix: str, depth: int, upstream_needed: bool) -> int:
    """Index of the first block the backward has to reach (depth = none)."""
    if upstream_needed:
        return 0
    for i in range(depth):
        if _block_needs(W, f"{prefix}{i}."):
            return i
    return depth


def encoder_bwd(W: Weights, sink: GradSink, dims: ViTDims, sv: EncoderSaved, dlatent: torch.Tensor):
    """dlatent: bf16 [N*Te, D], gradient w.r.t. the output of the final LayerNorm."""
    N, keep = sv.N, sv.keep
    Te = keep + 1
    stem_needed = (W.need("patch_embed.proj.weight") or W.need("patch_embed.proj.bias") or W.need("cls_token")
                   or W.need("pos_embed"))
    first = _first_needed_block(W, dims, dims.block_prefix, dims.depth, stem_needed)
    if first == dims.depth and not (W.need("norm.weight") or W.need("norm.bias")) and not stem_needed:
        return
    dx, dxb = _ln_bwd(W, sink, "norm.", dlatent, sv.x_final, sv.mean, sv.rstd, None,
                      upstream_bias=f"{dims.block_prefix}{dims.depth - 1}.mlp.fc2.bias" if dims.depth > 0 else None)
    for i in range(dims.depth - 1, first - 1, -1):
        dx, dxb = block_bwd(W, sink, f"{dims.block_prefix}{i}.", sv.blocks[i], dx, dxb, N, Te, dims.num_heads,
                            want_bf16=(i > first), prev_fc2_bias=f"{dims.block_prefix}{i - 1}.mlp.fc2.bias" if i > 0 else None,
                            fc2_bias_done=True)
        sv.blocks[i] = None                      # free the block's activations as soon as they are consumed
    if not stem_needed:
        return
    cls = W.f("cls_token")
    if W.need("cls_token"):
        dcls, acc = sink.buf("cls_token", cls)
    else:
        dcls, acc = torch.empty(cls.shape, dtype=torch.float32, device=cls.device), False
    dtok = ops.assemble_tokens_bwd(dx.view(N, Te, -1), dcls, accumulate=acc)
    sink.ready("cls_token")
    if W.need("pos_embed"):
        # learned position table (timm VisionTransformer, models.py:28-33 `x + self.pos_embed`): d pos = sum over images.
        # The MAE encoder's table is a frozen sin-cos buffer (models_mae.py:36) and never gets here.
        if not sv.all_patches:
            raise RuntimeError("a trainable pos_embed is only supported without masking (classifier path)")
        dpos, acc_p = sink.buf("pos_embed", W.f("pos_embed"))
        ops.reduce_rows_f32(dx.view(N, -1), dpos.view(-1), accumulate=acc_p)
        sink.ready("pos_embed")
    _wgrad(W, sink, "patch_embed.proj.weight", "patch_embed.proj.bias", dtok, sv.cols)


# ----------------------------------------------------------------------------------------------------------
# decoder: embed -> mask-token fill + unshuffle + pos -> blocks -> norm -> pred
# ----------------------------------------------------------------------------------------------------------
@dataclass
class DecoderSaved:
    latent: torch.Tensor = None          # bf16 [N*Te, D]
    blocks: list = field(default_factory=list)
    x_final: torch.Tensor = None
    mean: torch.Tensor = None
    rstd: torch.Tensor = None
    hn: torch.Tensor = None              # bf16 [N*T, Dd] output of decoder_norm
    ids_keep: torch.Tensor = None
    mask: torch.Tensor = None
    N: int = 0
    Te: int = 0


def decoder_fwd(W: Weights, dims: ViTDims, latent: torch.Tensor, ids_restore, ids_keep, mask, N: int, save: bool):
    """latent bf16 [N*Te, D] -> pred_full fp32 [N, 1+L, P] (row 0 of each image is the cls row the reference
    drops at models_mae.py:194; callers slice it off as a view)."""
    Dd, L = dims.decoder_embed_dim, dims.num_patches
    Te = latent.shape[0] // N
    T = L + 1
    y = ops.gemm(latent, W.b("decoder_embed.weight"), GEMM_NT, EPI_BIAS_BF16, bias=W.f("decoder_embed.bias"))
    x = ops.decoder_unshuffle_fwd(y.view(N, Te, Dd), W.f("mask_token"), ids_restore, W.f("decoder_pos_embed")).view(N * T, Dd)
    sv = DecoderSaved(latent=latent, ids_keep=ids_keep, mask=mask, N=N, Te=Te) if save else None
    for i in range(dims.decoder_depth):
        x, s = block_fwd(W, f"decoder_blocks.{i}.", x, N, T, dims.decoder_num_heads, dims.ln_eps, save)
        if save:
            sv.blocks.append(s)
    hn, mean, rstd = ops.layernorm_fwd(x, W.f("decoder_norm.weight"), W.f("decoder_norm.bias"), dims.ln_eps)
    pred_full = ops.gemm(hn, W.b("decoder_pred.weight"), GEMM_NT, EPI_BIAS_F32, bias=W.f("decoder_pred.bias"))
    if save:
        sv.x_final, sv.mean, sv.rstd, sv.hn = x, mean, rstd, hn
    return pred_full.view(N, T, dims.patch_dim), sv


def decoder_bwd(W: Weights, sink: GradSink, dims: ViTDims, sv: DecoderSaved, dpred_full: torch.Tensor, want_dlatent: bool):
    """dpred_full: bf16 [N, 1+L, P] (cls rows zero).  Returns dlatent bf16 [N*Te, D] or None."""
    N, Te = sv.N, sv.Te
    L, Dd = dims.num_patches, dims.decoder_embed_dim
    T = L + 1
    dp = dpred_full.view(N * T, -1)
    _wgrad(W, sink, "decoder_pred.weight", "decoder_pred.bias", dp, sv.hn)
    dhn = ops.gemm(dp, W.b("decoder_pred.weight"), GEMM_NN, EPI_BIAS_BF16)
    dx, dxb = _ln_bwd(W, sink, "decoder_norm.", dhn, sv.x_final, sv.mean, sv.rstd, None,
                      upstream_bias=f"decoder_blocks.{dims.decoder_depth - 1}.mlp.fc2.bias" if dims.decoder_depth > 0 else None)
    for i in range(dims.decoder_depth - 1, -1, -1):
        dx, dxb = block_bwd(W, sink, f"decoder_blocks.{i}.", sv.blocks[i], dx, dxb, N, T, dims.decoder_num_heads,
                            want_bf16=(i > 0), prev_fc2_bias=f"decoder_blocks.{i - 1}.mlp.fc2.bias" if i > 0 else None,
                            fc2_bias_done=True)
        sv.blocks[i] = None
    mt = W.f("mask_token")
    if W.need("mask_token"):
        dmt, acc = sink.buf("mask_token", mt)
    else:
        dmt, acc = torch.empty(mt.shape, dtype=torch.float32, device=mt.device), False
    dy = ops.decoder_unshuffle_bwd(dx.view(N, T, Dd), sv.ids_keep, sv.mask, dmt, accumulate=acc)
    sink.ready("mask_token")
    dy2 = dy.view(N * Te, Dd)
    _wgrad(W, sink, "decoder_embed.weight", "decoder_embed.bias", dy2, sv.latent)
    if not want_dlatent:
        return None
    return ops.gemm(dy2, W.b("decoder_embed.weight"), GEMM_NN, EPI_BIAS_BF16)


# ----------------------------------------------------------------------------------------------------------
# the whole MAE step: forward -> (loss, pred_full, mask, ids_restore), backward from d(loss)
# ----------------------------------------------------------------------------------------------------------
@dataclass
class MAESaved:
    enc: EncoderSaved = None
    dec: DecoderSaved = None
    imgs: torch.Tensor = None
    pred_full: torch.Tensor = None
    mask: torch.Tensor = None
    loss_ws: torch.Tensor = None
    norm_pix: bool = False


def mae_fwd(W: Weights, dims: ViTDims, imgs: torch.Tensor, noise: torch.Tensor, len_keep: int, norm_pix: bool, save: bool):
    N = imgs.shape[0]
    ids_restore, mask, ids_keep = ops.mask_from_noise(noise, len_keep)
    latent, enc = encoder_fwd(W, dims, imgs, ids_keep, len_keep, save)
    pred_full, dec = decoder_fwd(W, dims, latent, ids_restore, ids_keep, mask, N, save)
    loss, ws = ops.masked_mse_fwd(imgs, pred_full, True, mask, norm_pix, dims.patch_size)
    sv = MAESaved(enc, dec, imgs, pred_full, mask, ws, norm_pix) if save else None
    return loss, pred_full, mask, ids_restore, sv


def mae_bwd(W: Weights, sink: GradSink, dims: ViTDims, sv: MAESaved, dloss: torch.Tensor):
    """dloss: fp32 0-dim device tensor (d objective / d loss)."""
    dpred = ops.masked_mse_bwd(sv.imgs, sv.pred_full, True, sv.mask, sv.norm_pix, dims.patch_size, dloss, sv.loss_ws)
    dlatent = decoder_bwd(W, sink, dims, sv.dec, dpred, want_dlatent=True)
    encoder_bwd(W, sink, dims, sv.enc, dlatent)


# ----------------------------------------------------------------------------------------------------------
# classifier: encoder on all patches -> pooled token -> lin_head   (models.py:211-222)
# ----------------------------------------------------------------------------------------------------------
@dataclass
class ClsSaved:
    enc: EncoderSaved = None
    xn: torch.Tensor = None
    spatial: bool = False
    head: bool = True


def classifier_fwd(W: Weights, dims: ViTDims, imgs: torch.Tensor, spatial: bool, save: bool):
    """-> (logits fp32 [N, ncls], ClsSaved | None); needs lin_head (head=True)."""
    N = imgs.shape[0]
    L = dims.num_patches
    xn, enc = encoder_fwd(W, dims, imgs, None, L, save)
    logits = ops.pool_head_fwd(xn, W.f("lin_head.weight"), W.f("lin_head.bias"), N, L + 1, dims.embed_dim, spatial)
    return logits, (ClsSaved(enc, xn, spatial) if save else None)


def classifier_bwd(W: Weights, sink: GradSink, dims: ViTDims, sv: ClsSaved, dlogits: torch.Tensor):
    N = sv.enc.N
    T = dims.num_patches + 1
    wname, bname = "lin_head.weight", "lin_head.bias"
    w = W.f(wname)
    if W.need(wname) or W.need(bname):
        dW, acc = sink.buf(wname, w)
        db, _ = sink.buf(bname, W.f(bname))
    else:
        dW = torch.empty(w.shape, dtype=torch.float32, device=w.device)
        db = torch.empty(w.shape[0], dtype=torch.float32, device=w.device)
        acc = False
    backbone = any(W.need(n) for n in W.names() if not n.startswith("lin_head."))
    dxn = ops.pool_head_bwd(sv.xn, w, dlogits, dW, db, N, T, dims.embed_dim, sv.spatial, want_dx=backbone, accumulate=acc)
    sink.ready(wname)
    sink.ready(bname)
    if backbone:
        encoder_bwd(W, sink, dims, sv.enc, dxn)
