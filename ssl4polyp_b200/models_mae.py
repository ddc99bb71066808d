"""Drop-in `MaskedAutoencoderViT` whose forward/backward run on hand-written sm_100a kernels.

Mirrors the public surface of /root/reference/src/ssl4polyp/models/mae/models_mae.py:
  constructor signature :25-28, `patchify` :95, `unpatchify` :109, `random_masking` :123, `forward_encoder` :150,
  `forward_decoder` :172, `forward_loss` :198, `forward(imgs, mask_ratio=0.75) -> (loss, pred, mask)` :216,
  factories `mae_vit_{base,large,huge}_patch{16,16,14}` (+ `_dec512d8b` names) :223-250,
and keeps the state-dict key names and shapes of §3.2 so reference checkpoints load unchanged.

What is different underneath: `forward` is ONE autograd node.  Its forward enqueues our kernels (bit-exact masking
from the single `torch.rand(N, L)` draw the reference makes, patch-embed GEMM on the kept patches only, fused
attention, tcgen05 GEMMs with bias/GELU/residual epilogues, patchify-free masked MSE) and its backward is the
hand-written schedule in `vit_core.mae_bwd`.  There is no CPU or eager-PyTorch fallback: a CPU tensor raises.
"""
from __future__ import annotations

from functools import partial

import torch
import torch.nn as nn

from . import ops, vit_core
from .layers import Block, PatchEmbed
from .param_store import ParamStore
from .util.pos_embed import get_2d_sincos_pos_embed


def _as_input(imgs: torch.Tensor) -> torch.Tensor:
    if not imgs.is_cuda:
        raise RuntimeError("ssl4polyp_b200: input must be a CUDA tensor (no CPU path)")
    if imgs.dtype != torch.float32:
        imgs = imgs.float()
    return imgs.contiguous()


class _MAEStep(torch.autograd.Function):
    """(imgs, noise, *params) -> (loss, pred, mask, ids_restore); one node for the whole model."""

    @staticmethod
    def forward(ctx, model, imgs, noise, len_keep, save, *params):
        loss, pred_full, mask, ids_restore, sv = vit_core.mae_fwd(
            model._store, model._dims, imgs, noise, len_keep, model.norm_pix_loss, save)
        ctx.model, ctx.sv = model, sv
        pred = pred_full[:, 1:, :]
        ctx.mark_non_differentiable(pred, mask, ids_restore)
        return loss, pred, mask, ids_restore

    @staticmethod
    def backward(ctx, dloss, *_unused):
        model, sv = ctx.model, ctx.sv
        if sv is None:
            raise RuntimeError("backward through a forward that ran without gradient tracking")
        ctx.sv = None
        sink = vit_core.FreshGrads()
        vit_core.mae_bwd(model._store, sink, model._dims, sv, dloss.detach().to(torch.float32).contiguous())
        return (None, None, None, None, None) + model._grads_in_order(sink)


class _EncoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, imgs, noise, len_keep, save, *params):
        ids_restore, mask, ids_keep = ops.mask_from_noise(noise, len_keep)
        latent, sv = vit_core.encoder_fwd(model._store, model._dims, imgs, ids_keep, len_keep, save)
        ctx.model, ctx.sv = model, sv
        ctx.mark_non_differentiable(mask, ids_restore)
        N = imgs.shape[0]
        return latent.view(N, len_keep + 1, -1).float(), mask, ids_restore

    @staticmethod
    def backward(ctx, dlatent, *_unused):
        model, sv = ctx.model, ctx.sv
        ctx.sv = None
        sink = vit_core.FreshGrads()
        dl = ops.cast_bf16(dlatent.detach().float().contiguous().view(-1, dlatent.shape[-1]))
        vit_core.encoder_bwd(model._store, sink, model._dims, sv, dl)
        return (None, None, None, None, None) + model._grads_in_order(sink)


class _DecoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, latent, ids_restore, save, *params):
        N, Te, D = latent.shape
        keep = Te - 1
        L = ids_restore.shape[1]
        # index plumbing (not arithmetic): invert the permutation to recover ids_keep and the mask
        ids_shuffle = torch.empty_like(ids_restore)
        ids_shuffle.scatter_(1, ids_restore, torch.arange(L, device=ids_restore.device).expand(N, L))
        ids_keep = ids_shuffle[:, :keep].contiguous()
        mask = (ids_restore >= keep).to(torch.float32)
        lat = ops.cast_bf16(latent.detach().float().contiguous().view(N * Te, D))
        pred_full, sv = vit_core.decoder_fwd(model._store, model._dims, lat, ids_restore.contiguous(), ids_keep, mask, N, save)
        ctx.model, ctx.sv = model, sv
        ctx.need_dlatent = latent.requires_grad
        return pred_full[:, 1:, :]

    @staticmethod
    def backward(ctx, dpred):
        model, sv = ctx.model, ctx.sv
        ctx.sv = None
        N, L, P = dpred.shape
        dfull = torch.zeros(N, L + 1, P, dtype=torch.bfloat16, device=dpred.device)
        dfull[:, 1:, :] = dpred
        sink = vit_core.FreshGrads()
        dlat = vit_core.decoder_bwd(model._store, sink, model._dims, sv, dfull, want_dlatent=True)
        dlat = dlat.view(N, sv.Te, -1).float() if ctx.need_dlatent else None
        return (None, dlat, None, None) + model._grads_in_order(sink)


class _LossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, imgs, pred, mask, norm_pix, p):
        loss, ws = ops.masked_mse_fwd_strided(imgs, pred, mask, norm_pix, p)
        ctx.save_for_backward(imgs, pred, mask, ws)
        ctx.norm_pix, ctx.p = norm_pix, p
        return loss

    @staticmethod
    def backward(ctx, dloss):
        imgs, pred, mask, ws = ctx.saved_tensors
        d = ops.masked_mse_bwd_strided(imgs, pred, mask, ctx.norm_pix, ctx.p, dloss.detach().float().contiguous(), ws)
        return None, d.float(), None, None, None


class MaskedAutoencoderViT(nn.Module):
    """Masked Autoencoder with a ViT backbone (reference models_mae.py:22), B200-native."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3,
                 embed_dim=1024, depth=24, num_heads=16,
                 decoder_embed_dim=512, decoder_depth=8, decoder_num_heads=16,
                 mlp_ratio=4., norm_layer=nn.LayerNorm, norm_pix_loss=False):
        super().__init__()
        # encoder
        self.patch_embed = PatchEmbed(img_size, patch_size, in_chans, embed_dim)
        n_patches = self.patch_embed.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, n_patches + 1, embed_dim), requires_grad=False)   # fixed sin-cos
        self.blocks = nn.ModuleList([Block(embed_dim, num_heads, mlp_ratio, qkv_bias=True, norm_layer=norm_layer)
                                     for _ in range(depth)])
        self.norm = norm_layer(embed_dim)
        # decoder
        self.decoder_embed = nn.Linear(embed_dim, decoder_embed_dim, bias=True)
        self.mask_token = nn.Parameter(torch.zeros(1, 1, decoder_embed_dim))
        self.decoder_pos_embed = nn.Parameter(torch.zeros(1, n_patches + 1, decoder_embed_dim), requires_grad=False)
        self.decoder_blocks = nn.ModuleList([Block(decoder_embed_dim, decoder_num_heads, mlp_ratio, qkv_bias=True,
                                                   norm_layer=norm_layer) for _ in range(decoder_depth)])
        self.decoder_norm = norm_layer(decoder_embed_dim)
        self.decoder_pred = nn.Linear(decoder_embed_dim, patch_size ** 2 * in_chans, bias=True)
        self.norm_pix_loss = norm_pix_loss

        self._dims = vit_core.ViTDims(
            img_size=img_size, patch_size=patch_size, in_chans=in_chans, embed_dim=embed_dim, depth=depth,
            num_heads=num_heads, decoder_embed_dim=decoder_embed_dim, decoder_depth=decoder_depth,
            decoder_num_heads=decoder_num_heads, ln_eps=float(self.norm.eps))
        self._store = ParamStore(self)
        self.initialize_weights()

    # ---- initialisation (reference models_mae.py:65-93) ---------------------------------------------------
    def initialize_weights(self):
        grid = int(self.patch_embed.num_patches ** .5)
        table = get_2d_sincos_pos_embed(self.pos_embed.shape[-1], grid, cls_token=True)
        self.pos_embed.data.copy_(torch.from_numpy(table).float().unsqueeze(0))
        dtable = get_2d_sincos_pos_embed(self.decoder_pos_embed.shape[-1], grid, cls_token=True)
        self.decoder_pos_embed.data.copy_(torch.from_numpy(dtable).float().unsqueeze(0))
        w = self.patch_embed.proj.weight.data
        torch.nn.init.xavier_uniform_(w.view([w.shape[0], -1]))     # conv weight initialised like an nn.Linear
        torch.nn.init.normal_(self.cls_token, std=.02)
        torch.nn.init.normal_(self.mask_token, std=.02)
        self.apply(self._init_weights)

    def _init_weights(self, m):
        if isinstance(m, nn.Linear):
            torch.nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.LayerNorm):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    # ---- plumbing ---------------------------------------------------------------------------------------
    def _param_list(self):
        self._store.sync()
        names = list(self._store.params.keys())
        self._param_names = names
        return [self._store.params[n] for n in names]

    def _grads_in_order(self, sink):
        out = []
        for n in self._param_names:
            p = self._store.params[n]
            out.append(sink.grads.get(n) if p.requires_grad else None)
        return tuple(out)

    def _want_grad(self, *tensors):
        return torch.is_grad_enabled() and (any(p.requires_grad for p in self._store.params.values())
                                            or any(t.requires_grad for t in tensors if isinstance(t, torch.Tensor)))

    def _len_keep(self, mask_ratio):
        return int(self.patch_embed.num_patches * (1 - mask_ratio))       # Python-float arithmetic, as models_mae.py:130

    # ---- patch layout (reference :95-121) ---------------------------------------------------------------
    def patchify(self, imgs):
        """imgs (N, 3, H, W) -> x (N, L, patch_size**2 * 3)"""
        p = self.patch_embed.patch_size[0]
        assert imgs.shape[2] == imgs.shape[3] and imgs.shape[2] % p == 0
        return ops.patchify(_as_input(imgs), p)

    def unpatchify(self, x):
        """x (N, L, patch_size**2 * 3) -> imgs (N, 3, H, W)"""
        p = self.patch_embed.patch_size[0]
        h = int(x.shape[1] ** .5)
        assert h * h == x.shape[1]
        return ops.unpatchify(_as_input(x), p, self._dims.in_chans)

    def random_masking(self, x, mask_ratio):
        """Per-sample random masking by per-sample shuffling (argsort of uniform noise), reference :123-148.
        x: [N, L, D] -> (x_masked [N, keep, D], mask [N, L] fp32 (0 keep, 1 remove), ids_restore [N, L] int64)."""
        N, L, D = x.shape
        len_keep = int(L * (1 - mask_ratio))
        noise = torch.rand(N, L, device=x.device)
        ids_restore, mask, ids_keep = ops.mask_from_noise(noise, len_keep)
        x_masked = ops.gather_tokens(_as_input(x), ids_keep)
        return x_masked, mask, ids_restore

    # ---- the three stages, each one autograd node (reference :150-214) ----------------------------------
    def forward_encoder(self, x, mask_ratio):
        x = _as_input(x)
        self.patch_embed.check_input(x)
        params = self._param_list()
        noise = torch.rand(x.shape[0], self.patch_embed.num_patches, device=x.device)
        return _EncoderFn.apply(self, x, noise, self._len_keep(mask_ratio), self._want_grad(), *params)

    def forward_decoder(self, x, ids_restore):
        params = self._param_list()
        return _DecoderFn.apply(self, x, ids_restore, self._want_grad(x), *params)

    def forward_loss(self, imgs, pred, mask):
        """imgs [N,3,H,W], pred [N,L,p*p*3], mask [N,L] (0 keep, 1 remove) -> masked mean of per-patch MSE."""
        return _LossFn.apply(_as_input(imgs), pred, mask, self.norm_pix_loss, self.patch_embed.patch_size[0])

    def forward(self, imgs, mask_ratio=0.75):
        imgs = _as_input(imgs)
        self.patch_embed.check_input(imgs)
        params = self._param_list()
        noise = torch.rand(imgs.shape[0], self.patch_embed.num_patches, device=imgs.device)
        loss, pred, mask, _ = _MAEStep.apply(self, imgs, noise, self._len_keep(mask_ratio), self._want_grad(), *params)
        return loss, pred, mask

    def forward_with_noise(self, imgs, noise, mask_ratio=0.75):
        """Same as `forward` with the uniform noise supplied by the caller (parity tests, CUDA-graph replay).
        Returns (loss, pred, mask, ids_restore)."""
        imgs = _as_input(imgs)
        params = self._param_list()
        return _MAEStep.apply(self, imgs, noise.contiguous(), self._len_keep(mask_ratio), self._want_grad(), *params)


def mae_vit_base_patch16_dec512d8b(**kwargs):
    return MaskedAutoencoderViT(patch_size=16, embed_dim=768, depth=12, num_heads=12,
                                decoder_embed_dim=512, decoder_depth=8, decoder_num_heads=16,
                                mlp_ratio=4, norm_layer=partial(nn.LayerNorm, eps=1e-6), **kwargs)


def mae_vit_large_patch16_dec512d8b(**kwargs):
    return MaskedAutoencoderViT(patch_size=16, embed_dim=1024, depth=24, num_heads=16,
                                decoder_embed_dim=512, decoder_depth=8, decoder_num_heads=16,
                                mlp_ratio=4, norm_layer=partial(nn.LayerNorm, eps=1e-6), **kwargs)


def mae_vit_huge_patch14_dec512d8b(**kwargs):
    return MaskedAutoencoderViT(patch_size=14, embed_dim=1280, depth=32, num_heads=16,
                                decoder_embed_dim=512, decoder_depth=8, decoder_num_heads=16,
                                mlp_ratio=4, norm_layer=partial(nn.LayerNorm, eps=1e-6), **kwargs)


# recommended architectures (decoder: 512 dim, 8 blocks)
mae_vit_base_patch16 = mae_vit_base_patch16_dec512d8b
mae_vit_large_patch16 = mae_vit_large_patch16_dec512d8b
mae_vit_huge_patch14 = mae_vit_huge_patch14_dec512d8b
