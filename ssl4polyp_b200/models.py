"""Drop-in classifier backbones of /root/reference/src/ssl4polyp/models/models.py on the sm_100a kernels:

  `ViT_from_MAE`            (models.py:143-222)  MAE encoder without masking + `lin_head`
  `VisionTransformer_from_Any` (models.py:26-140)   timm-0.4.12-shaped ViT (learned pos_embed) + `lin_head`

Same constructor arguments, attribute names (`blocks`, `lin_head`, `head`, `frozen`, `dense`, `out_token`,
`patch_embed.{patch_size,num_patches,proj}`, `pos_embed`, `cls_token`, `norm`) and state-dict keys, so
`classification/finetune.py:49-91` and the checkpoint hand-off (`load_my_state_dict`, models.py:186-194) work unchanged.
`forward(imgs) -> logits` is one autograd node over `vit_core.classifier_fwd/bwd`; frozen parameters prune the
backward (no wgrad for them, and the chain stops below the first trainable block).

Out of scope: the `dense=` DPT segmentation decoder — never enabled by the reference's call
sites (`utils/__init__.py:29-43`, `train_classification.py:5645-5655`); a truthy `dense` raises.
"""
from __future__ import annotations

from functools import partial

import torch
import torch.nn as nn

from . import ops, vit_core
from .layers import Block, PatchEmbed
from .models_mae import MaskedAutoencoderViT, _as_input
from .param_store import ParamStore


class _ClassifierFn(torch.autograd.Function):
    """(imgs, *params) -> logits fp32 [N, num_classes]."""

    @staticmethod
    def forward(ctx, model, imgs, save, *params):
        logits, sv = vit_core.classifier_fwd(model._store, model._dims, imgs, model.out_token == "spatial", save)
        ctx.model, ctx.sv = model, sv
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        model, sv = ctx.model, ctx.sv
        if sv is None:
            raise RuntimeError("backward through a forward that ran without gradient tracking")
        ctx.sv = None
        sink = vit_core.FreshGrads()
        vit_core.classifier_bwd(model._store, sink, model._dims, sv, dlogits.detach().float().contiguous())
        return (None, None, None) + model._grads_in_order(sink)


class _FeaturesFn(torch.autograd.Function):
    """(imgs, *params) -> norm(blocks(tokens)) as fp32 [N, 1+L, D] (the headless `forward_encoder`)."""

    @staticmethod
    def forward(ctx, model, imgs, save, *params):
        N = imgs.shape[0]
        xn, sv = vit_core.encoder_fwd(model._store, model._dims, imgs, None, model._dims.num_patches, save)
        ctx.model, ctx.sv = model, sv
        return xn.view(N, model._dims.num_patches + 1, -1).float()

    @staticmethod
    def backward(ctx, dx):
        model, sv = ctx.model, ctx.sv
        ctx.sv = None
        sink = vit_core.FreshGrads()
        d = ops.cast_bf16(dx.detach().float().contiguous().view(-1, dx.shape[-1]))
        vit_core.encoder_bwd(model._store, sink, model._dims, sv, d)
        return (None, None, None) + model._grads_in_order(sink)


class _ClassifierMixin:
    """forward()/forward_encoder() shared by the two backbones."""

    def _features(self, imgs):
        imgs = _as_input(imgs)
        self.patch_embed.check_input(imgs)
        params = self._param_list()
        return _FeaturesFn.apply(self, imgs, self._want_grad(), *params)

    def _classify(self, imgs, has_head: bool):
        if self.dense:
            raise NotImplementedError("dense (DPT) decoding is outside the B200 hot path (see module docstring)")
        if has_head and self.out_token in ("cls", "spatial"):
            imgs = _as_input(imgs)
            self.patch_embed.check_input(imgs)
            params = self._param_list()
            return _ClassifierFn.apply(self, imgs, self._want_grad(), *params)
        x = self._features(imgs)
        if self.out_token == "cls":
            x = x[:, 0]
        elif self.out_token == "spatial":
            x = x[:, 1:].mean(1)
        if has_head:
            x = torch.nn.functional.linear(x, self.lin_head.weight, self.lin_head.bias)   # unusual out_token only
        return x


class ViT_from_MAE(_ClassifierMixin, MaskedAutoencoderViT):
    def __init__(self, weight_path, head, num_classes, frozen, dense, embed_dim, depth, num_heads, out_token):
        super().__init__(patch_size=16, embed_dim=embed_dim, depth=depth, num_heads=num_heads,
                         decoder_embed_dim=512, decoder_depth=8, decoder_num_heads=16, mlp_ratio=4,
                         norm_layer=partial(nn.LayerNorm, eps=1e-6))
        if weight_path is not None:
            weights = torch.load(weight_path, map_location="cpu", weights_only=False)["model"]   # checkpoints carry argparse.Namespace ("args", util/misc.py:312-318)
            self.load_my_state_dict(weights)
        del self.mask_token
        del self.decoder_embed
        del self.decoder_blocks
        del self.decoder_norm
        del self.decoder_pred
        self.head = head
        if head:
            self.lin_head = nn.Linear(embed_dim, num_classes)
        self.frozen = frozen
        self.dense = dense
        if dense:
            raise NotImplementedError("dense (DPT) decoding is outside the B200 hot path (see module docstring)")
        self.out_token = out_token
        self._dims.decoder_depth = 0

    def load_my_state_dict(self, state_dict):
        own = self.state_dict()
        n = 0
        for name, value in state_dict.items():
            if name not in own:
                continue
            own[name].copy_(value)
            n += 1
        self._store.mark_dirty()
        print(f"Successfully loaded params for {n} items")

    def forward_encoder(self, x):
        return self._features(x)

    def forward(self, imgs):
        return self._classify(imgs, bool(self.head))


class VisionTransformer_from_Any(_ClassifierMixin, nn.Module):
    """timm-0.4.12-shaped ViT-B/16 (learned `pos_embed`, `cls_token`, `qkv_bias=True`, LayerNorm eps 1e-6) + `lin_head`,
    the `--pretraining random / ImageNet_class` backbone of the reference (models.py:26-140).  State-dict keys are
    timm's (`cls_token, pos_embed, patch_embed.proj.*, blocks.i.*, norm.*`) plus `lin_head.*`.

    Kept on purpose: timm 0.4.12 holds `blocks` in an `nn.Sequential`, so `classification/finetune.py:76-77`
    (`isinstance(blocks, (nn.ModuleList, list, tuple))`) does not unfreeze tail blocks for this backbone in the
    `head+1` / `head+2` modes; the reference behaves the same way.
    `ImageNet_weights=True` downloads an .npz in the reference (models.py:52-56); there is no network here, so pass a
    local torch checkpoint path through `load_pretrained` instead.
    """

    def __init__(self, head, num_classes, frozen, dense, embed_dim, depth, num_heads, out_token, ImageNet_weights=False):
        super().__init__()
        norm_layer = partial(nn.LayerNorm, eps=1e-6)
        self.num_features = self.embed_dim = embed_dim
        self.patch_embed = PatchEmbed(224, 16, 3, embed_dim)
        n_patches = self.patch_embed.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, n_patches + 1, embed_dim))
        self.pos_drop = nn.Dropout(p=0.0)
        self.blocks = nn.Sequential(*[Block(embed_dim, num_heads, 4., qkv_bias=True, norm_layer=norm_layer) for _ in range(depth)])
        self.norm = norm_layer(embed_dim)
        self.pre_logits = nn.Identity()
        self.head = nn.Identity()                                  # models.py:58
        self.head_bool = head
        if head:
            self.lin_head = nn.Linear(embed_dim, num_classes)
        self.frozen = frozen
        self.dense = dense
        if dense:
            raise NotImplementedError("dense (DPT) decoding is outside the B200 hot path (see module docstring)")
        self.out_token = out_token
        # timm 0.4.12 VisionTransformer.init_weights(''): trunc_normal(.02) tables and Linear weights, zero biases
        nn.init.trunc_normal_(self.pos_embed, std=.02)
        nn.init.trunc_normal_(self.cls_token, std=.02)
        for m in list(self.blocks.modules()) + [self.norm]:
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.LayerNorm):
                nn.init.zeros_(m.bias)
                nn.init.ones_(m.weight)
        self._dims = vit_core.ViTDims(img_size=224, patch_size=16, in_chans=3, embed_dim=embed_dim, depth=depth,
                                      num_heads=num_heads, ln_eps=1e-6)
        self._store = ParamStore(self)
        if ImageNet_weights:
            raise RuntimeError("ImageNet_weights=True needs the augreg .npz download (models.py:52-56); no network here: "
                               "construct with ImageNet_weights=False and call load_pretrained(<local torch checkpoint>)")

    # ---- arena plumbing (same contract as MaskedAutoencoderViT) ----------------------------------------------
    _param_list = MaskedAutoencoderViT._param_list
    _grads_in_order = MaskedAutoencoderViT._grads_in_order
    _want_grad = MaskedAutoencoderViT._want_grad

    def load_pretrained(self, checkpoint_path):
        """torch checkpoints only (models.py:87-122: `state_dict` / `model` / `model_state` / `weights` / `params` or a bare
        state dict, non-strict)."""
        if str(checkpoint_path).lower().endswith(".npz"):
            raise NotImplementedError("numpy (.npz) checkpoints need timm's _load_weights, which is not part of the hot path")
        ckpt = torch.load(checkpoint_path, map_location="cpu", weights_only=False)
        sd = None
        if isinstance(ckpt, dict):
            for key in ("state_dict", "model", "model_state", "weights", "params"):
                if isinstance(ckpt.get(key), dict):
                    sd = ckpt[key]
                    break
            if sd is None and all(isinstance(k, str) for k in ckpt.keys()):
                sd = ckpt
        if not isinstance(sd, dict):
            print(f"Warning: unsupported checkpoint format for {checkpoint_path!s}: {type(ckpt)}")
            return
        missing, unexpected = self.load_state_dict(sd, strict=False)
        self._store.mark_dirty()
        if missing:
            print(f"Missing keys when loading pretrained weights: {missing}")
        if unexpected:
            print(f"Unexpected keys when loading pretrained weights: {unexpected}")

    def forward_features(self, x):
        return self._features(x)

    def forward(self, x):
        return self._classify(x, bool(self.head_bool))
