"""Flat parameter arena: every parameter of a model lives in ONE contiguous fp32 buffer in HBM (each tensor
aligned to 256 elements = 1 KiB), next to a bf16 shadow of the same layout that the tensor-core GEMMs read.

Why: (1) one cast kernel refreshes all bf16 GEMM operands after an optimiser step instead of ~100 small ones;
(2) the fused AdamW kernel (`ops.adamw_flat`) and the bucketed gradient all-reduce (`engine.py`) walk flat ranges;
(3) TMA descriptors of the weights stay valid for the life of the arena.  `nn.Parameter` objects keep their
identity (optimisers and DDP hold references to them): only `.data` is re-pointed at a view of the arena, so
`state_dict()` keys, shapes and `load_state_dict` behave exactly as in the reference.
"""
from __future__ import annotations

import torch

from . import ops
from .vit_core import Weights

ALIGN = 256     # elements; also the weight-decay flag granularity of ops.adamw_flat


class ParamStore(Weights):
    def __init__(self, module: torch.nn.Module):
        self.module = module
        self.flat: torch.Tensor | None = None          # fp32 [total]
        self.shadow: torch.Tensor | None = None        # bf16 [total]
        self.offsets: dict[str, tuple[int, int]] = {}  # name -> (start, numel)
        self.params: dict[str, torch.nn.Parameter] = {}
        self.total = 0
        self._version_sum = -1
        self._shadow_views: dict[str, torch.Tensor] = {}
        self.shadow_fresh_token = None                 # set by the fused optimiser when it rewrote the shadow itself

    # ---- layout -------------------------------------------------------------------------------------
    def _named(self):
        seen, out = set(), []
        for name, p in self.module.named_parameters():
            if id(p) in seen:
                continue
            seen.add(id(p))
            out.append((name, p))
        return out

    def materialize(self) -> None:
        named = self._named()
        if not named:
            raise RuntimeError("model has no parameters")
        dev = named[0][1].device
        if dev.type != "cuda":
            raise RuntimeError("ssl4polyp_b200 runs on CUDA (sm_100a) only: move the model with .to('cuda') first "
                               "(there is no CPU path)")
        for name, p in named:
            if p.device != dev:
                raise RuntimeError(f"parameter {name} is on {p.device}, expected {dev}")
            if p.dtype != torch.float32:
                raise RuntimeError(f"parameter {name} has dtype {p.dtype}; master weights must be fp32 "
                                   "(bf16 operands are derived by the library)")
        self.offsets.clear(); self.params.clear(); self._shadow_views.clear()
        off = 0
        for name, p in named:
            n = p.numel()
            self.offsets[name] = (off, n)
            self.params[name] = p
            off += (n + ALIGN - 1) // ALIGN * ALIGN
        self.total = off
        flat = torch.zeros(off, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for name, p in named:
                a, n = self.offsets[name]
                view = flat[a:a + n].view(p.shape)
                view.copy_(p.data)
                p.data = view
        self.flat = flat
        self.shadow = torch.empty(off, dtype=torch.bfloat16, device=dev)
        self._version_sum = -1

    def _intact(self) -> bool:
        if self.flat is None:
            return False
        base, esz = self.flat.data_ptr(), 4
        named = self._named()
        if len(named) != len(self.offsets):
            return False
        for name, p in named:
            o = self.offsets.get(name)
            if o is None or self.params.get(name) is not p or p.data_ptr() != base + o[0] * esz:
                return False
        return True

    def sync(self) -> None:
        """Call before every forward: (re)build the arena if the model was moved/re-parameterised, and refresh the
        bf16 shadow if any parameter was written since the last refresh."""
        if not self._intact():
            self.materialize()
        v = 0
        for p in self.params.values():
            v += p._version
        if v != self._version_sum:
            ops.cast_bf16(self.flat, self.shadow)
            self._version_sum = v

    def mark_dirty(self) -> None:
        self._version_sum = -1

    def mark_shadow_fresh(self) -> None:
        """The fused optimiser wrote flat and shadow together: skip the next cast."""
        v = 0
        for p in self.params.values():
            v += p._version
        self._version_sum = v

    # ---- Weights interface --------------------------------------------------------------------------
    def f(self, name):
        return self.params[name].data

    def b(self, name):
        t = self._shadow_views.get(name)
        if t is None or t.data_ptr() != self.shadow.data_ptr() + self.offsets[name][0] * 2:
            a, n = self.offsets[name]
            shape = self.params[name].shape
            t = self.shadow[a:a + n].view(shape[0], -1)
            self._shadow_views[name] = t
        return t

    def need(self, name):
        p = self.params.get(name)
        return p is not None and p.requires_grad

    def has(self, name):
        return name in self.params

    def names(self):
        return self.params.keys()
