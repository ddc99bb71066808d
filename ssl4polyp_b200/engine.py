"""Fused data-parallel training step for the MAE / classifier hot path.

What `engine_pretrain.train_one_epoch` does per iteration in the reference
(/root/reference/src/ssl4polyp/models/mae/engine_pretrain.py:47-81: lr schedule, autocast forward, backward through DDP,
~600 `.any()` host syncs of `detect_grad_anomalies`, GradScaler step, `cuda.synchronize`, scalar all-reduce) is done here as:

  forward  : vit_core.mae_fwd on our kernels (bf16 tensor-core operands, fp32 residual/statistics/loss)
  backward : vit_core.mae_bwd writing weight gradients straight into ONE flat fp32 arena laid out like the
             parameter arena (param_store.ParamStore)
  DP       : the arena is cut into buckets in reverse-forward order; the moment the last gradient of a bucket has been
             enqueued an event is recorded and NCCL all-reduces that bucket on a side stream while backward continues
             (one process per GPU; the SUM is turned into the mean by `grad_scale = 1/world` inside the optimiser)
  update   : ops.adamw_flat — one pass over p/g/m/v for all ~250 tensors, decoupled weight decay by timm's rule
             (ndim == 1 or name.endswith('.bias') -> no decay; main_pretrain.py:217, timm optim_factory.add_weight_decay),
             bf16 shadow weights refreshed in the same pass, non-finite gradients counted on the device
             (util/misc.py:387-400) — no host sync anywhere in the step.
"""
from __future__ import annotations

import contextlib
import gc
import math
import os

import torch
import torch.distributed as dist

from . import ops, vit_core
from .dp import BucketReducer, broadcast_parameters, plan_buckets_tapered
from .param_store import ALIGN


class ArenaSink(vit_core.GradSink):
    """Gradients go straight into the engine's flat arena.  With `overlap` the weight-gradient work of the backward
    (wgrad GEMM + split-K reduce + bias column sums: ~45 % of the backward's kernel time, none of it on the dgrad chain)
    is enqueued on the engine's side stream: its CTAs fill the SMs that the critical chain's kernels leave idle at their
    tails and between launches, and the ~300 short reduction kernels disappear from the critical path."""

    LAG = 2                                     # side work of a block is joined this many blocks later

    def __init__(self, engine: "TrainEngine"):
        self.e = engine
        self.accumulate = False
        self.overlap = False
        self._kept = []                         # operands the side stream may still be reading (this block)
        self._lagging = []                      # [(event on the side stream, operands)] of earlier blocks

    def buf(self, name, like):
        return self.e.grad_views[name], self.accumulate

    def ready(self, name):
        self.e._on_ready(name)

    def side(self, *operands):
        if not self.overlap:
            return contextlib.nullcontext()
        aux = self.e.aux_stream
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        aux.wait_event(ev)                      # the operands were produced on the main stream
        # They were also ALLOCATED there: once the schedule drops its references the caching allocator may hand their
        # memory to a later main-stream kernel while the side stream still reads it -> keep them until the join.
        self._kept.extend(operands)
        return torch.cuda.stream(aux)

    def block_boundary(self):
        if not self.overlap:
            return
        if self._kept:
            ev = torch.cuda.Event()
            ev.record(self.e.aux_stream)
            self._lagging.append((ev, self._kept))
            self._kept = []
        while len(self._lagging) > self.LAG:
            ev, _ = self._lagging.pop(0)
            torch.cuda.current_stream().wait_event(ev)

    def join(self):
        """All side work done before whatever the main stream enqueues next (the optimiser, or the caller)."""
        if not self.overlap:
            return
        torch.cuda.current_stream().wait_stream(self.e.aux_stream)
        self._kept, self._lagging = [], []


class TrainEngine:
    def __init__(self, model, lr: float, weight_decay: float = 0.05, betas=(0.9, 0.95), eps: float = 1e-8,
                 process_group=None, bucket_mb: float = 64.0, no_decay_rule=None):
        self.model = model
        store = model._store
        store.sync()
        self.store = store
        broadcast_parameters(store.flat, process_group)       # DDP's constructor broadcast (main_pretrain.py:212-214)
        store.mark_dirty()
        store.sync()                                           # refresh the bf16 shadow from the broadcast weights
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        dev = store.flat.device
        self.device = dev
        n = store.total
        self.grads = self._alloc_grad_arena(n, dev, process_group)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.nonfinite = torch.zeros(1, dtype=torch.int32, device=dev)
        self.t = 0
        # per-step scalars live on the device so a captured step can be replayed: hyper = [lr, optimiser steps done]
        self.hyper = torch.tensor([float(lr), 0.0], dtype=torch.float32, device=dev)
        self._lr_on_device = float(lr)
        self._graph = None
        self._graph_warm = 0
        self._graph_key = None
        self._static_in = None
        self._static_out = None
        self.grad_views = {}
        flags = torch.zeros(n // ALIGN, dtype=torch.uint8)
        rule = no_decay_rule or (lambda name, p: p.ndim == 1 or name.endswith(".bias"))
        for name, (a, cnt) in store.offsets.items():
            p = store.params[name]
            self.grad_views[name] = self.grads[a:a + cnt].view(p.shape)
            if p.requires_grad and not rule(name, p):
                flags[a // ALIGN:(a + cnt + ALIGN - 1) // ALIGN] = 1
        self.wd_flags = flags.to(dev)
        # ---- data parallel (dp.py): bucket plan over the gradient arena + overlapped all-reduce
        trainable = [n for n in store.offsets if store.params[n].requires_grad]
        bucket_mb = float(os.environ.get("SSL4P_DP_BUCKET_MB", bucket_mb))
        self.buckets = plan_buckets_tapered(store.offsets, trainable, int(bucket_mb * (1 << 20) / 4), ALIGN)
        self.reducer = BucketReducer(self.grads, self.buckets, process_group)
        self.world = self.reducer.world
        if self.world > 1 and dev.type == "cuda":
            # optional: persistent grids leave SSL4P_DP_SMS SMs to the NCCL kernels of the overlapped all-reduce (see
            # dp.configure_overlap).  Measured on 2 x B200 (scripts/dp_sweep.sh): 0 -> 26.4 ms, 2 -> 28.9, 4 -> 27.5,
            # 16 -> 27.4 ms/step, so sharing the SMs is the default.
            from . import _lib
            _lib.load().ssl4p_reserve_sms(int(os.environ.get("SSL4P_DP_SMS", "0")))
        self.sink = ArenaSink(self)
        self._reduce_now = True
        # second stream for the weight-gradient half of the backward (ArenaSink.side); SSL4P_WGRAD_OVERLAP=0 disables it
        self.aux_stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self.sink.overlap = self.aux_stream is not None and os.environ.get("SSL4P_WGRAD_OVERLAP", "1") != "0"
        # third stream: AdamW bucket by bucket behind the backward / the bucket all-reduces; SSL4P_BUCKET_ADAMW=0 -> one pass
        self.opt_stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self.bucket_adamw = self.opt_stream is not None and os.environ.get("SSL4P_BUCKET_ADAMW", "1") != "0"

    def _on_ready(self, name):
        if self._reduce_now:
            self.reducer.ready(name)

    # ---- one optimisation step -------------------------------------------------------------------------------
    def backward_and_step(self, run_backward, lr=None, accumulate: bool = False, do_step: bool = True, apply: bool = True):
        """run_backward(sink) enqueues the backward; gradients are reduced across ranks as buckets complete when
        `do_step` (the reference all-reduces on every micro-step, engine_pretrain.py:64-72; summing locally and
        reducing once on the boundary gives the same mean with 1/accum of the traffic).  apply=False stops after the
        reduction: the gradient arena then holds the SUM over ranks (checks: `PretrainEngine.gradients`)."""
        self.sink.accumulate = accumulate
        self._reduce_now = do_step
        self.reducer.reset()
        main = torch.cuda.current_stream()
        self.reducer.producer_streams = [main, self.aux_stream] if self.sink.overlap else None
        stepping = do_step and apply
        if stepping and not torch.cuda.is_current_stream_capturing():
            self._set_lr(self.lr if lr is None else lr)          # before the first bucket's AdamW can run
            self.t += 1
        self.reducer.on_bucket = self._adamw_bucket if (stepping and self.bucket_adamw) else None
        run_backward(self.sink)
        self.sink.join()
        if not do_step:
            return
        self.reducer.finish()
        if not apply:
            return
        b1, b2 = self.betas
        if self.bucket_adamw:
            main.wait_stream(self.opt_stream)
            ops.adamw_advance_step(self.hyper)
        else:
            ops.adamw_flat_dev(self.store.flat, self.grads, self.exp_avg, self.exp_avg_sq, self.store.shadow, self.wd_flags,
                               self.hyper, b1, b2, self.eps, self.wd, grad_scale=1.0 / self.world, nonfinite=self.nonfinite)
        self.store.mark_shadow_fresh()

    @staticmethod
    def _alloc_grad_arena(n, dev, group):
        """The gradient arena.  SSL4P_DP_SYMM=1 (experiment, world > 1): allocated as symmetric memory and exchanged with
        torch's in-switch (NVLS multimem) all-reduce instead of NCCL - see dp.BucketReducer."""
        import torch.distributed as dist
        if (os.environ.get("SSL4P_DP_SYMM", "0") == "1" and dev.type == "cuda" and dist.is_available() and dist.is_initialized()
                and dist.get_world_size(group) > 1):
            import torch.distributed._symmetric_memory as symm_mem
            g = group if group is not None else dist.group.WORLD
            t = symm_mem.empty(n, dtype=torch.float32, device=dev)
            symm_mem.rendezvous(t, g.group_name)
            t.zero_()
            t._ssl4p_symm_group = g.group_name
            return t
        return torch.zeros(n, dtype=torch.float32, device=dev)

    def _adamw_bucket(self, i: int, deps) -> None:
        """AdamW over bucket i's arena range on the optimiser stream, after `deps` (its all-reduce and the backward's
        position one bucket later: dp.BucketReducer).  Parameters outside every bucket are frozen: nothing to update."""
        b = self.buckets[i]
        for ev in deps:
            self.opt_stream.wait_event(ev)
        b1, b2 = self.betas
        with torch.cuda.stream(self.opt_stream):
            ops.adamw_flat_dev_range(self.store.flat, self.grads, self.exp_avg, self.exp_avg_sq, self.store.shadow, self.wd_flags,
                                     b.start, b.end, self.hyper, b1, b2, self.eps, self.wd, grad_scale=1.0 / self.world,
                                     nonfinite=self.nonfinite)

    def _set_lr(self, lr: float) -> None:
        if lr != self._lr_on_device:
            self.hyper[0:1].fill_(float(lr))
            self._lr_on_device = float(lr)

    # ---- CUDA-graph replay of a whole step ---------------------------------------------------------------------
    def input_buffer(self, *shapes_dtypes):
        """Static device tensors the captured step reads its batch from; write the batch here (e.g. the H2D copy)
        to skip the staging copy in `step`."""
        key = tuple((tuple(s), d) for s, d in shapes_dtypes)
        if self._static_in is None or self._graph_key != key:
            self._static_in = [torch.empty(s, dtype=d, device=self.device) for s, d in shapes_dtypes]
            self._graph_key, self._graph, self._graph_warm = key, None, 0
        return self._static_in

    def _replay(self, eager_step, inputs, lr):
        """Run `eager_step(*static_inputs)` from a CUDA graph (the first two calls run eagerly, the third
        captures).  One cudaGraphLaunch replaces the ~500 kernel launches of the step."""
        static = self.input_buffer(*[(t.shape, t.dtype) for t in inputs])
        for dst, src in zip(static, inputs):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self._set_lr(self.lr if lr is None else lr)
        if self._graph is None:
            if self._graph_warm < 2:                          # lazy attribute / NCCL / workspace initialisation happens
                self._graph_warm += 1                         # in ordinary eager steps on the caller's own batches
                return eager_step(*static, lr=lr)
            # Python's cyclic collector must not run during the capture: if it frees a CUDAGraph / pinned buffer of an engine
            # that went out of use earlier (cudaGraphExecDestroy, cudaFreeHost), the capture is invalidated
            # (cudaErrorStreamCaptureInvalidated; torch.cuda.graph no longer collects on entry).  Collect now, then hold it off.
            gc.collect()
            torch.cuda.synchronize()
            gc_was_enabled = gc.isenabled()
            gc.disable()
            try:
                g = torch.cuda.CUDAGraph()
                # thread_local: CUDA calls of OTHER threads (a DataLoader's pin-memory thread allocating, NCCL's watchdog) do
                # not touch this capture; in the default global mode any cudaHostAlloc elsewhere invalidates it
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    self._static_out = eager_step(*static, lr=lr)   # lr is read from self.hyper on the device when replayed
            finally:
                if gc_was_enabled:
                    gc.enable()
            self._graph = g
        self._graph.replay()
        self.t += 1
        return self._static_out

    def nonfinite_count(self) -> int:
        """Number of optimiser blocks that saw a non-finite gradient so far (host sync: call off the hot path)."""
        return int(self.nonfinite.item())

    # ---- optimiser state in torch.optim.AdamW's layout (checkpoint hand-off, util/misc.py:306-352) -----------
    def _group_plan(self):
        """[(group extras, [parameter names])] in the order the reference builds its optimizer param_groups."""
        named = [(n, self.store.params[n]) for n in self.store.offsets]
        return timm_weight_decay_plan(named, self.wd)

    def optimizer_state_dict(self):
        """`torch.optim.AdamW.state_dict()` of the optimizer the reference would have built over this model: the same
        param_groups (order, 'params' index lists, hyper-parameters) and per-parameter `step` / `exp_avg` / `exp_avg_sq`,
        so the reference's `optimizer.load_state_dict` (util/misc.py:349) accepts our checkpoints.  `param_names`
        (index -> state-dict name) and `t` are extra keys torch ignores."""
        groups, names, state = [], [], {}
        b1, b2 = self.betas
        for extras, gnames in self._group_plan():
            idx = list(range(len(names), len(names) + len(gnames)))
            names += gnames
            g = {"lr": self.lr, "betas": (b1, b2), "eps": self.eps, "weight_decay": self.wd, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": True}
            g.update(extras)
            g["params"] = idx
            groups.append(g)
        for i, nme in enumerate(names):
            if not self.store.params[nme].requires_grad or self.t == 0:
                continue                                       # torch creates state lazily, on the first step with a gradient
            a, cnt = self.store.offsets[nme]
            shape = self.store.params[nme].shape
            state[i] = {"step": torch.tensor(float(self.t)),
                        "exp_avg": self.exp_avg[a:a + cnt].view(shape).clone(),
                        "exp_avg_sq": self.exp_avg_sq[a:a + cnt].view(shape).clone()}
        return {"state": state, "param_groups": groups, "param_names": names, "t": self.t}

    def load_optimizer_state_dict(self, sd):
        """Accepts our own checkpoints and the reference's (`torch.optim.AdamW.state_dict()` without `param_names`): the
        index -> name map is then rebuilt from the reference's group construction order; every shape is checked."""
        names = sd.get("param_names")
        if names is None:
            names = [n for _, g in self._group_plan() for n in g]
            n_idx = sum(len(g["params"]) for g in sd["param_groups"])
            if n_idx != len(names):
                raise RuntimeError(f"optimizer state has {n_idx} parameters in {len(sd['param_groups'])} groups; this model's "
                                   f"reference optimizer would have {len(names)} - not a checkpoint of this architecture")
        self.t = int(sd.get("t", 0))
        steps = set()
        for i, st in sd["state"].items():
            nme = names[int(i)]
            if nme not in self.store.offsets:
                raise RuntimeError(f"optimizer state names an unknown parameter {nme!r}")
            a, cnt = self.store.offsets[nme]
            shape = tuple(self.store.params[nme].shape)
            for k in ("exp_avg", "exp_avg_sq"):
                if tuple(st[k].shape) != shape:
                    raise RuntimeError(f"optimizer state {k} of {nme} has shape {tuple(st[k].shape)}, expected {shape}")
            self.exp_avg[a:a + cnt].copy_(st["exp_avg"].reshape(-1))
            self.exp_avg_sq[a:a + cnt].copy_(st["exp_avg_sq"].reshape(-1))
            steps.add(int(float(st["step"])))
        if "t" not in sd and steps:
            if len(steps) > 1:
                raise RuntimeError(f"per-parameter step counts differ ({sorted(steps)}); the fused AdamW keeps one counter")
            self.t = steps.pop()
        self.hyper[1:2].fill_(float(self.t))


def timm_weight_decay_plan(named_params, weight_decay: float):
    """timm 0.4.12 `optim_factory.add_weight_decay(model, wd)` as called at main_pretrain.py:217: walk
    `named_parameters()`, skip frozen tensors, `ndim == 1 or name.endswith(".bias")` -> the wd-0 group; the result is
    `[{'params': no_decay, 'weight_decay': 0.}, {'params': decay, 'weight_decay': wd}]` in THAT order, which fixes the
    parameter indices of `optimizer.state_dict()`."""
    no_decay, decay = [], []
    for name, p in named_params:
        if not p.requires_grad:
            continue
        (no_decay if (p.ndim == 1 or name.endswith(".bias")) else decay).append(name)
    return [({"weight_decay": 0.0}, no_decay), ({"weight_decay": float(weight_decay)}, decay)]


def finetune_group_plan(named_params, lr: float):
    """train_classification.py:5751-5768 over finetune.py:94-126: group "head" = the parameters of `lin_head` (frozen
    or not), group "backbone" = every other parameter in `model.parameters()` order; a model without a separate head
    puts everything in "head"."""
    head = [n for n, _ in named_params if n.startswith("lin_head.")]
    backbone = [n for n, _ in named_params if not n.startswith("lin_head.")]
    if not head and backbone:
        head, backbone = backbone, []
    plan = [({"lr": lr, "name": "head"}, head)]
    if backbone:
        plan.append(({"lr": lr, "name": "backbone"}, backbone))
    return plan


class PretrainEngine(TrainEngine):
    """MAE pretraining step: `step(imgs) -> loss` (device scalar, this rank's masked-patch mean: models_mae.py:213)."""

    def __init__(self, model, lr, weight_decay=0.05, betas=(0.9, 0.95), mask_ratio=0.75, **kw):
        super().__init__(model, lr, weight_decay, betas, **kw)
        self.mask_ratio = mask_ratio
        self._one = torch.ones((), dtype=torch.float32, device=self.device)

    def step(self, imgs, lr=None, noise=None, accum_iter: int = 1, micro_index: int = 0, graph: bool = False):
        """graph=True replays the whole step (noise draw, forward, backward, bucket all-reduces, AdamW) from a CUDA
        graph; the returned loss tensor is then a static buffer overwritten by the next call."""
        if graph:
            if accum_iter != 1:
                raise RuntimeError("graph replay does not accumulate gradients")
            if noise is not None:                             # caller-supplied noise becomes a second static input
                return self._replay(lambda x, z, lr: self.step(x, lr=lr, noise=z), [imgs, noise], lr)
            return self._replay(lambda x, lr: self.step(x, lr=lr), [imgs], lr)
        m = self.model
        if noise is None:
            noise = torch.rand(imgs.shape[0], m.patch_embed.num_patches, device=imgs.device)   # models_mae.py:132
        loss, pred_full, mask, ids_restore, sv = vit_core.mae_fwd(
            self.store, m._dims, imgs, noise, m._len_keep(self.mask_ratio), m.norm_pix_loss, True)
        dloss = self._one if accum_iter == 1 else self._one / accum_iter
        last = micro_index == accum_iter - 1
        self.backward_and_step(lambda sink: vit_core.mae_bwd(self.store, sink, m._dims, sv, dloss), lr=lr,
                               accumulate=micro_index > 0, do_step=last)
        return loss

    def gradients(self, imgs, noise, reduce: bool):
        """Forward + backward WITHOUT the optimiser update -> (loss, gradient arena clone).  reduce=True sends the
        gradients through the same bucketed all-reduce as `step` and returns the mean over ranks (what AdamW would have
        consumed); reduce=False returns this rank's own gradient.  Used by bench.py's `dp_check`."""
        m = self.model
        loss, _, _, _, sv = vit_core.mae_fwd(self.store, m._dims, imgs, noise, m._len_keep(self.mask_ratio), m.norm_pix_loss, True)
        self.backward_and_step(lambda sink: vit_core.mae_bwd(self.store, sink, m._dims, sv, self._one), do_step=reduce, apply=False)
        g = self.grads.clone()
        return loss, (g / self.world if reduce else g)


class ClassifierEngine(TrainEngine):
    """Fine-tune step (train_classification.py:4525-4546): logits -> BCE-with-logits on logit1-logit0 (2 classes,
    :3347-3374) or cross-entropy -> backward -> AdamW.  The loss over B x ncls scalars stays in torch."""

    def __init__(self, model, lr, weight_decay=0.0, betas=(0.9, 0.999), pos_weight=None, class_weights=None, **kw):
        # train_classification.py:5751-5768: plain torch.optim.AdamW over head + backbone groups with ONE weight decay for
        # every parameter (biases and LayerNorm included; default 0.0) - not timm's no-decay rule of the pretraining script
        kw.setdefault("no_decay_rule", lambda name, p: False)
        super().__init__(model, lr, weight_decay, betas, **kw)
        self.pos_weight = pos_weight
        self.class_weights = class_weights      # CrossEntropyLoss(weight=...) for > 2 classes (:6103)

    def _group_plan(self):
        return finetune_group_plan([(n, self.store.params[n]) for n in self.store.offsets], self.lr)

    def step(self, imgs, target, lr=None, graph: bool = False):
        if graph:
            return self._replay(lambda x, y, lr: self.step(x, y, lr=lr), [imgs, target], lr)
        m = self.model
        logits, sv = vit_core.classifier_fwd(self.store, m._dims, imgs, m.out_token == "spatial", True)
        lg = logits.detach().requires_grad_(True)
        if logits.shape[1] == 2:
            z = lg[:, 1] - lg[:, 0]
            loss = torch.nn.functional.binary_cross_entropy_with_logits(z, target.float(), pos_weight=self.pos_weight)
        else:
            loss = torch.nn.functional.cross_entropy(lg, target, weight=self.class_weights)
        (dlogits,) = torch.autograd.grad(loss, lg)
        self.backward_and_step(lambda sink: vit_core.classifier_bwd(self.store, sink, m._dims, sv, dlogits.contiguous()), lr=lr)
        return loss.detach()


class HostPipeline:
    """Feeds an engine from HOST batches without ever letting the GPU wait for the host:

      * batch i+1 is copied host -> device on a side stream (pinned source, double-buffered destination) while step i runs;
      * the step itself is the captured CUDA graph (`engine.step(..., graph=True)`) whenever the batch has the static shape;
      * the step's loss goes device -> host asynchronously into a pinned slot and is handed to the caller ONE step late,
        when its event has completed - the reference's `loss.item()` + `torch.cuda.synchronize()` + scalar all-reduce per
        iteration (engine_pretrain.py:55,74,81) without the pipeline bubble.  What the caller observes is unchanged: every
        loss value, in order; a non-finite loss is seen one step later than in the reference.

    `submit(batch..., lr=...)` enqueues one step and returns the losses that completed meanwhile (possibly none);
    `drain()` returns the rest.  `depth` steps may be in flight."""

    def __init__(self, engine: TrainEngine, graph: bool = True, depth: int = 2):
        self.e, self.graph, self.depth = engine, graph, max(2, depth)
        dev = engine.device
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.bufs: list = [None] * self.depth
        self.ready = [torch.cuda.Event() for _ in range(self.depth)]
        self.consumed = [torch.cuda.Event() for _ in range(self.depth)]
        self.loss_host = torch.zeros(self.depth, dtype=torch.float32).pin_memory()
        self.loss_done = [torch.cuda.Event() for _ in range(self.depth)]
        self.i = 0                       # steps submitted
        self.j = 0                       # losses handed out
        self.h2d_bytes = 0
        self._static_shapes = None

    def _harvest(self, block_until: int) -> list:
        """Losses of steps [j, ...) whose D2H has completed; blocks for every step < block_until."""
        out = []
        while self.j < self.i:
            ev = self.loss_done[self.j % self.depth]
            if self.j < block_until:
                ev.synchronize()
            elif not ev.query():
                break
            out.append(float(self.loss_host[self.j % self.depth]))
            self.j += 1
        return out

    def submit(self, *batch, lr=None, **step_kw) -> list:
        done = self._harvest(self.i - self.depth + 1)          # frees this slot's loss word and device buffers
        k = self.i % self.depth
        cur = torch.cuda.current_stream()
        on_host = not batch[0].is_cuda
        if on_host:
            shapes = [(tuple(t.shape), t.dtype) for t in batch]
            if self.bufs[k] is None or [(tuple(t.shape), t.dtype) for t in self.bufs[k]] != shapes:
                self.bufs[k] = [torch.empty(s, dtype=d, device=self.e.device) for s, d in shapes]
                # The caching allocator hands out blocks in the order of the ALLOCATING (compute) stream: this block may
                # still be in use by kernels already enqueued there (activations of an earlier step).  The copy stream
                # must not write it before they have run.
                self.copy_stream.wait_stream(cur)
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.consumed[k])
                for dst, src in zip(self.bufs[k], batch):
                    dst.copy_(src, non_blocking=True)
                    self.h2d_bytes += src.numel() * src.element_size()
                self.ready[k].record(self.copy_stream)
            cur.wait_event(self.ready[k])
            dev_batch = self.bufs[k]
        else:
            dev_batch = list(batch)
        shapes = [tuple(t.shape) for t in dev_batch]
        if self._static_shapes is None:
            self._static_shapes = shapes
        graph = self.graph and shapes == self._static_shapes and not step_kw.get("accum_iter", 1) > 1
        loss = self.e.step(*dev_batch, lr=lr, graph=graph, **step_kw)
        self.consumed[k].record(cur)
        self.loss_host[k:k + 1].copy_(loss.detach().reshape(1), non_blocking=True)
        self.loss_done[k].record(cur)
        self.i += 1
        return done

    def drain(self) -> list:
        return self._harvest(self.i)


def cosine_lr(it_epoch: float, lr: float, min_lr: float, warmup_epochs: float, epochs: float) -> float:
    """Per-iteration half-cosine with linear warm-up (reference util/lr_sched.py:9-21)."""
    if it_epoch < warmup_epochs:
        return lr * it_epoch / warmup_epochs
    return min_lr + (lr - min_lr) * 0.5 * (1. + math.cos(math.pi * (it_epoch - warmup_epochs) / (epochs - warmup_epochs)))


def finetune_cosine_factor(epoch: int, total_epochs: int, warmup_epochs: int = 0) -> float:
    """Per-EPOCH lr multiplier of the fine-tune script's `--scheduler cosine` (train_classification.py:3943-3958, a
    LambdaLR): linear warm-up (epoch+1)/warmup, then half cosine over the remaining epochs."""
    if warmup_epochs > 0 and epoch < warmup_epochs:
        return float(epoch + 1) / float(max(1, warmup_epochs))
    progress = (epoch - warmup_epochs) / float(max(1, total_epochs - warmup_epochs))
    progress = min(max(progress, 0.0), 1.0)
    return 0.5 * (1.0 + math.cos(math.pi * progress))


class PlateauLR:
    """`--scheduler plateau` (train_classification.py:3959-3970): torch's ReduceLROnPlateau(mode="max", factor, patience,
    min_lr) restated on a plain float, since the engine owns the learning rate (device scalar), not a torch optimizer.
    `lr = sched.step(metric)` once per epoch; pass the result to `engine.step(..., lr=lr)`."""

    def __init__(self, lr: float, factor: float = 0.5, patience: int = 2, min_lr: float = 1e-6, threshold: float = 1e-4, eps: float = 1e-8):
        self.lr, self.factor, self.patience, self.min_lr, self.threshold, self.eps = lr, factor, patience, min_lr, threshold, eps
        self.best, self.num_bad = -math.inf, 0

    def step(self, metric: float) -> float:
        if metric > self.best * (1.0 + self.threshold) if self.best > -math.inf else True:     # mode="max", threshold_mode="rel"
            self.best, self.num_bad = metric, 0
        else:
            self.num_bad += 1
        if self.num_bad > self.patience:
            new_lr = max(self.lr * self.factor, self.min_lr)
            if self.lr - new_lr > self.eps:
                self.lr = new_lr
            self.num_bad = 0
        return self.lr
