"""Data-parallel gradient exchange for the flat gradient arena: bucket plan + overlapped all-reduce.

The reference wraps the model in `torch.nn.parallel.DistributedDataParallel`
(/root/reference/src/ssl4polyp/models/mae/main_pretrain.py:212-214, classification/train_classification.py:5743-5750):
bucketed SUM all-reduce fired from autograd hooks, averaged by world size.  Here the backward is our own schedule, so
the exchange is driven by it directly: the arena is cut into contiguous buckets starting from its END (the arena is in
forward order, so its end is what backward produces first); when the last gradient of a bucket has been enqueued, an
event is recorded on the compute stream and NCCL all-reduces that slice on a side stream while backward carries on.
The division by world size is folded into the optimiser (`grad_scale`).  One process per GPU; no other collective is
on the data path.  The same class runs over gloo on CPU tensors (tests: world_size 2) with async work handles.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch
import torch.distributed as dist


@dataclass
class Bucket:
    start: int
    end: int
    names: set = field(default_factory=set)


def plan_buckets(offsets: dict, trainable: list, limit_elems: int, align: int = 256) -> list:
    """offsets: name -> (start, numel) in the arena; trainable: names that get a gradient.
    Returns buckets ordered as backward completes them (highest arena offsets first).  Buckets are contiguous and
    tile [first trainable start, last trainable end): frozen tensors in between ride along (their gradient is zero)."""
    names = sorted(trainable, key=lambda n: offsets[n][0], reverse=True)
    buckets: list[Bucket] = []
    cur = None
    for nme in names:
        a, cnt = offsets[nme]
        if cur is None:
            cur = Bucket(start=a, end=(a + cnt + align - 1) // align * align)
        cur.start = a
        cur.names.add(nme)
        if cur.end - cur.start >= limit_elems:
            buckets.append(cur)
            cur = None
    if cur is not None:
        buckets.append(cur)
    for i in range(1, len(buckets)):
        buckets[i].end = buckets[i - 1].start
    return buckets


def plan_buckets_tapered(offsets: dict, trainable: list, limit_elems: int, align: int = 256, tail_elems: int | None = None) -> list:
    """`plan_buckets`, then the LAST bucket (lowest offsets: what backward finishes last, i.e. the only all-reduce and
    optimiser slice with nothing left to hide behind) is re-cut into pieces of `tail_elems` (default limit/4), so the
    exposed tail of a step is one small bucket instead of one full-size one."""
    buckets = plan_buckets(offsets, trainable, limit_elems, align)
    tail_elems = tail_elems or max(limit_elems // 4, 1)
    if not buckets or tail_elems >= limit_elems:
        return buckets
    last = buckets.pop()
    names = sorted(last.names, key=lambda n: offsets[n][0], reverse=True)
    sub = plan_buckets({n: offsets[n] for n in names}, names, tail_elems, align)
    sub[0].end = last.end
    return buckets + sub


def broadcast_parameters(flat: torch.Tensor, group=None, src: int = 0) -> None:
    """Every rank starts from rank `src`'s parameter arena - what `DistributedDataParallel.__init__` does in the reference
    (main_pretrain.py:212-214): ranks are seeded `seed + rank` BEFORE the model is built (:146-148), so without this
    broadcast each replica would start from a different initialisation and the averaged gradients would be applied to N
    different models."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(flat, src=src, group=group)


def configure_overlap(n_sms: int) -> None:
    """Call BEFORE `init_process_group`: cap NCCL at `n_sms` CTAs (= SMs).  Together with `ssl4p_reserve_sms(n_sms)`
    (done by the engine) the all-reduce and our persistent kernels run on disjoint SM sets.  The gradient all-reduce
    needs little bandwidth (447 MB per ~25 ms step for ViT-B), so a handful of SMs is plenty."""
    import os
    if n_sms > 0:
        os.environ.setdefault("NCCL_MAX_CTAS", str(n_sms))
        os.environ.setdefault("NCCL_MAX_NCHANNELS", str(n_sms))


# timing experiments only (scripts/dp_sweep.sh): enqueue everything except the collective itself
_SKIP_COLLECTIVE = bool(int(__import__("os").environ.get("SSL4P_DP_SKIP_COLLECTIVE", "0")))


class BucketReducer:
    """Drives the per-bucket work of a step as the backward completes buckets:

      ready(name)  the kernels producing gradient `name` have been enqueued (on one of `producer_streams`);
      bucket i complete -> (world > 1) SUM all-reduce of its arena slice on the communication stream, ordered after the
                           producers by events;
      on_bucket(j, deps)   optional hook (the engine's per-bucket AdamW), called for bucket j ONE bucket late - when the
                           next bucket completes, or in finish() - with the events it has to wait for: bucket j's
                           all-reduce, plus the producers' position at the time of the call.  The lag is what makes an
                           in-place parameter update safe: every backward kernel that still READS the weights of bucket j
                           (the dgrad GEMMs of its layers) was enqueued before the next bucket could complete.
    With world == 1 there is no collective; readiness tracking and the hook work the same way."""

    def __init__(self, flat: torch.Tensor, buckets: list, group=None):
        self.flat = flat
        self.buckets = buckets
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.cuda = flat.is_cuda
        self.comm_stream = torch.cuda.Stream(device=flat.device) if (self.cuda and self.world > 1) else None
        self.symm_group = getattr(flat, "_ssl4p_symm_group", None)     # symmetric-memory arena: in-switch all-reduce (experiment)
        self.bucket_of = {n: i for i, b in enumerate(buckets) for n in b.names}
        self.launch_order: list[int] = []
        self.producer_streams = None            # streams the gradients are produced on (None: the current stream)
        self.on_bucket = None
        self.reset()

    def reset(self):
        self._remaining = [len(b.names) for b in self.buckets]
        self._seen = set()
        self._launched = set()
        self._handles = []
        self._done = {}                         # bucket -> [events] after which its (reduced) gradients are final
        self._pending_hook = None               # bucket whose on_bucket call is lagging
        self.launch_order = []

    def _producer_events(self):
        evs = []
        for st in (self.producer_streams or [torch.cuda.current_stream()]):
            ev = torch.cuda.Event()
            ev.record(st)
            evs.append(ev)
        return evs

    def ready(self, name: str) -> None:
        """The kernels producing gradient `name` have been enqueued."""
        if (self.world == 1 and self.on_bucket is None) or name in self._seen:
            return
        i = self.bucket_of.get(name)
        if i is None:
            return
        self._seen.add(name)
        self._remaining[i] -= 1
        if self._remaining[i] == 0:
            self._launch(i)

    def _launch(self, i: int) -> None:
        b = self.buckets[i]
        view = self.flat[b.start:b.end]
        self._launched.add(i)
        self.launch_order.append(i)
        if self.cuda:
            evs = self._producer_events()
            if self.world > 1:
                for ev in evs:
                    self.comm_stream.wait_event(ev)
                with torch.cuda.stream(self.comm_stream):
                    if not _SKIP_COLLECTIVE:
                        if self.symm_group is not None:
                            torch.ops.symm_mem.multimem_all_reduce_(view, "sum", self.symm_group)
                        else:
                            dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group)
                    done = torch.cuda.Event()
                    done.record(self.comm_stream)
                self._done[i] = [done]
            else:
                self._done[i] = evs
            if self.on_bucket is not None:
                if self._pending_hook is not None:
                    j = self._pending_hook
                    self.on_bucket(j, self._done[j] + evs)
                self._pending_hook = i
        elif self.world > 1:
            self._handles.append(dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group, async_op=True))

    def finish(self) -> None:
        """Reduce whatever has not been launched (parameters that received no gradient this step), run the lagging hook,
        then make the current stream wait for all bucket reductions."""
        if self.world == 1 and self.on_bucket is None:
            return
        for i in range(len(self.buckets)):
            if i not in self._launched:
                self._launch(i)
        if self.cuda:
            if self.on_bucket is not None and self._pending_hook is not None:
                j = self._pending_hook
                self.on_bucket(j, self._done[j] + self._producer_events())
                self._pending_hook = None
            if self.comm_stream is not None:
                torch.cuda.current_stream().wait_stream(self.comm_stream)
        else:
            for h in self._handles:
                h.wait()
            self._handles = []
