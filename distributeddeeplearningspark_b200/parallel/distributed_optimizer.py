"""``DistributedOptimizer`` — synchronous data parallelism with a static bucket plan.

Horovod-shaped wrapper required by BASELINE.json (the reference itself only has
dist-keras' asynchronous ADAG, see ``trainers``).  Because the gradient set of a
model is static, the bucket plan is computed once at construction and is
identical on every rank — no coordinator thread, no negotiation.
Gradients live in the flat (symmetric) gradient buffer, so a bucket is just a
contiguous range; the moment the last gradient of a bucket has been accumulated
its allreduce is launched on a side stream, overlapping the rest of backward.
``step()`` waits for the buckets and runs the fused optimizer kernel.
"""
from __future__ import annotations

from enum import Enum
from typing import Dict, Iterable, List, Optional, Tuple

import torch

from .. import runtime
from ..optim.flat import FlatSpace
from ..optim.keras_optimizers import _FlatOptimizer
from ..utils.misc import env_int
from .comm import Average, ReduceOp, Sum


class Compression(Enum):
    none = 'none'
    bf16 = 'bf16'


class _Bucket:
    __slots__ = ('lo', 'hi', 'params', 'pending', 'done_event', 'launched')

    def __init__(self, lo, hi, params):
        self.lo, self.hi, self.params = lo, hi, params
        self.pending = len(params)
        self.done_event = None
        self.launched = False


def plan_buckets(segments: List[Tuple[int, int]], bucket_elems: int,
                 total: Optional[int] = None) -> List[Tuple[int, int, List[int]]]:
    """Static plan: walk parameters in reverse registration order (≈ the order backward produces
    gradients) and cut a bucket whenever it reaches ``bucket_elems``.  Returns
    ``[(lo, hi, [param indices])]`` over the flat space, in launch order."""
    out = []
    cur: List[int] = []
    cur_elems = 0
    for i in reversed(range(len(segments))):
        cur.append(i)
        cur_elems += segments[i][1]
        if cur_elems >= bucket_elems:
            out.append(cur); cur = []; cur_elems = 0
    if cur:
        out.append(cur)
    plan = []
    for idxs in out:
        lo = min(segments[i][0] for i in idxs)
        hi_i = max(idxs)
        hi = segments[hi_i][0] + segments[hi_i][1]
        plan.append((lo, hi, idxs))
    # make ranges tile the flat buffer exactly (alignment padding belongs to the bucket below it)
    plan_sorted = sorted(range(len(plan)), key=lambda k: plan[k][0])
    for a, b in zip(plan_sorted[:-1], plan_sorted[1:]):
        lo, _, idxs = plan[a]
        plan[a] = (lo, plan[b][0], idxs)
    if total is not None and plan_sorted:
        # the top bucket also owns the alignment padding after the last parameter, so every range is a multiple of
        # the 16-byte vector the collectives move (an odd tail would bounce through the scratch region)
        top = plan_sorted[-1]
        lo, hi, idxs = plan[top]
        plan[top] = (lo, max(hi, total), idxs)
    return plan


class DistributedOptimizer:
    def __init__(self, optimizer, named_parameters: Optional[Iterable] = None,
                 compression: Compression = Compression.none, backward_passes_per_step: int = 1,
                 op: ReduceOp = Average, bucket_bytes: Optional[int] = None, overlap: bool = True):
        if not runtime.is_initialized():
            runtime.init()
        self.comm = runtime.communicator()
        self.size = runtime.size()
        self.inner = optimizer
        self.compression = Compression(compression) if not isinstance(compression, Compression) else compression
        self.backward_passes_per_step = int(backward_passes_per_step)
        self.op = op
        self._passes = 0
        self._handles: List = []
        if isinstance(optimizer, _FlatOptimizer):
            self.flat = optimizer.flat
            self._generic = False
        else:  # any torch.optim.Optimizer: gradients get their own flat space
            params = [p for g in optimizer.param_groups for p in g['params'] if p.requires_grad]
            self.flat = _GradOnlyFlat(params, self.comm)
            self._generic = True
        self.device = self.flat.device
        # the gradient buffer may carry a metric tail (FlatSpace.metric_slots) that rides in the top bucket
        self._grad_numel = getattr(self.flat, 'grad_numel', self.flat.numel)
        bucket_bytes = bucket_bytes or env_int('DDLS_BUCKET_BYTES', 2 << 20)
        self.buckets = [_Bucket(lo, hi, idxs) for lo, hi, idxs in
                        plan_buckets(self.flat.segments(), max(1, bucket_bytes // 4), self._grad_numel)]
        self._bucket_of: Dict[int, _Bucket] = {}
        for b in self.buckets:
            for i in b.params:
                self._bucket_of[id(self.flat.params[i])] = b
        self.overlap = overlap and self.device.type == 'cuda' and self.size > 1
        self._comm_stream = torch.cuda.Stream(device=self.device) if self.overlap else None
        self._hooks = []
        if self.size > 1 and not self._generic:
            for p in self.flat.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        if isinstance(self.inner, _FlatOptimizer):
            self.inner.grad_scale = 1.0 / self.backward_passes_per_step
        self._bf16 = None
        if self.compression is Compression.bf16 and self.size > 1:
            self._bf16 = self.comm.alloc(self._grad_numel, torch.bfloat16)
        # Single-bucket models whose gradients fit the flag-in-data allreduce get allreduce + optimizer in ONE kernel
        # (K12/K13 of SURVEY.md §2.5): the hook then has nothing to launch, step() issues the fused kernel.
        self._fused = (self.size > 1 and not self._generic and len(self.buckets) == 1
                       and self.backward_passes_per_step == 1 and env_int('DDLS_FUSED_OPT', 1) != 0
                       and hasattr(self.comm, 'allreduce_opt_') and getattr(self.inner, 'kind', '') in ('adam', 'adagrad', 'sgd')
                       and not getattr(self.inner, 'momentum', 0.0)
                       and self.buckets[0].lo == 0 and self.buckets[0].hi == self._grad_numel
                       and self.comm.can_fuse_optimizer(self.flat.flat_grads))

    # ------------------------------------------------------------------ hooks
    def _on_grad(self, p: torch.Tensor) -> None:
        if self.size == 1 or self._passes + 1 < self.backward_passes_per_step:
            return  # nothing to reduce / keep accumulating locally
        b = self._bucket_of.get(id(p))
        if b is None or b.launched:
            return
        b.pending -= 1
        if b.pending == 0:
            self._launch(b)

    def _reduce_range(self, lo: int, hi: int) -> None:
        g = self.flat.flat_grads[lo:hi]
        if self._bf16 is not None:
            w = self._bf16[lo:hi]
            fused = getattr(self.comm, 'allreduce_bf16_wire_', None)
            if fused is not None:
                fused(g, w, self.op)
            else:
                w.copy_(g)
                self.comm.allreduce_(w, self.op)
                g.copy_(w)
        else:
            self.comm.allreduce_(g, self.op)

    def _launch(self, b: _Bucket) -> None:
        b.launched = True
        if self._fused and not getattr(self.inner, '_nspecs', 0):
            return                                   # reduced and applied by the fused kernel in step()
        if self.overlap:
            cur = torch.cuda.current_stream(self.device)
            ev = torch.cuda.Event()
            ev.record(cur)
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                self._reduce_range(b.lo, b.hi)
                b.done_event = torch.cuda.Event()
                b.done_event.record(self._comm_stream)
        else:
            self._reduce_range(b.lo, b.hi)

    # -------------------------------------------------------------- public API
    def synchronize(self) -> None:
        """Launch any bucket whose hook did not fire (unused parameters) and wait for all."""
        if self.size > 1:
            if self._generic:
                self.flat.pack()
            for b in self.buckets:
                if not b.launched:
                    self._launch(b)
            if self.overlap:
                cur = torch.cuda.current_stream(self.device)
                for b in self.buckets:
                    if b.done_event is not None:
                        cur.wait_event(b.done_event)
            if self._generic:
                self.flat.unpack()
        for b in self.buckets:
            b.pending = len(b.params); b.launched = False; b.done_event = None

    def step(self, closure=None):
        self._passes += 1
        if self._passes < self.backward_passes_per_step:
            return None
        self._passes = 0
        if not self._generic and not self.flat.grads_attached():
            self.flat.gather_stray_grads()
            for b in self.buckets:  # hooks saw foreign tensors; reduce everything now
                b.launched = False
        if self._fused and closure is None and not getattr(self.inner, '_nspecs', 0):
            for b in self.buckets:
                b.pending = len(b.params); b.launched = False; b.done_event = None
            self.comm.allreduce_opt_(self.flat.flat_grads, self.inner, self.op, wire_bf16=self._bf16 is not None)
            return None
        self.synchronize()
        return self.inner.step() if closure is None else self.inner.step(closure)

    @property
    def metrics(self) -> Optional[torch.Tensor]:
        """Metric slots piggy-backed on the top gradient bucket (K20): write local values before ``step()``, read the
        reduced values (same reduction as the gradients: average by default) after it.  ``None`` unless the wrapped
        optimizer was built with ``metric_slots=k``."""
        return getattr(self.flat, 'metrics', None)

    def zero_grad(self, set_to_none: bool = False) -> None:
        if self._generic:
            self.inner.zero_grad(set_to_none=set_to_none)
        else:
            self.flat.zero_grads()

    def state_dict(self):
        return self.inner.state_dict()

    def load_state_dict(self, sd):
        return self.inner.load_state_dict(sd)

    @property
    def param_groups(self):
        return self.inner.param_groups

    def __getattr__(self, name):
        return getattr(self.__dict__['inner'], name)


class _GradOnlyFlat:
    """Flat gradient staging for a generic ``torch.optim.Optimizer`` (pack → allreduce → unpack)."""

    def __init__(self, params, comm):
        self.params = list(params)
        self.device = self.params[0].device
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + 3) // 4 * 4
        self.numel = off
        self.flat_grads = comm.alloc(off, torch.float32) if self.device.type == 'cuda' \
            else torch.zeros(off, dtype=torch.float32, device=self.device)

    def segments(self):
        return [(o, p.numel()) for p, o in zip(self.params, self.offsets)]

    def grads_attached(self):
        return True

    def pack(self):
        for p, o in zip(self.params, self.offsets):
            v = self.flat_grads[o:o + p.numel()]
            if p.grad is None:
                v.zero_()
            else:
                v.copy_(p.grad.reshape(-1))

    def unpack(self):
        for p, o in zip(self.params, self.offsets):
            g = self.flat_grads[o:o + p.numel()].view(p.shape)
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)

    def zero_grads(self):
        self.flat_grads.zero_()
