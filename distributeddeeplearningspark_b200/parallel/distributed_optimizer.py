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
    __slots__ = ('lo', 'hi', 'params', 'pending', 'done_event', 'launched', 'll_ok', 'sharded')

    def __init__(self, lo, hi, params):
        self.lo, self.hi, self.params = lo, hi, params
        self.pending = len(params)
        self.done_event = None
        self.launched = False
        self.ll_ok = False
        self.sharded = False


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
        # Fused data-parallel step (K12/K13/K17 of SURVEY.md §2.5): every bucket's allreduce AND optimizer update is ONE
        # kernel, launched the moment the bucket's last gradient lands —
        #   'll'  small buckets: flag-in-data one-shot allreduce, every rank applies the (replicated) update;
        #   'dp'  large buckets or buckets holding bf16 shadow weights: in-switch reduce-scatter, optimizer on this rank's
        #         1/N slice (slots sharded by rank), multicast all-gather of the new parameters.
        # step() then only waits for the buckets; the wrapped optimizer's own kernel never runs.
        self._fused = (self.size > 1 and not self._generic and env_int('DDLS_FUSED_OPT', 1) != 0
                       and hasattr(self.comm, 'allreduce_opt_') and getattr(self.inner, 'kind', '') in ('adam', 'adagrad', 'sgd')
                       and not getattr(self.inner, 'momentum', 0.0) and self._plan_fused())
        # One rank, several buckets, CUDA: there is nothing to reduce, but the same machinery runs the optimizer bucket by
        # bucket on the side stream as gradients become final ('local' mode) — the update of the big Dense bucket then
        # overlaps the convolution backward and only the small tail bucket's update is left at the end of the step.
        self._local = (self.size == 1 and not self._generic and self.device.type == 'cuda' and len(self.buckets) > 1
                       and env_int('DDLS_LOCAL_BUCKETS', 1) != 0 and self.backward_passes_per_step == 1
                       and hasattr(self.inner, 'can_update_range') and self.inner.can_update_range())
        if self._local:
            self._fused = True
            self.overlap = overlap
            self._comm_stream = torch.cuda.Stream(device=self.device)
            for p in self.flat.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        self._n_launched = 0
        if self._fused:
            self.inner._shard_gather = self.gather_slots
        # one bucket: nothing to overlap with, the kernel runs on the caller's stream
        if len(self.buckets) == 1:
            self.overlap = False

    def _plan_fused(self) -> bool:
        """Can every bucket run as a fused kernel?  Sets ``_dp_ok`` (sharded kernel usable) and ``_ll_ok`` per bucket."""
        comm = self.comm
        self._dp_ok = hasattr(comm, 'dp_step_') and comm.can_shard_step(self.inner) and self._bf16 is None
        # above this size the sharded kernel (in-switch reduce-scatter) beats the flag-in-data one (see ll_max_bytes)
        self._dp_min_bytes = env_int('DDLS_DP_STEP_MIN_BYTES', 0) or (comm.ll_max_bytes(self.size) if hasattr(comm, 'll_max_bytes') else 512 << 10)
        for b in self.buckets:
            b.ll_ok = comm.can_fuse_optimizer(self.flat.flat_grads[b.lo:b.hi])
            if not (b.ll_ok or self._dp_ok):
                return False
        return True

    def _bucket_mode(self, b) -> str:
        """'ll' or 'dp' for bucket ``b`` (decided at launch time: bf16 shadows are registered after construction);
        'local' on a single rank."""
        if self._local:
            return 'local'
        shadowed = any(beg < b.hi and beg + rows * cols > b.lo for beg, rows, cols, _, _ in getattr(self.inner, '_shadows', ()))
        if b.ll_ok and not shadowed and ((b.hi - b.lo) * 4 <= self._dp_min_bytes or not self._dp_ok):
            return 'll'
        if not self._dp_ok:
            raise RuntimeError('bucket holds bf16 shadow weights but the sharded step kernel is unavailable '
                               '(parameters, gradients and slots must live in the symmetric heap; bf16 wire unsupported)')
        return 'dp'

    def _fused_bucket(self, b) -> None:
        advance = self._n_launched == len(self.buckets)      # the last bucket of the step publishes step+1
        if self._local:
            self.inner.update_range(b.lo, min(b.hi, self.flat.numel), advance)
            return
        if self._bucket_mode(b) == 'll':
            n_opt = max(0, min(b.hi, self.flat.numel) - b.lo)
            self.comm.allreduce_opt_(self.flat.flat_grads[b.lo:b.hi], self.inner, self.op, wire_bf16=self._bf16 is not None,
                                     lo=b.lo, n_opt=n_opt, advance=advance)
        else:
            self.comm.dp_step_(self.inner, b.lo, b.hi, self.op, advance=advance)
            b.sharded = True

    def shard_bounds(self, b, rank: Optional[int] = None):
        """[lo, hi) flat element range of bucket ``b`` whose optimizer slots ``rank`` owns under the sharded kernel."""
        rank = self.comm.rank if rank is None else rank
        nvec = (b.hi - b.lo) // 4
        per = (nvec + self.size - 1) // self.size
        lo = min(nvec, rank * per); hi = min(nvec, lo + per)
        return b.lo + 4 * lo, min(b.lo + 4 * hi, self.flat.numel)

    def gather_slots(self) -> None:
        """Collective: make every rank's optimizer slots complete.  Buckets that ran through the sharded kernel keep
        m / v / accumulator up to date only on the owner of each slice; before a checkpoint or a state broadcast the
        slices are exchanged (non-owned parts are zeroed and the ranges summed over the ranks)."""
        if not self._fused or self.size == 1:
            return
        for b in self.buckets:
            if not getattr(b, 'sharded', False):
                continue
            lo, hi = self.shard_bounds(b)
            top = min(b.hi, self.flat.numel)
            for t in self.inner.slots().values():
                t[b.lo:lo].zero_(); t[hi:top].zero_()
                self.comm.allreduce_(t[b.lo:top], Sum)

    # ------------------------------------------------------------------ hooks
    def _on_grad(self, p: torch.Tensor) -> None:
        if (self.size == 1 and not self._local) or self._passes + 1 < self.backward_passes_per_step:
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
        self._n_launched += 1
        run = self._fused_bucket if self._fused else (lambda bb: self._reduce_range(bb.lo, bb.hi))
        if self.overlap:
            cur = torch.cuda.current_stream(self.device)
            ev = torch.cuda.Event()
            ev.record(cur)
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                run(b)
                b.done_event = torch.cuda.Event()
                b.done_event.record(self._comm_stream)
        else:
            run(b)

    # -------------------------------------------------------------- public API
    def synchronize(self) -> None:
        """Launch any bucket whose hook did not fire (unused parameters) and wait for all."""
        if self.size > 1 or self._local:
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
        self._n_launched = 0

    def step(self, closure=None):
        self._passes += 1
        if self._passes < self.backward_passes_per_step:
            return None
        self._passes = 0
        if not self._generic and not self.flat.grads_attached():
            self.flat.gather_stray_grads()
            for b in self.buckets:  # hooks saw foreign tensors; reduce everything now
                b.launched = False
        self.synchronize()
        if self._fused:                      # every bucket kernel has reduced AND applied its range
            return None if closure is None else closure()
        return self.inner.step() if closure is None else self.inner.step(closure)

    @property
    def metrics(self) -> Optional[torch.Tensor]:
        """Metric slots piggy-backed on the top gradient bucket (K20): write local values BEFORE ``backward()`` — on a
        multi-bucket model the bucket that carries the tail is launched from the gradient hook during backward, so a value
        written between ``backward()`` and ``step()`` (this step's loss, say) misses the collective or races with it; report
        it one step late instead.  Read the reduced values (same reduction as the gradients: average by default) after
        ``step()``.  ``None`` unless the wrapped optimizer was built with ``metric_slots=k``."""
        return getattr(self.flat, 'metrics', None)

    def zero_grad(self, set_to_none: bool = False) -> None:
        if self._generic:
            self.inner.zero_grad(set_to_none=set_to_none)
        else:
            self.flat.zero_grads()

    def state_dict(self):
        """Collective when slots are sharded (every rank must call it): gathers them first."""
        self.gather_slots()
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
