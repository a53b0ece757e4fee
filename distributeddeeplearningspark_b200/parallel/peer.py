"""Peer-memory communicator: symmetric VMM heap + hand-written NVLink/NVSwitch collectives.

``PeerMemoryComm`` is the product path for gradient allreduce / parameter broadcast on GPUs
(``csrc/symm.cpp`` + ``csrc/comm.cu``).  Set-up per rank:

1. ``cuMemCreate`` one physical heap, export a POSIX fd, map it locally;
2. swap fds with every peer over Unix sockets (``SCM_RIGHTS`` — the ranks are separate processes
   spawned by the Spark-style launcher, so there is no common CUDA parent to inherit from);
3. map every peer's heap; if the fabric supports it, create/import one multicast object, add all
   devices, bind every heap and map the multicast view (``multimem.*``);
4. the first 4 KB of each heap hold the flag words and the epoch counter.

Tensors handed out by ``alloc`` live in the heap at the same offset on every rank (allocation calls
must be made in the same order on all ranks), so collectives address peers as ``base[r] + offset``.
``torch.distributed`` is used only to bootstrap (socket paths, barriers, python objects).

``EmulatedPeerGroup`` runs the same kernels for ``world`` emulated ranks on ONE GPU inside a single
cooperative launch (blockIdx.y = rank) — used by the single-GPU tests; it never launches kernels that
wait on each other separately.
"""
from __future__ import annotations

import array
import ctypes
import os
import socket
import tempfile
import time
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from .. import _native as N
from ..utils.misc import env_int, env_str
from .comm import Average, Communicator, ReduceOp, Sum

MAX_RANKS = 8
CONTROL_BYTES = 8192
FLAGS_OFF = 0
SEQ_OFF = 4096
ALIGN = 256

AR_ONESHOT, AR_TWOSHOT, AR_NVLS, AR_ONESHOT_LL = 0, 1, 2, 3
WIRE_BF16, LL_MC_PUSH = 1, 2                # wire_flags of ddls_comm_allreduce / ddls_comm_allreduce_opt
LL_MAX_BYTES_DEFAULT = 512 << 10     # flag-in-data one-shot up to this message size (NYISO LSTM: 267 KB of gradients)


class CommDesc(ctypes.Structure):
    _fields_ = [('base', ctypes.c_void_p * MAX_RANKS), ('mc_base', ctypes.c_void_p), ('world', ctypes.c_int),
                ('rank', ctypes.c_int), ('emulate', ctypes.c_int), ('timeout_ms', ctypes.c_uint),
                ('flags_off', ctypes.c_uint64), ('seq_off', ctypes.c_uint64)]


P, I, U64, F = N.c_void_p, N.c_int, ctypes.c_uint64, N.c_float
DP = ctypes.POINTER(CommDesc)
N.register('ddls_multicast_supported', [I])
N.register('ddls_symm_granularity', [I, I, ctypes.POINTER(ctypes.c_size_t)])
N.register('ddls_symm_create', [I, N.c_size_t, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(P), ctypes.POINTER(I)])
N.register('ddls_symm_import', [I, I, N.c_size_t, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(P)])
N.register('ddls_symm_unmap', [P, N.c_size_t, ctypes.c_ulonglong])
N.register('ddls_mc_create', [I, N.c_size_t, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(I)])
N.register('ddls_mc_import', [I, ctypes.POINTER(ctypes.c_ulonglong)])
N.register('ddls_mc_add_device', [ctypes.c_ulonglong, I])
N.register('ddls_mc_bind_and_map', [ctypes.c_ulonglong, ctypes.c_ulonglong, N.c_size_t, I, ctypes.POINTER(P)])
N.register('ddls_comm_allreduce', [DP, I, U64, U64, U64, U64, U64, F, I, I, P])
N.register('ddls_comm_allreduce_opt', [DP, U64, U64, U64, U64, U64, F, I, I, P, P, P, P, F, F, F, F, I, I, P])
N.register('ddls_comm_dp_step', [DP, U64, U64, U64, U64, U64, U64, U64, F, I, P, F, F, F, F, I, I, P, I, I, P])
N.register('ddls_comm_broadcast', [DP, U64, U64, I, I, I, P])
N.register('ddls_comm_allgather', [DP, U64, U64, U64, I, P])
N.register('ddls_comm_barrier', [DP, P])
N.register('ddls_comm_adag_commit', [DP, P, P, U64, U64, U64, F, I, P, P])


class _RawBuffer:
    """Exposes a raw device address range to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.ptr, self.nbytes, self.owner = ptr, nbytes, owner
        self.__cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (ptr, False), 'version': 3,
                                         'strides': None}


def _tensor_from_ptr(ptr: int, nbytes: int, dtype: torch.dtype, device: torch.device, owner=None) -> torch.Tensor:
    raw = torch.as_tensor(_RawBuffer(ptr, nbytes, owner), device=device)
    return raw.view(dtype)


class _HeapBase:
    """Bump allocator + collective launchers shared by the real and the emulated communicator."""

    def _init_heap(self, heap_bytes: int):
        self.heap_bytes = heap_bytes
        self._top = CONTROL_BYTES
        self._stage_off = 0
        self._stage_elems = 0
        self._stage_bytes = 0
        self._ll_off = self._ll_elems = self._ll_bytes = 0

    def _init_ll_stage(self, world: int) -> None:
        """Reserve the flag-in-data staging region.  The caller zeroes it (and, across processes, synchronises)
        BEFORE any collective can run: a packet's flag word must never hold anything but an epoch number."""
        self._ll_elems = env_int('DDLS_LL_MAX_BYTES', LL_MAX_BYTES_DEFAULT) // 4
        self._ll_bytes = 2 * world * self._ll_elems * 8
        self._ll_off = self._alloc_off(self._ll_bytes)

    def _alloc_off(self, nbytes: int) -> int:
        off = (self._top + ALIGN - 1) // ALIGN * ALIGN
        if off + nbytes > self.heap_bytes:
            raise MemoryError(f'symmetric heap exhausted: need {nbytes} B at {off} of {self.heap_bytes} '
                              f'(raise DDLS_HEAP_MB)')
        self._top = off + nbytes
        return off

    def _ensure_stage(self, n_elems: int, world: int, algo: int = AR_ONESHOT):
        """(offset, bytes) of the staging region for push collectives: two fixed parity halves of [world][n] slots.
        The flag-in-data algorithm has its own zero-initialised region (8 bytes per element), never shared with
        the others, so a packet's flag word can only ever hold an epoch number."""
        if algo == AR_ONESHOT_LL:
            if n_elems > self._ll_elems:
                raise ValueError(f'flag-in-data allreduce is limited to {self._ll_elems * 4} bytes (DDLS_LL_MAX_BYTES)')
            return self._ll_off, self._ll_bytes
        if n_elems > self._stage_elems:
            want = max(n_elems, 1 << 20)
            if self._stage_bytes and self._stage_off + self._stage_bytes == self._top:
                self._top = self._stage_off          # the old region is the last allocation: grow it in place
            # (otherwise the bump allocator cannot give the old region back; size DDLS_HEAP_MB for the largest message)
            self._stage_bytes = 2 * world * want * 4
            self._stage_off = self._alloc_off(self._stage_bytes)
            self._stage_elems = want
        return self._stage_off, self._stage_bytes

    def ll_max_bytes(self, world: int) -> int:
        """Largest message the flag-in-data one-shot still wins at ``world`` ranks.  Every rank receives (world-1) packets
        per element, the in-switch reduction one: measured crossovers (profiles/sweep_n{2,4,8}.json, graph-timed) are
        above 256 KB at 2 and at 4 GPUs (NYISO LSTM step at 4 GPUs: 89.1 us with the flag-in-data kernel, 93.7 us with the
        sharded one) and ~64-100 KB at 8, i.e. about 800 KB of inbound payload."""
        return min(self._ll_elems * 4, env_int('DDLS_LL_INBOUND_BYTES', 800 << 10) // max(1, world - 1))

    def select_algo(self, nbytes: int, symmetric_out: bool, have_mc: bool) -> int:
        """Measured selection (profiles/sweep_n{2,4,8}.json): flag-in-data one-shot while its (world-1)-fold inbound
        traffic is small, then the in-switch reduction; without multicast, fenced one-shot up to 1 MB and two-shot above.
        At two ranks the switch has nothing to save (each rank's own data loops through it), so two-shot takes over."""
        world = getattr(self, 'size', None) or getattr(self, 'world', 2)
        one_max = env_int('DDLS_ONESHOT_MAX_BYTES', 1 << 20)
        forced = env_str('DDLS_ALLREDUCE_ALGO', '')
        if forced:
            return {'oneshot': AR_ONESHOT, 'twoshot': AR_TWOSHOT, 'nvls': AR_NVLS, 'll': AR_ONESHOT_LL}[forced]
        if nbytes <= (self.ll_max_bytes(world) if have_mc else self._ll_elems * 4):
            return AR_ONESHOT_LL
        if not symmetric_out:
            return AR_ONESHOT
        if have_mc and not (world == 2 and nbytes >= (2 << 20)):
            return AR_NVLS
        return AR_ONESHOT if nbytes <= one_max else AR_TWOSHOT


class PeerMemoryComm(Communicator, _HeapBase):
    name = 'peer-memory'

    def __init__(self, rank: int, size: int, device: torch.device, local_rank: int, local_size: int):
        Communicator.__init__(self, rank, size, device)
        if size > MAX_RANKS:
            raise RuntimeError(f'peer-memory communicator supports up to {MAX_RANKS} ranks on one NVSwitch domain')
        if local_size != size:
            raise RuntimeError('peer-memory communicator needs all ranks on one node')
        self.lib = N.lib()
        self.dev_index = device.index or 0
        for peer in range(torch.cuda.device_count()):
            if peer != self.dev_index and not torch.cuda.can_device_access_peer(self.dev_index, peer):
                raise RuntimeError(f'GPU {self.dev_index} cannot access GPU {peer} peer-to-peer')
        torch.cuda.set_device(device)
        gran = ctypes.c_size_t(0)
        N.call('ddls_symm_granularity', self.dev_index, size, ctypes.byref(gran))
        want = env_int('DDLS_HEAP_MB', 768) << 20
        heap_bytes = (want + gran.value - 1) // gran.value * gran.value
        self._init_heap(heap_bytes)
        # 1. my heap
        h = ctypes.c_ulonglong(0); p = P(0); fd = I(-1)
        N.call('ddls_symm_create', self.dev_index, heap_bytes, ctypes.byref(h), ctypes.byref(p), ctypes.byref(fd))
        self._handles: List[int] = [0] * size
        self._bases: List[int] = [0] * size
        self._handles[rank], self._bases[rank] = h.value, p.value
        # 2. swap fds
        peer_fds = self._exchange_fds(fd.value, tag='heap')
        os.close(fd.value)
        # 3. map peers
        for r, pfd in peer_fds.items():
            hh = ctypes.c_ulonglong(0); pp = P(0)
            N.call('ddls_symm_import', self.dev_index, pfd, heap_bytes, ctypes.byref(hh), ctypes.byref(pp))
            os.close(pfd)
            self._handles[r], self._bases[r] = hh.value, pp.value
        # zero my control region before anyone signals
        self._heap_u8 = _tensor_from_ptr(self._bases[rank], heap_bytes, torch.uint8, device, owner=self)
        self._heap_u8[:CONTROL_BYTES].zero_()
        torch.cuda.synchronize(device)
        self._pg_barrier()
        # 4. multicast
        self.mc_base = 0
        self._mc_handle = 0
        if env_int('DDLS_DISABLE_MULTICAST', 0) == 0 and self.lib.ddls_multicast_supported(self.dev_index):
            try:
                self._setup_multicast(heap_bytes)
            except Exception as e:  # noqa: BLE001 - multicast is an optimisation, P2P paths remain
                from ..utils.logging import get_logger
                get_logger().warning('multicast set-up failed (%s); NVLS paths disabled', e)
                self.mc_base = 0
        self.desc = CommDesc()
        for r in range(size):
            self.desc.base[r] = self._bases[r]
        self.desc.mc_base = self.mc_base or None
        self.desc.world, self.desc.rank, self.desc.emulate = size, rank, 0
        self.desc.timeout_ms = env_int('DDLS_COMM_TIMEOUT_MS', 20000)
        self.desc.flags_off, self.desc.seq_off = FLAGS_OFF, SEQ_OFF
        self._scratch_off = self._alloc_off(64 << 20)   # staging for tensors that are not heap-resident
        self._scratch_bytes = 64 << 20
        self._init_ll_stage(size)
        # CTAs of the bandwidth algorithms (0 = by size, up to 64): fewer leave more SMs to the backward kernels the
        # gradient allreduce overlaps with
        self._ar_blocks = env_int('DDLS_AR_BLOCKS', 0)
        # flag-in-data packets pushed with one store through the multicast mapping instead of one store per peer
        self._ll_mc_push = env_int('DDLS_LL_MC_PUSH', 1) != 0
        self._heap_u8[self._ll_off:self._ll_off + self._ll_bytes].zero_()
        torch.cuda.synchronize(device)
        self._pg_barrier()

    def _pg_barrier(self) -> None:
        """Host-side barrier of the bootstrap process group (gloo by default: no NCCL on the product path)."""
        if dist.get_backend() == 'nccl':
            dist.barrier(device_ids=[self.dev_index])
        else:
            dist.barrier()

    # ------------------------------------------------------------------ fd exchange
    def _exchange_fds(self, my_fd: int, tag: str, only_from: Optional[int] = None) -> Dict[int, int]:
        """All-to-all (or one-to-all from ``only_from``) transfer of a file descriptor."""
        tmpdir = tempfile.gettempdir()
        token = [None]
        if self.rank == 0:
            token[0] = f'{os.getpid()}_{int(time.time() * 1e6) & 0xffffffff:x}'
        dist.broadcast_object_list(token, src=0)
        path = lambda r: os.path.join(tmpdir, f'ddls_{token[0]}_{tag}_{r}.sock')
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        try:
            os.unlink(path(self.rank))
        except FileNotFoundError:
            pass
        srv.bind(path(self.rank))
        srv.listen(self.size)
        self._pg_barrier()
        received: Dict[int, int] = {}
        senders = [only_from] if only_from is not None else list(range(self.size))
        # send my fd to every peer that expects it
        if only_from is None or only_from == self.rank:
            for r in range(self.size):
                if r == self.rank:
                    continue
                c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                c.connect(path(r))
                socket.send_fds(c, [self.rank.to_bytes(4, 'little')], [my_fd])
                c.close()
        expect = sum(1 for s in senders if s != self.rank)
        srv.settimeout(60.0)
        for _ in range(expect):
            conn, _ = srv.accept()
            msg, fds, _, _ = socket.recv_fds(conn, 4, 1)
            received[int.from_bytes(msg, 'little')] = fds[0]
            conn.close()
        self._pg_barrier()
        srv.close()
        try:
            os.unlink(path(self.rank))
        except FileNotFoundError:
            pass
        return received

    def _setup_multicast(self, heap_bytes: int) -> None:
        mch = ctypes.c_ulonglong(0)
        if self.rank == 0:
            fd = I(-1)
            N.call('ddls_mc_create', self.size, heap_bytes, ctypes.byref(mch), ctypes.byref(fd))
            self._exchange_fds(fd.value, tag='mc', only_from=0)
            os.close(fd.value)
        else:
            got = self._exchange_fds(-1, tag='mc', only_from=0)
            N.call('ddls_mc_import', got[0], ctypes.byref(mch))
            os.close(got[0])
        N.call('ddls_mc_add_device', mch.value, self.dev_index)
        self._pg_barrier()       # every device joined before anyone binds
        p = P(0)
        N.call('ddls_mc_bind_and_map', mch.value, self._handles[self.rank], heap_bytes, self.dev_index, ctypes.byref(p))
        self._pg_barrier()
        self._mc_handle, self.mc_base = mch.value, p.value

    # ------------------------------------------------------------------ memory
    def alloc(self, numel: int, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        off = self._alloc_off(max(nbytes, 16))
        t = self._heap_u8[off:off + nbytes].view(dtype)
        t.zero_()
        return t

    def offset_of(self, t: torch.Tensor) -> Optional[int]:
        a = t.data_ptr()
        base = self._bases[self.rank]
        if base <= a and a + t.numel() * t.element_size() <= base + self.heap_bytes:
            return a - base
        return None

    # ------------------------------------------------------------------ collectives
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def allreduce_(self, t: torch.Tensor, op: ReduceOp = Average, prescale: float = 1.0, wire_bf16: bool = False,
                   algo: Optional[int] = None) -> torch.Tensor:
        if t.dtype != torch.float32:
            tmp = t.float()
            self.allreduce_(tmp, op, prescale, wire_bf16, algo)
            t.copy_(tmp)
            return t
        n = t.numel()
        scale = prescale * ((1.0 / self.size) if op is Average else 1.0)
        off = self.offset_of(t) if (n % 4 == 0 and t.is_contiguous()) else None
        if off is not None and off % 16 == 0:
            a = algo if algo is not None else self.select_algo(n * 4, True, bool(self.mc_base))
            if a == AR_NVLS and wire_bf16 and n % 8 != 0:
                a = AR_TWOSHOT                            # the bf16 NVLS form moves 8-element vectors
            if a == AR_NVLS:                              # fp32 NVLS is in place; the bf16 wire needs n bf16 of staging
                stage, sbytes = self._ensure_stage((n + 2 * self.size - 1) // (2 * self.size), self.size, a) if wire_bf16 else (0, 0)
            else:
                stage, sbytes = self._ensure_stage(n, self.size, a)
            blocks = self._ar_blocks if a in (AR_NVLS, AR_TWOSHOT) else 0
            N.call('ddls_comm_allreduce', ctypes.byref(self.desc), a, off, off, stage, sbytes, n, scale,
                   self._wire_flags(wire_bf16), blocks, self._stream())
            torch.autograd.graph.increment_version(t)     # raw-pointer kernels bypass torch's version counter
            return t
        # not heap-resident: bounce through the scratch region (chunked)
        flat = t.reshape(-1)
        cap = self._scratch_bytes // 4
        scratch = self._heap_u8[self._scratch_off:self._scratch_off + self._scratch_bytes].view(torch.float32)
        for lo in range(0, n, cap):
            m = min(cap, n - lo)
            m4 = (m + 3) // 4 * 4
            scratch[:m].copy_(flat[lo:lo + m])
            if m4 > m:
                scratch[m:m4].zero_()
            a = algo if algo is not None else self.select_algo(m4 * 4, True, bool(self.mc_base))
            stage, sbytes = self._ensure_stage(m4, self.size, a) if a != AR_NVLS else (0, 0)
            N.call('ddls_comm_allreduce', ctypes.byref(self.desc), a, self._scratch_off, self._scratch_off, stage, sbytes,
                   m4, scale, self._wire_flags(wire_bf16), 0, self._stream())
            flat[lo:lo + m].copy_(scratch[:m])
        return t

    def can_fuse_optimizer(self, g: torch.Tensor) -> bool:
        """Is ``g`` (a heap-resident fp32 gradient range) small enough for the fused allreduce + optimizer kernel?"""
        n = g.numel()
        return (g.dtype == torch.float32 and n % 4 == 0 and g.is_contiguous() and n <= self._ll_elems
                and self.offset_of(g) is not None and self.offset_of(g) % 16 == 0)

    def _wire_flags(self, wire_bf16: bool) -> int:
        return (WIRE_BF16 if wire_bf16 else 0) | (LL_MC_PUSH if (self._ll_mc_push and self.mc_base) else 0)

    def allreduce_opt_(self, g: torch.Tensor, opt, op: ReduceOp = Average, wire_bf16: bool = False,
                       lo: int = 0, n_opt: Optional[int] = None, advance: bool = True) -> None:
        """Allreduce the gradient range ``g`` (= flat_grads[lo:lo+len(g)]) and apply ``opt``'s update (Keras Adam /
        Adagrad / SGD over the matching range of the flat parameter space) in ONE kernel: the thread that has summed a
        gradient vector updates its parameters.  ``n_opt`` elements of the range belong to the optimizer (the rest is a
        metric tail that is only reduced); ``advance``: publish step+1 (the last bucket of a step)."""
        kind = {'adam': 1, 'adagrad': 2, 'sgd': 3}[opt.kind]
        scale = opt.grad_scale * ((1.0 / self.size) if op is Average else 1.0)
        s1 = opt.m if kind == 1 else (opt.acc if kind == 2 else None)
        s2 = opt.v if kind == 1 else None
        n_opt = (opt.flat.numel - lo) if n_opt is None else n_opt
        sl = lambda t: None if t is None else t[lo:]
        N.call('ddls_comm_allreduce_opt', ctypes.byref(self.desc), self.offset_of(g), self._ll_off, self._ll_bytes,
               g.numel(), n_opt, scale, self._wire_flags(wire_bf16), kind, opt.flat.flat_params[lo:].data_ptr(),
               N.ptr(sl(s1)), N.ptr(sl(s2)), opt._step_buf.data_ptr(), opt.lr, getattr(opt, 'beta_1', 0.0),
               getattr(opt, 'beta_2', 0.0), getattr(opt, 'epsilon', 0.0), int(opt.zero_grads_in_step), int(advance),
               self._stream())
        torch.autograd.graph.increment_version(g)

    def can_shard_step(self, opt) -> bool:
        """Do ``opt``'s parameters, gradients and slots all live in the symmetric heap (what ``dp_step_`` needs)?"""
        ts = [opt.flat.flat_params, opt.flat.flat_grads] + list(opt.slots().values())
        return all(t is not None and t.dtype == torch.float32 and self.offset_of(t) is not None
                   and self.offset_of(t) % 16 == 0 for t in ts)

    def dp_step_(self, opt, lo: int, hi: int, op: ReduceOp = Average, advance: bool = True, blocks: int = 0) -> None:
        """Sharded data-parallel step of the gradient bucket ``flat_grads[lo:hi]`` (``ddls_comm_dp_step``): in-switch
        reduce-scatter, Keras update of this rank's 1/size slice, multicast all-gather of the new parameters, local
        gradient clearing and bf16 shadow refresh — one kernel.  The slots of the bucket are valid only on the owner
        of each slice afterwards (``shard_bounds``)."""
        kind = {'adam': 1, 'adagrad': 2, 'sgd': 3}[opt.kind]
        f = opt.flat
        scale = opt.grad_scale * ((1.0 / self.size) if op is Average else 1.0)
        s1 = opt.m if kind == 1 else (opt.acc if kind == 2 else None)
        s2 = opt.v if kind == 1 else None
        n_opt = max(0, min(hi, f.numel) - lo)
        off = lambda t: 0 if t is None else self.offset_of(t) + 4 * lo
        N.call('ddls_comm_dp_step', ctypes.byref(self.desc), off(f.flat_grads), off(f.flat_params), off(s1), off(s2),
               hi - lo, n_opt, lo, scale, kind, opt._step_buf.data_ptr(), opt.lr, getattr(opt, 'beta_1', 0.0),
               getattr(opt, 'beta_2', 0.0), getattr(opt, 'epsilon', 0.0), int(opt.zero_grads_in_step), int(advance),
               opt._specs, opt._nspecs, blocks or self._ar_blocks, self._stream())
        torch.autograd.graph.increment_version(f.flat_grads)

    def allreduce_bf16_wire_(self, g: torch.Tensor, _w: torch.Tensor, op: ReduceOp = Average) -> torch.Tensor:
        """fp32 gradients, bf16 on the wire, fp32 accumulate — the cast is fused into the collective."""
        return self.allreduce_(g, op, wire_bf16=True)

    def broadcast_(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        nbytes = t.numel() * t.element_size()
        off = self.offset_of(t) if t.is_contiguous() else None
        if off is not None and off % 16 == 0 and nbytes % 16 == 0:
            N.call('ddls_comm_broadcast', ctypes.byref(self.desc), off, nbytes, root, 1, 0, self._stream())
            torch.autograd.graph.increment_version(t)
            return t
        flat = t.contiguous().view(-1).view(torch.uint8) if t.is_contiguous() else t.contiguous().view(torch.uint8).view(-1)
        scratch = self._heap_u8[self._scratch_off:self._scratch_off + self._scratch_bytes]
        for lo in range(0, nbytes, self._scratch_bytes):
            m = min(self._scratch_bytes, nbytes - lo)
            m16 = (m + 15) // 16 * 16
            if self.rank == root:
                scratch[:m].copy_(flat[lo:lo + m])
            N.call('ddls_comm_broadcast', ctypes.byref(self.desc), self._scratch_off, m16, root, 1, 0, self._stream())
            if self.rank != root:
                flat[lo:lo + m].copy_(scratch[:m])
        if not t.is_contiguous():
            t.copy_(flat.view(t.dtype).view(t.shape))
        return t

    def allgather(self, t: torch.Tensor) -> torch.Tensor:
        """[size, *t.shape]: every rank's tensor, through the scratch region of the heap (chunked when it does not fit)."""
        t = t.contiguous()
        nbytes = t.numel() * t.element_size()
        src = t.view(-1).view(torch.uint8)
        out = torch.empty((self.size, nbytes), dtype=torch.uint8, device=t.device)
        cap = (self._scratch_bytes // (self.size + 1)) // 16 * 16
        scratch = self._heap_u8[self._scratch_off:self._scratch_off + self._scratch_bytes]
        for lo in range(0, max(nbytes, 1), cap):
            m = min(cap, nbytes - lo)
            chunk = (m + 15) // 16 * 16
            scratch[:m].copy_(src[lo:lo + m])
            N.call('ddls_comm_allgather', ctypes.byref(self.desc), self._scratch_off, self._scratch_off + chunk, chunk, 0,
                   self._stream())
            # (the next chunk's kernel starts with a flag round: no peer pushes before every rank has copied this one out)
            out[:, lo:lo + m].copy_(scratch[chunk:chunk + chunk * self.size].view(self.size, chunk)[:, :m])
        return out.view(t.dtype).reshape((self.size,) + tuple(t.shape))

    def barrier(self) -> None:
        N.call('ddls_comm_barrier', ctypes.byref(self.desc), self._stream())

    def allgather_object(self, obj):
        out = [None] * self.size
        dist.all_gather_object(out, obj)
        return out

    def broadcast_object(self, obj, root: int = 0):
        box = [obj]
        dist.broadcast_object_list(box, src=root)
        return box[0]

    def close(self) -> None:
        """Drain the device, wait for every rank (nobody may still be reading this heap), then unmap the multicast view,
        the peers' heaps and this rank's own heap and release their handles.  Tensors handed out by ``alloc`` are dead
        afterwards.  If the ranks cannot be synchronised (a peer died) the mappings are left to process exit."""
        if getattr(self, '_closed', False):
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
            self._pg_barrier()
        except Exception:  # noqa: BLE001 - a dead peer: do not unmap under a possibly running kernel
            return
        self._heap_u8 = None
        size = self.heap_bytes
        try:
            if self.mc_base:
                N.call('ddls_symm_unmap', self.mc_base, size, self._mc_handle)
                self.mc_base, self._mc_handle = 0, 0
            for r in range(self.size):
                if r != self.rank and self._bases[r]:
                    N.call('ddls_symm_unmap', self._bases[r], size, self._handles[r])
                    self._bases[r] = self._handles[r] = 0
            self._pg_barrier()                       # every importer has dropped its mapping of my heap
            if self._bases[self.rank]:
                N.call('ddls_symm_unmap', self._bases[self.rank], size, self._handles[self.rank])
                self._bases[self.rank] = self._handles[self.rank] = 0
        except Exception as e:  # noqa: BLE001
            from ..utils.logging import get_logger
            get_logger().warning('peer-memory close: %s', e)


class EmulatedPeerGroup(_HeapBase):
    """``world`` emulated ranks on one GPU: one plain device buffer per rank plays the heap, and every
    collective is a single cooperative launch whose blockIdx.y enumerates the ranks."""

    def __init__(self, world: int, device: torch.device, heap_bytes: int = 64 << 20, timeout_ms: int = 5000):
        self.world, self.device = world, torch.device(device)
        self._init_heap(heap_bytes)
        self.heaps = [torch.zeros(heap_bytes, dtype=torch.uint8, device=self.device) for _ in range(world)]
        self.desc = CommDesc()
        for r in range(world):
            self.desc.base[r] = self.heaps[r].data_ptr()
        self.desc.mc_base = None
        self.desc.world, self.desc.rank, self.desc.emulate = world, 0, 1
        self.desc.timeout_ms = timeout_ms
        self.desc.flags_off, self.desc.seq_off = FLAGS_OFF, SEQ_OFF
        self._init_ll_stage(world)                       # heaps are torch.zeros: already clean
        N.lib()

    def alloc(self, numel: int, dtype=torch.float32) -> Tuple[int, List[torch.Tensor]]:
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        off = self._alloc_off(max(nbytes, 16))
        return off, [h[off:off + nbytes].view(dtype) for h in self.heaps]

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def allreduce(self, off: int, n: int, algo: int, scale: float = 1.0, wire_bf16: bool = False, blocks: int = 0):
        stage, sbytes = self._ensure_stage(n, self.world, algo)
        N.call('ddls_comm_allreduce', ctypes.byref(self.desc), algo, off, off, stage, sbytes, n, scale, int(wire_bf16),
               blocks, self._stream())

    def dp_step(self, grad_off: int, param_off: int, s1_off: int, s2_off: int, n: int, n_opt: int, elem0: int, scale: float,
                kind: int, step: torch.Tensor, lr: float, b1: float = 0.0, b2: float = 0.0, eps: float = 1e-7,
                zero_grads: bool = True, advance: bool = True, specs=None, nspecs: int = 0, blocks: int = 0):
        N.call('ddls_comm_dp_step', ctypes.byref(self.desc), grad_off, param_off, s1_off, s2_off, n, n_opt, elem0, scale, kind,
               step.data_ptr(), lr, b1, b2, eps, int(zero_grads), int(advance), specs, nspecs, blocks, self._stream())

    def broadcast(self, off: int, nbytes: int, root: int, blocks: int = 0):
        N.call('ddls_comm_broadcast', ctypes.byref(self.desc), off, nbytes, root, 0, blocks, self._stream())

    def allgather(self, in_off: int, out_off: int, chunk_bytes: int, blocks: int = 0):
        N.call('ddls_comm_allgather', ctypes.byref(self.desc), in_off, out_off, chunk_bytes, blocks, self._stream())

    def barrier(self):
        N.call('ddls_comm_barrier', ctypes.byref(self.desc), self._stream())
