"""Allreduce latency / bandwidth sweep, 4 KB .. 256 MB: this library's one-shot / two-shot / NVLS kernels vs NCCL.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        -m distributeddeeplearningspark_b200.bench.allreduce_sweep [--max-mb 256] [--out gpurun_out/sweep.json]

BASELINE.json config 5.  Every number is device-timed with CUDA events on the launching stream and is
the MAX over ranks; buffers are warm; each size is timed as the mean of ``iters`` back-to-back calls.
Roofline per algorithm (link = 770 GB/s measured per direction per GPU, B200_PROFILING.md):
one-shot moves (N-1)·S in per rank, two-shot 2(N-1)/N·S, NVLS ≈ S.
"""
from __future__ import annotations

import argparse
import json
import os

import torch
import torch.distributed as dist

LINK_GBS = 770.0


_COMM = None          # this library's communicator: host barriers and max-over-ranks go through it, not through NCCL
_NCCL = None          # NCCL group for the comparison arm only


def _max_over_ranks(v):
    return max(_COMM.allgather_object(float(v)))


def _barrier(dev):
    torch.cuda.synchronize(dev); _COMM.barrier(); torch.cuda.synchronize(dev)


def _time(fn, iters, dev):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        fn()
    _barrier(dev)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize(dev)
    return _max_over_ranks(a.elapsed_time(b) / iters * 1e3)   # us


def _time_graph(fn, dev, calls=16, replays=20):
    """Device time per call with host launch overhead removed: `calls` collectives per captured graph."""
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream(dev).wait_stream(side)
    _barrier(dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(calls):
            fn()
    g.replay()
    _barrier(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(replays):
        g.replay()
    b.record()
    torch.cuda.synchronize(dev)
    return _max_over_ranks(a.elapsed_time(b) / (replays * calls) * 1e3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--max-mb', type=int, default=256)
    ap.add_argument('--out', default=None)
    ap.add_argument('--graph', action='store_true', help='also time inside a CUDA graph (8 calls per replay)')
    args = ap.parse_args()
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.parallel.peer import AR_NVLS, AR_ONESHOT, AR_ONESHOT_LL, AR_TWOSHOT, PeerMemoryComm
    os.environ.setdefault('DDLS_HEAP_MB', str(768 + 44 * min(args.max_mb, 64) + 2 * args.max_mb))   # buffer + one-shot staging (2 x N x message, <= 16 MB messages) at N = 8
    ddls.init()
    comm = ddls.runtime.communicator()
    assert isinstance(comm, PeerMemoryComm), f'peer-memory communicator unavailable: {type(comm).__name__}'
    global _COMM, _NCCL
    _COMM = comm
    _NCCL = dist.group.WORLD if dist.get_backend() == 'nccl' else dist.new_group(backend='nccl')
    r, n, dev = ddls.rank(), ddls.size(), ddls.device()
    from distributeddeeplearningspark_b200.utils import ClockSampler
    sampler = ClockSampler(gpu_index=dev.index or 0, period_ms=100).start() if r == 0 else None
    sizes = []
    s = 4 << 10
    while s <= args.max_mb << 20:
        sizes.append(s); s *= 4
    buf = comm.alloc((args.max_mb << 20) // 4)
    nccl_buf = torch.empty((args.max_mb << 20) // 4, device=dev)
    rows = []
    for nbytes in sizes:
        numel = nbytes // 4
        t, tn = buf[:numel], nccl_buf[:numel]
        iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (32 << 20) else 10)
        row = {'bytes': nbytes, 'nccl_us': _time(lambda: dist.all_reduce(tn, group=_NCCL), iters, dev)}
        algos = [('oneshot', AR_ONESHOT)] + ([('twoshot', AR_TWOSHOT)] if nbytes <= (64 << 20) else []) + \
            ([('nvls', AR_NVLS)] if comm.mc_base else [])
        if nbytes <= comm._ll_elems * 4:
            algos.insert(0, ('ll', AR_ONESHOT_LL))
        for name, a in algos:
            if name == 'oneshot' and nbytes > (16 << 20):
                continue
            us = _time(lambda: comm.allreduce_(t, ddls.Sum, algo=a), iters, dev)
            link_bytes = {'ll': 2 * (n - 1) * nbytes, 'oneshot': (n - 1) * nbytes, 'twoshot': 2 * (n - 1) / n * nbytes,
                          'nvls': nbytes}[name]
            row[name + '_us'] = us
            if args.graph and nbytes <= (4 << 20):
                row[name + '_graph_us'] = _time_graph(lambda: comm.allreduce_(t, ddls.Sum, algo=a), dev)
            row[name + '_roofline_frac'] = (link_bytes / (LINK_GBS * 1e3)) / us      # t_link / t_measured
        if args.graph and nbytes <= (4 << 20):
            try:
                row['nccl_graph_us'] = _time_graph(lambda: dist.all_reduce(tn, group=_NCCL), dev)
            except Exception as e:  # noqa: BLE001
                row['nccl_graph_us'] = None
        best = min((row[k], k) for k in row if k.endswith('_us') and not k.startswith('nccl') and 'graph' not in k)
        row['best'] = best[1][:-3]; row['speedup_vs_nccl'] = row['nccl_us'] / best[0]
        row['busbw_gbs_best'] = 2 * (n - 1) / n * nbytes / best[0] / 1e3
        rows.append(row)
        if r == 0:
            print(json.dumps(row), flush=True)
    clocks = sampler.stop() if sampler is not None else None
    if r == 0 and args.out:
        os.makedirs(os.path.dirname(args.out) or '.', exist_ok=True)
        with open(args.out, 'w') as f:
            json.dump({'n_gpus': n, 'multicast': bool(comm.mc_base), 'link_gbs': LINK_GBS, 'clocks': clocks,
                       'timing': 'CUDA events on the launching stream, warm, max over ranks; *_graph_us: 16 calls per captured '
                                 'graph (no host launch overhead on either side)', 'rows': rows}, f, indent=1)
    ddls.shutdown()


if __name__ == '__main__':
    main()
