#!/usr/bin/env python
"""Headline benchmark: training throughput of the reference's workloads on N B200 GPUs of one node.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 200 --warmup 20
    python bench.py --gpus 8 --steps 200 --warmup 20     # same, ranks started by this framework's spark.run launcher

Default model/config = BASELINE.json config 2: the MNIST Keras-style CNN of
/root/reference/ddl_mnist_aztk.py:180-196, synthetic 28x28 images, per-rank batch 128, Adam,
categorical cross-entropy, synchronous data parallelism (``DistributedOptimizer``).  ``--model
nyiso_lstm|nyiso_gru|stacked_lstm`` selects BASELINE.json configs 3/4.

Prints ONE JSON line on rank 0:
  value        whole-job images(or samples)/s, device-timed with CUDA events, max over ranks, inputs
               already on the device and rotating through a pool larger than L2
  e2e          the same metric through the public API including, every step, the pinned-host -> device
               copy of that step's inputs and a device -> host read of the loss
  gpu_launches hand-written kernels of this library launched inside the timed region
``--impl reference`` reports why the unmodified reference cannot run here; ``--impl standin`` runs the
PyTorch + cuDNN/cuBLAS + NCCL stand-in from ``baseline/``.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Published numbers (derived from ddl_nyiso_hdi.ipynb:608-609,817-818): aggregate samples/s of the
# reference's 4-CPU-worker HDInsight runs.  The MNIST script records no number.
PUBLISHED = {'nyiso_gru': 228000.0 / 88.475, 'nyiso_lstm': 228000.0 / 99.254, 'mnist_cnn': None, 'stacked_lstm': None}

MODEL_DEFAULTS = {
    'mnist_cnn': dict(batch=128, unit='images/s', metric='mnist_cnn_train_images_per_sec', seq=784),
    'nyiso_lstm': dict(batch=32, unit='samples/s', metric='nyiso_lstm_train_samples_per_sec', seq=25),
    'nyiso_gru': dict(batch=32, unit='samples/s', metric='nyiso_gru_train_samples_per_sec', seq=25),
    'stacked_lstm': dict(batch=512, unit='samples/s', metric='stacked_lstm_h1024_train_samples_per_sec', seq=168),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference', 'standin'])
    ap.add_argument('--model', default=os.environ.get('DDLS_BENCH_MODEL', 'mnist_cnn'), choices=sorted(MODEL_DEFAULTS))
    ap.add_argument('--batch-per-rank', type=int, default=None)
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-dataset', action='store_true', help='feed step(x, y) from the device pool instead of the in-graph gather')
    ap.add_argument('--no-secondary', action='store_true', help='skip the NYISO LSTM run and the allreduce latency table')
    ap.add_argument('--nccl-compare', action='store_true', help='also time NCCL in the allreduce latency table')
    ap.add_argument('--comm', default=None, help='peer (default on GPUs) | nccl')
    return ap.parse_args()


def reference_arm(args):
    """The unmodified reference cannot be installed or run on this image (recorded in DESIGN.md):
    /root/reference has no setup.py/pyproject.toml, and it needs pyspark + a JVM, dist-keras, Keras and
    TensorFlow, none of which are installed or in /opt/wheelhouse."""
    why = ("reference is three Spark/dist-keras/Keras scripts with no setup.py or pyproject.toml (pip: 'not "
           "installable'); pyspark, a JVM, distkeras, keras and tensorflow are absent from the image and wheelhouse")
    try:
        sys.path.insert(0, os.path.join(ROOT, 'baseline', '_ref'))
        import pyspark  # noqa: F401
        import distkeras  # noqa: F401
        why = 'pyspark/distkeras import succeeded unexpectedly; reference arm not wired'
    except Exception:
        pass
    print(json.dumps({'impl': 'reference', 'unavailable': why}))
    return 0


def make_batches(model_name, batch, n_batches, seed, device):
    """Synthetic data of the reference's shapes, generated once on the host (pinned)."""
    import numpy as np
    import torch
    from distributeddeeplearningspark_b200.data import mnist_like_arrays, nyiso_windows_arrays
    if model_name == 'mnist_cnn':
        px, lab = mnist_like_arrays(batch * n_batches, seed=seed)
        return {'pixels': torch.from_numpy(px), 'labels': torch.from_numpy(lab)}
    seq = MODEL_DEFAULTS[model_name]['seq']
    x, y = nyiso_windows_arrays(batch * n_batches, seq_len=seq, seed=seed)
    return {'x': torch.from_numpy(x), 'y': torch.from_numpy(y)}


def measure(args, model_name, ddls, K, W, with_e2e=True, with_clocks=True):
    """Train ``model_name`` for W untimed + K timed steps (kernel-only, then end to end) and return the result fields."""
    import torch
    from distributeddeeplearningspark_b200 import _native
    from distributeddeeplearningspark_b200.trainers import StepEngine
    from distributeddeeplearningspark_b200.utils import ClockSampler
    from distributeddeeplearningspark_b200.ops.preprocess import mnist_preprocess
    rank, size, dev = ddls.rank(), ddls.size(), ddls.device()
    md = MODEL_DEFAULTS[model_name]
    B = (args.batch_per_rank if (args.batch_per_rank and model_name == args.model) else None) or md['batch']
    use_graph = not args.no_graph
    comm = ddls.runtime.communicator()

    def max_over_ranks(v: float) -> float:
        return max(comm.allgather_object(float(v))) if size > 1 else float(v)

    torch.manual_seed(1234)
    model = getattr(ddls.models, model_name)(seed=1234).to(dev)
    loss_name, opt_name = model.loss_name, model.optimizer_name
    base_opt = ddls.optim.get_optimizer(opt_name, model.parameters(), comm=comm)
    base_opt.zero_grads_in_step = True
    ddls.broadcast_parameters(base_opt, root_rank=0)
    dopt = ddls.DistributedOptimizer(base_opt) if size > 1 else base_opt
    engine = StepEngine(model, optimizer=dopt, loss=loss_name, device=dev, use_cuda_graph=use_graph)

    # ---- data: a device pool larger than L2 for the kernel-only number; pinned host for e2e ------------
    is_mnist = model_name == 'mnist_cnn'
    bytes_per_batch = B * (784 * 4 + 10 * 4) if is_mnist else B * (md['seq'] + 1) * 4
    pool_n = max(8, min(65536, (160 << 20) // bytes_per_batch + 1))      # > 126 MB of L2 for every model
    host = make_batches(model_name, B, pool_n, seed=rank, device=dev)
    if is_mnist:
        px = host['pixels'].to(dev); lab = host['labels'].to(dev)
        x_all, y_all = mnist_preprocess(px, lab)
        del px
    else:
        x_all, y_all = host['x'].to(dev), host['y'].to(dev)
    x_all, y_all = x_all[:pool_n * B].contiguous(), y_all[:pool_n * B].contiguous()
    pool_bytes = x_all.numel() * x_all.element_size() + y_all.numel() * y_all.element_size()
    # The training set stays on the device (the reference caches its DataFrame partitions in executor memory); every
    # step gathers its minibatch from this pool INSIDE the captured graph (device-side cursor), so consecutive steps
    # read different, L2-cold inputs and a step is one graph launch.
    use_dataset = engine.use_cuda_graph and not args.no_dataset
    if use_dataset:
        engine.attach_dataset(x_all, y_all, B)
    pool_x = list(x_all.split(B)); pool_y = list(y_all.split(B))

    def train_step(i):
        return engine.step_dataset() if use_dataset else engine.step(pool_x[i % pool_n], pool_y[i % pool_n])

    def sync_all():
        torch.cuda.synchronize(dev)
        if size > 1:
            ddls.barrier()
            torch.cuda.synchronize(dev)

    # ---- launches per step (counted on one eager step; a graph replays exactly these) --------------------
    for i in range(max(W, 5)):                      # includes the three eager warm-ups and the graph capture
        train_step(i)
    sync_all()
    _log(f'{model_name}: warm')
    c0 = _native.launch_count()
    if use_dataset:
        engine._gather()
    engine._eager_step(pool_x[0], pool_y[0])
    per_step_launches = _native.launch_count() - c0
    sync_all()

    # ---- kernel-only: device-resident inputs, CUDA events, max over ranks ---------------------------------
    sampler = ClockSampler(gpu_index=dev.index or 0, period_ms=20).start() if (rank == 0 and with_clocks) else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    ev0.record()
    for i in range(K):
        train_step(i)
    ev1.record()
    sync_all()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    _log(f'{model_name}: kernel-only timed')
    last_loss = float(engine.step(pool_x[0], pool_y[0]).item())

    # ---- end to end: pinned host -> H2D every step, device -> host loss read every step -------------------
    e2e = None
    if with_e2e and not args.no_e2e:
        n_host = min(pool_n, 512)
        if is_mnist:
            hp = host['pixels'][:n_host * B].reshape(n_host, B, 784).pin_memory()
            hl = host['labels'][:n_host * B].reshape(n_host, B).pin_memory()
            d_px = torch.empty((B, 784), dtype=torch.uint8, device=dev)
            d_lab = torch.empty((B,), dtype=torch.int64, device=dev)
            h2d = B * 784 + B * 8
        else:
            hx = host['x'][:n_host * B].reshape(n_host, B, md['seq'], 1).pin_memory()
            hy = host['y'][:n_host * B].reshape(n_host, B, 1).pin_memory()
            h2d = B * (md['seq'] + 1) * 4
        # Public API under test: StepEngine.run_stream — pinned host batches -> copy stream (H2D + on-device
        # preprocessing) -> captured step graph; the loss of every step is read back (one step behind the launches).
        if is_mnist:
            def stage(hb, xs, ys):
                d_px.copy_(hb[0], non_blocking=True); d_lab.copy_(hb[1], non_blocking=True)
                mnist_preprocess(d_px, d_lab, out_x=xs, out_y=ys)

            def host_batches(n):
                return ((hp[i % n_host], hl[i % n_host]) for i in range(n))
        else:
            stage = None

            def host_batches(n):
                return ((hx[i % n_host], hy[i % n_host]) for i in range(n))

        engine.run_stream(host_batches(max(W, 6)), pool_x[0], pool_y[0], stage=stage)   # eager warm-ups + both captures
        sync_all()
        t0 = time.perf_counter()
        e2e_losses = engine.run_stream(host_batches(K), pool_x[0], pool_y[0], stage=stage)
        sync_all()
        assert len(e2e_losses) == K
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e = {'value': B * size * K / e2e_s, 'unit': md['unit'], 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4,
               'ms_per_step': 1e3 * e2e_s / K}

    clocks = sampler.stop() if sampler is not None else None
    value = B * size * K / (ms / 1e3)
    pub = PUBLISHED.get(model_name)
    dopt = engine.optimizer                      # (a lone rank: the engine wraps the optimizer for per-bucket updates)
    fused = bool(getattr(dopt, '_fused', False))
    return {
        'metric': md['metric'], 'value': value, 'unit': md['unit'], 'n_gpus': size, 'steps': K, 'warmup': W,
        'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': (value / pub) if pub else None, 'dtype': 'bf16 tensor-core operands, fp32 accumulate/master',
        'data': 'synthetic', 'impl': 'ours',
        'config': {'model': model_name, 'global_batch': B * size, 'per_rank_batch': B, 'seq_len': md['seq'],
                   'parallelism': f'dp{size}', 'optimizer': opt_name, 'loss': loss_name,
                   'comm': type(comm).__name__,
                   'dp_step': ([dopt._bucket_mode(b) for b in dopt.buckets] if fused else ('unfused' if size > 1 else 'single rank')),
                   'cuda_graph': bool(engine.use_cuda_graph and (engine._graph is not None or
                                                                 getattr(engine, '_ds_graph', None) is not None)),
                   'l2': (f'inputs rotate through a {pool_bytes >> 20} MB device pool (> 126 MB L2)'
                          if pool_bytes > (126 << 20) else f'{pool_bytes >> 20} MB input pool (smaller than L2)'),
                   'input_feed': ('minibatch gathered inside the step graph from the device-resident pool '
                                  '(device-side cursor)') if use_dataset else 'step(x, y) from the device pool'},
        'e2e': e2e, 'gpu_launches': int(per_step_launches * K), 'gpu_launches_per_step': int(per_step_launches),
        'clocks': clocks, 'final_loss': last_loss,
    }


def allreduce_latency(ddls, sizes=(4 << 10, 256 << 10, 4 << 20), with_nccl=False):
    """Device time per allreduce call (fp32 SUM, in place) of this library's selected algorithm, timed inside a CUDA graph
    of 16 calls (no host launch overhead), max over ranks.  The third leg of BASELINE.json's metric ("allreduce us").
    ``with_nccl`` (``--nccl-compare``) adds NCCL timed the same way; by default the bench never creates an NCCL
    communicator (the full comparison, 4 KB - 256 MB, is ``bench/allreduce_sweep.py`` -> ``profiles/sweep_n*.json``)."""
    import torch
    import torch.distributed as dist
    rank, size, dev = ddls.rank(), ddls.size(), ddls.device()
    comm = ddls.runtime.communicator()
    if size == 1 or not hasattr(comm, 'select_algo'):
        return None
    names = {0: 'oneshot', 1: 'twoshot', 2: 'nvls', 3: 'flag-in-data'}

    def graph_us(fn, calls=16, replays=10):
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                fn()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev); comm.barrier(); torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(calls):
                fn()
        g.replay(); torch.cuda.synchronize(dev); comm.barrier(); torch.cuda.synchronize(dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(replays):
            g.replay()
        b.record(); torch.cuda.synchronize(dev)
        return max(comm.allgather_object(a.elapsed_time(b) / (replays * calls) * 1e3))

    out = {}
    nccl_group = None
    if with_nccl:
        try:
            nccl_group = dist.group.WORLD if dist.get_backend() == 'nccl' else dist.new_group(backend='nccl')
        except Exception:  # noqa: BLE001
            nccl_group = None
    for nbytes in sizes:
        n = nbytes // 4
        t = comm.alloc(n)
        row = {'ours_us': graph_us(lambda: comm.allreduce_(t, ddls.Sum)),
               'algorithm': names[comm.select_algo(nbytes, True, bool(comm.mc_base))]}
        if nccl_group is not None:
            tn = torch.zeros(n, device=dev)
            try:
                row['nccl_us'] = graph_us(lambda: dist.all_reduce(tn, group=nccl_group))
            except Exception as e:  # noqa: BLE001
                row['nccl_us'] = None
        out[str(nbytes)] = row
    return out


def _log(msg):
    if os.environ.get('RANK', '0') == '0':
        print(f'[bench] {msg}', file=sys.stderr, flush=True)


def run_rank(args):
    """One rank of the benchmark (this process is rank RANK of WORLD_SIZE): returns the result dict on rank 0."""
    import distributeddeeplearningspark_b200 as ddls
    if args.comm:
        os.environ['DDLS_COMM'] = args.comm
    ddls.init()
    rank = ddls.rank()
    K, W = args.steps, max(3, args.warmup)

    if args.impl == 'standin':
        sys.path.insert(0, os.path.join(ROOT, 'baseline'))
        import standin
        md = MODEL_DEFAULTS[args.model]
        standin.run(args, ddls, md, args.batch_per_rank or md['batch'], K, W)
        return None

    out = measure(args, args.model, ddls, K, W)
    _log(f'{args.model}: {out["value"]:.0f} {out["unit"]}, {out["ms_per_step"] * 1e3:.1f} us/step')
    # BASELINE.json's metric has three legs: MNIST images/s (the primary line), NYISO LSTM samples/s and allreduce latency.
    # The default run measures all three so that the driver's record carries them at every N.
    if args.model == 'mnist_cnn' and not args.no_secondary:
        sec = measure(args, 'nyiso_lstm', ddls, K, W, with_clocks=False)
        _log(f'nyiso_lstm: {sec["value"]:.0f} samples/s')
        pub = PUBLISHED['nyiso_lstm']
        out['secondary'] = {k: sec[k] for k in ('metric', 'value', 'unit', 'ms_per_step', 'steps', 'warmup', 'e2e',
                                                'gpu_launches', 'gpu_launches_per_step', 'final_loss')}
        out['secondary']['vs_published'] = sec['value'] / pub
        out['secondary']['published'] = {'value': pub, 'unit': 'samples/s', 'what': '4 CPU workers, ddl_nyiso_hdi.ipynb:817-818'}
        out['secondary']['config'] = sec['config']
        if ddls.size() > 1:
            out['allreduce_us'] = allreduce_latency(ddls, with_nccl=args.nccl_compare)
    ddls.shutdown()
    return out if rank == 0 else None


def _spark_rank(argv):
    """Entry of a rank started by ``ddls.spark.run`` (BASELINE.json config 2: "via Spark launcher")."""
    sys.argv = ['bench.py'] + list(argv)
    return run_rank(parse())


def main():
    args = parse()
    if args.impl == 'reference':
        return reference_arm(args)
    import faulthandler
    # a hung collective must not hang the caller for ever: dump every thread's stack and exit after DDLS_BENCH_WATCHDOG_S
    faulthandler.dump_traceback_later(int(os.environ.get('DDLS_BENCH_WATCHDOG_S', '900')), exit=True)

    import torch
    if not torch.cuda.is_available():
        print(json.dumps({'impl': args.impl, 'unavailable': 'no CUDA device visible'}))
        return 0
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if world == 1 and args.gpus > 1:
        # not started by torch.distributed.run: start the ranks through this framework's own Spark-style launcher
        # (barrier stage, one rank per GPU) — the path BASELINE.json config 2 names
        if args.impl != 'ours':
            print(json.dumps({'impl': args.impl, 'unavailable':
                              f'--gpus {args.gpus} must be launched with torch.distributed.run (one rank per GPU)'}))
            return 0
        import distributeddeeplearningspark_b200 as ddls
        results = ddls.spark.run(_spark_rank, args=(sys.argv[1:],), num_proc=args.gpus)
        out = results[0]
        out['config']['launcher'] = 'spark.run (local barrier backend)'
        print(json.dumps(out))
        return 0
    out = run_rank(args)
    if out is not None:
        out['config']['launcher'] = 'torch.distributed.run' if world > 1 else 'single process'
        print(json.dumps(out))
    return 0


if __name__ == '__main__':
    sys.exit(main())
