"""NCCL/cuDNN/cuBLAS stand-in baseline (NOT Horovod, NOT the reference).

The reference's own stack (Spark + dist-keras + Keras + TensorFlow, CPU) and the Horovod build named by
BASELINE.json cannot be installed offline.  This is the same-box comparator built from what
the image has: plain PyTorch modules (cuDNN conv / cuDNN LSTM-GRU, cuBLAS GEMMs, bf16 autocast),
``DistributedDataParallel`` (bucketed NCCL allreduce overlapped with backward) and ``torch.optim``, with the WHOLE training
step captured in a CUDA graph (so the comparison with this framework's captured step is launch-overhead-free on both
sides; ``--no-graph`` times the eager form).
Same synthetic data shapes, same batch sizes, same JSON line as ``bench.py`` (``"impl": "standin"``).
"""
from __future__ import annotations

import json
import time

import torch
import torch.nn as nn
import torch.nn.functional as F


class CNN(nn.Module):
    def __init__(self):
        super().__init__()
        self.c1 = nn.Conv2d(1, 32, 3); self.c2 = nn.Conv2d(32, 32, 3)
        self.d1 = nn.Linear(4608, 225); self.d2 = nn.Linear(225, 10)

    def forward(self, x):  # x: (B,1,28,28)
        x = F.relu(self.c1(x)); x = F.relu(self.c2(x)); x = F.max_pool2d(x, 2)
        x = x.flatten(1)
        return self.d2(F.relu(self.d1(x)))


class RNNReg(nn.Module):
    def __init__(self, cell, hidden, layers=1):
        super().__init__()
        self.rnn = (nn.LSTM if cell == 'lstm' else nn.GRU)(1, hidden, num_layers=layers, batch_first=True)
        self.out = nn.Linear(hidden, 1)

    def forward(self, x):
        y, _ = self.rnn(x)
        return self.out(y[:, -1])


def run(args, ddls, md, B, K, W):
    import torch.distributed as dist
    rank, size, dev = ddls.rank(), ddls.size(), ddls.device()
    torch.manual_seed(1234)
    name = args.model
    if name == 'mnist_cnn':
        model = CNN().to(dev).to(memory_format=torch.channels_last)
        opt_cls, loss_fn = torch.optim.Adam, lambda o, y: F.cross_entropy(o.float(), y)
    elif name == 'stacked_lstm':
        model = RNNReg('lstm', 1024, 2).to(dev); opt_cls, loss_fn = torch.optim.Adam, lambda o, y: F.mse_loss(o.float(), y)
    else:
        model = RNNReg('lstm' if 'lstm' in name else 'gru', 128).to(dev)
        opt_cls = torch.optim.Adam if 'lstm' in name else torch.optim.Adagrad
        loss_fn = lambda o, y: F.mse_loss(o.float(), y)
    use_graph = not getattr(args, 'no_graph', False)
    nccl_group = None
    if size > 1:
        # the framework's own process group may be gloo (plumbing only): the stand-in gets an NCCL group of its own
        nccl_group = dist.group.WORLD if dist.get_backend() == 'nccl' else dist.new_group(backend='nccl')
    if size > 1:
        if use_graph:   # DDP must be built inside a side-stream context when its step is going to be captured
            s0 = torch.cuda.Stream(device=dev)
            with torch.cuda.stream(s0):
                model = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True,
                                                            process_group=nccl_group)
            torch.cuda.current_stream(dev).wait_stream(s0)
        else:
            model = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True,
                                                        process_group=nccl_group)
    okw = {'capturable': True} if (opt_cls is torch.optim.Adam and use_graph) else {}
    opt = opt_cls(model.parameters(), lr=1e-3 if opt_cls is torch.optim.Adam else 0.01, **okw)
    pool_n = 64
    if name == 'mnist_cnn':
        xs = [torch.rand(B, 1, 28, 28, device=dev).contiguous(memory_format=torch.channels_last) for _ in range(pool_n)]
        ys = [torch.randint(0, 10, (B,), device=dev) for _ in range(pool_n)]
    else:
        xs = [torch.rand(B, md['seq'], 1, device=dev) for _ in range(pool_n)]
        ys = [torch.rand(B, 1, device=dev) for _ in range(pool_n)]

    def eager_step(x, y):
        opt.zero_grad(set_to_none=not use_graph)
        with torch.autocast('cuda', dtype=torch.bfloat16):
            out = model(x)
        loss = loss_fn(out, y)
        loss.backward()
        opt.step()
        return loss

    graph_note = 'eager'
    step = eager_step
    if use_graph:
        try:
            sx, sy = xs[0].clone(), ys[0].clone()
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(11):                      # DDP needs ~11 warm-up iterations before capture
                    eager_step(sx, sy)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                static_loss = eager_step(sx, sy)

            def step(x, y):
                sx.copy_(x, non_blocking=True); sy.copy_(y, non_blocking=True)
                g.replay()
                return static_loss
            graph_note = 'whole step captured in one CUDA graph'
        except Exception as e:  # noqa: BLE001 - e.g. an optimizer without a capturable step
            torch.cuda.synchronize(dev)
            graph_note = f'eager (graph capture failed: {type(e).__name__}: {str(e)[:120]})'
            use_graph = False
            step = eager_step

    def sync():
        torch.cuda.synchronize(dev)
        if size > 1:
            ddls.barrier(); torch.cuda.synchronize(dev)

    for i in range(W):
        step(xs[i % pool_n], ys[i % pool_n])
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        step(xs[i % pool_n], ys[i % pool_n])
    e1.record(); sync()
    comm = ddls.runtime.communicator()
    ms = max(comm.allgather_object(float(e0.elapsed_time(e1)))) if size > 1 else float(e0.elapsed_time(e1))
    # end to end: pinned host inputs every step + loss read back
    hx = [x.cpu().pin_memory() for x in xs[:16]]; hy = [y.cpu().pin_memory() for y in ys[:16]]
    sync(); t0 = time.perf_counter()
    for i in range(K):
        x = hx[i % 16].to(dev, non_blocking=True); y = hy[i % 16].to(dev, non_blocking=True)
        float(step(x, y).item())
    sync()
    e2e_s = time.perf_counter() - t0
    if size > 1:
        e2e_s = max(comm.allgather_object(float(e2e_s)))
    if rank == 0:
        print(json.dumps({'metric': md['metric'], 'value': B * size * K / (ms / 1e3), 'unit': md['unit'], 'n_gpus': size,
                          'steps': K, 'warmup': W, 'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': 'weak',
                          'vs_baseline': None, 'dtype': 'bf16 autocast', 'data': 'synthetic',
                          'impl': 'standin (PyTorch + cuDNN/cuBLAS + NCCL DDP; not Horovod, not the reference)',
                          'config': {'model': name, 'global_batch': B * size, 'seq_len': md['seq'],
                                     'parallelism': f'dp{size}', 'step': graph_note},
                          'e2e': {'value': B * size * K / e2e_s, 'unit': md['unit'],
                                  'h2d_bytes_per_step': hx[0].numel() * hx[0].element_size() + hy[0].numel() * hy[0].element_size(),
                                  'd2h_bytes_per_step': 4},
                          'gpu_launches': 0}))
    ddls.shutdown()
    return 0
