"""Phase timing of the sharded data-parallel step kernel (ddls_comm_dp_step) on the MNIST-sized bucket, N GPUs:
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29591 tools/dp_step_trace.py
Prints, for rank 0, device time per call (CUDA events around a graph of 16 calls) and the in-kernel globaltimer stamps of
the LAST of 8 back-to-back calls (steady state: barrier A wait, slice work, barrier B wait, local pass)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import faulthandler
faulthandler.dump_traceback_later(int(os.environ.get('DUMP_AFTER', '60')), exit=True)
import torch
import distributeddeeplearningspark_b200 as ddls
from distributeddeeplearningspark_b200 import _native as N

ddls.init()
r, n, dev = ddls.rank(), ddls.size(), ddls.device()
comm = ddls.runtime.communicator()
N.register('ddls_comm_dp_step_trace', [N.c_void_p])
rows = int(os.environ.get('ROWS', '4608'))
shadow = os.environ.get('SHADOW', '1') == '1'
ps = [torch.nn.Parameter(torch.randn(rows, 225, device=dev))]
opt = ddls.optim.KerasAdam(ps, comm=comm)
opt.zero_grads_in_step = True
if shadow:
    opt.add_shadow(ps[0], 232)
nel = opt.flat.numel
for blocks in [int(b) for b in os.environ.get('BLOCKS', '0,32,64').split(',')]:
    def call():
        comm.dp_step_(opt, 0, nel, ddls.Average, advance=True, blocks=blocks)
    for _ in range(5):
        call()
    torch.cuda.synchronize(); ddls.barrier(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(16):
            call()
    g.replay(); torch.cuda.synchronize(); ddls.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        g.replay()
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) / 320 * 1e3
    tr = torch.zeros(5 * 128, dtype=torch.int64, device=dev)
    N.call('ddls_comm_dp_step_trace', tr.data_ptr())
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g2):
        for _ in range(8):
            call()
    N.call('ddls_comm_dp_step_trace', None)
    ddls.barrier(); torch.cuda.synchronize()
    g2.replay(); torch.cuda.synchronize()
    t = tr.view(128, 5).cpu()
    t = t[t[:, 0] > 0]
    t0 = int(t[:, 0].min())
    ph = (t - t0).float() / 1e3
    if r == 0:
        print(f'blocks={blocks or "auto"} ({t.shape[0]} CTAs) {nel * 4 / 1e6:.2f} MB x{n} shadow={shadow}: {us:.2f} us/call | mean stamps us: '
              f'entry {ph[:,0].mean():.2f} barA {ph[:,1].mean():.2f} slice {ph[:,2].mean():.2f} barB {ph[:,3].mean():.2f} '
              f'end {ph[:,4].mean():.2f} (max end {ph[:,4].max():.2f})', flush=True)
    ddls.barrier()
ddls.shutdown()
