"""Kernel timeline of one training step (CUDA-graph replay) on rank 0, 1..N GPUs, from CUPTI via torch.profiler:
start offset, duration and gap to the previous kernel.  Guidance only (profiler attached).
    python tools/step_timeline.py nyiso_lstm            # 1 GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29590 tools/step_timeline.py nyiso_lstm
"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributeddeeplearningspark_b200 as ddls
from distributeddeeplearningspark_b200.trainers import StepEngine
from torch.profiler import profile, ProfilerActivity

name = sys.argv[1] if len(sys.argv) > 1 else 'nyiso_lstm'
B = int(sys.argv[2]) if len(sys.argv) > 2 else (128 if name == 'mnist_cnn' else 32)
ddls.init()
rank, size, dev = ddls.rank(), ddls.size(), ddls.device()
m = getattr(ddls.models, name)(seed=0).to(dev)
opt = ddls.optim.get_optimizer(m.optimizer_name, m.parameters(), comm=ddls.runtime.communicator())
opt.zero_grads_in_step = True
ddls.broadcast_parameters(opt, root_rank=0)
dopt = ddls.DistributedOptimizer(opt) if size > 1 else opt
eng = StepEngine(m, optimizer=dopt, loss=m.loss_name, device=dev, use_cuda_graph=True)
if name == 'mnist_cnn':
    x = torch.rand(B, 28, 28, 1, device=dev); y = torch.eye(10, device=dev)[torch.randint(0, 10, (B,), device=dev)]
else:
    T = 168 if name == 'stacked_lstm' else 25
    x = torch.rand(B, T, 1, device=dev); y = torch.rand(B, 1, device=dev)
for _ in range(20):
    eng.step(x, y)
torch.cuda.synchronize()
ddls.barrier()
NSTEP = 12
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(NSTEP):
        eng.step(x, y)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type.name == 'CUDA']
ev.sort(key=lambda e: e.time_range.start)
if rank == 0:
    per = len(ev) // NSTEP
    step = ev[per * (NSTEP - 3): per * (NSTEP - 2)]            # one warm step
    t0 = step[0].time_range.start
    prev_end = t0
    print(f'{name} on {size} GPU(s): {per} device activities per step')
    rows = []
    for e in step:
        s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
        print(f'  +{s:8.2f} us  dur {d:7.2f}  gap {e.time_range.start - prev_end:6.2f}  {e.name[:90]}')
        rows.append({'start_us': s, 'dur_us': d, 'name': e.name})
        prev_end = max(prev_end, e.time_range.end)
    nxt = ev[per * (NSTEP - 2)].time_range.start
    print(f'  step period {nxt - t0:.2f} us, busy {sum(r["dur_us"] for r in rows):.2f} us')
    os.makedirs('gpurun_out', exist_ok=True)
    json.dump(rows, open(f'gpurun_out/timeline_{name}_n{size}.json', 'w'), indent=1)
ddls.shutdown()
