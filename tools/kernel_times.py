"""Warm per-kernel device times of one training step (CUDA-graph replays) via torch.profiler / CUPTI.
Guidance only — numbers taken under a profiler are never reported as bench values."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributeddeeplearningspark_b200 as ddls
from distributeddeeplearningspark_b200.trainers import StepEngine
from torch.profiler import profile, ProfilerActivity
from distributeddeeplearningspark_b200.utils import ClockSampler

name = sys.argv[1] if len(sys.argv) > 1 else 'mnist_cnn'
B = int(sys.argv[2]) if len(sys.argv) > 2 else (128 if name == 'mnist_cnn' else 32)
dev = torch.device('cuda', 0)
m = getattr(ddls.models, name)(seed=0).to(dev)
eng = StepEngine(m, optimizer=m.optimizer_name, loss=m.loss_name, device=dev, use_cuda_graph=True)
if name == 'mnist_cnn':
    x = torch.rand(B, 28, 28, 1, device=dev); y = torch.eye(10, device=dev)[torch.randint(0, 10, (B,), device=dev)]
else:
    T = 168 if name == 'stacked_lstm' else 25
    x = torch.rand(B, T, 1, device=dev); y = torch.rand(B, 1, device=dev)
for _ in range(10):
    eng.step(x, y)
torch.cuda.synchronize()
N = 20
sampler = ClockSampler(gpu_index=0, period_ms=20).start()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(N):
        eng.step(x, y)
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages():
    dt = getattr(e, 'device_time_total', None) or getattr(e, 'cuda_time_total', 0)
    if dt and e.device_type.name == 'CUDA' if hasattr(e, 'device_type') else dt:
        rows.append((dt / N, e.count / N, e.key))
rows.sort(reverse=True)
tot = 0
for t, c, k in rows:
    if t < 0.05: continue
    tot += t
    print(f'{t:9.2f} us/step  x{c:4.1f}  {k[:100]}')
print(f'{tot:9.2f} us/step total kernel time')
os.makedirs('gpurun_out', exist_ok=True)
json.dump({'model': name, 'batch': B, 'how': 'CUPTI device times of 20 warm CUDA-graph replays (profiler attached: guidance, not bench values)',
           'clocks': sampler.stop(), 'total_us_per_step': tot,
           'kernels': [{'us_per_step': t, 'calls_per_step': c, 'kernel': k} for t, c, k in rows]},
          open(f'gpurun_out/kernel_times_{name}.json', 'w'), indent=1)
