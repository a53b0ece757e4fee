"""A short eager (no CUDA graph) training run for `ncu`: python tools/ncu_target.py <model> [steps].
Run it once without ncu first (B200_PROFILING.md), then e.g.
    ncu --set full --clock-control none --import-source on -k regex:'adam_kernel|umma_gemm' -s 27 -c 9 -o gpurun_out/prof \
        python tools/ncu_target.py mnist_cnn"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributeddeeplearningspark_b200 as ddls
from distributeddeeplearningspark_b200.trainers import StepEngine

name = sys.argv[1] if len(sys.argv) > 1 else 'mnist_cnn'
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
dev = torch.device('cuda', 0)
m = getattr(ddls.models, name)(seed=0).to(dev)
eng = StepEngine(m, optimizer=m.optimizer_name, loss=m.loss_name, device=dev, use_cuda_graph=False)
if name == 'mnist_cnn':
    x = torch.rand(128, 28, 28, 1, device=dev); y = torch.eye(10, device=dev)[torch.randint(0, 10, (128,), device=dev)]
else:
    x = torch.rand(32, 25, 1, device=dev); y = torch.rand(32, 1, device=dev)
for _ in range(steps):
    loss = eng.step(x, y)
torch.cuda.synchronize()
print(f'{name}: {steps} eager steps, loss {float(loss):.5f}')
