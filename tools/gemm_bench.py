"""Times the generic GEMM entry on the stacked-LSTM shapes (TMA kernel vs the cp.async kernel vs cuBLAS).  Every timed
call is bracketed by its own CUDA events with an L2 flush (256 MB overwrite) before it, so no operand is L2-resident from
the previous iteration; the median of the iterations is reported; clocks are sampled during the run."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributeddeeplearningspark_b200.ops.gemm import gemm_bf16
from distributeddeeplearningspark_b200.utils import ClockSampler, flush_l2

dev = torch.device('cuda', 0)
def t(fn, iters=7):
    for _ in range(3): fn()                     # warm-up: lazy module load, tensor maps, allocator
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        flush_l2(dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    ms.sort()
    return ms[len(ms) // 2]

sampler = ClockSampler(gpu_index=0, period_ms=100).start()

TB, H = 512 * 168, 1024
shapes = [('zx   [TB,H]x[H,4H]', (TB, H), (H, 4 * H), False, True, torch.bfloat16),
          ('dX   [TB,4H]x[H,4H]^T', (TB, 4 * H), (H, 4 * H), False, False, torch.bfloat16),
          ('dW   [TB,H]^T x [TB,4H]', (TB, H), (TB, 4 * H), True, True, torch.float32),
          ('step [512,H]x[H,4H]', (512, H), (H, 4 * H), False, True, torch.float32)]
rows = []
for name, sa, sb, amn, bmn, od in shapes:
    a = torch.randn(*sa, device=dev).bfloat16(); b = torch.randn(*sb, device=dev).bfloat16()
    M, K = (sa[1], sa[0]) if amn else sa
    N = sb[1] if bmn else sb[0]
    out = torch.empty(M, N, device=dev, dtype=od)
    fl = 2.0 * M * N * K
    t_tma = t(lambda: gemm_bf16(a, b, a_mn_major=amn, b_mn_major=bmn, out=out))
    t_old = t(lambda: gemm_bf16(a, b, a_mn_major=amn, b_mn_major=bmn, out=out, no_tma=True), iters=3)
    A2 = a.t() if amn else a; B2 = b if bmn else b.t()
    t_cub = t(lambda: torch.matmul(A2, B2))
    ref = torch.matmul(A2, B2).float()
    err = float((out.float() - ref).norm() / ref.norm())
    row = dict(name=name, M=M, N=N, K=K, tma_ms=t_tma, cp_async_ms=t_old, cublas_ms=t_cub, tma_tflops=fl / t_tma / 1e9,
               cublas_tflops=fl / t_cub / 1e9, rel_err_vs_cublas=err)
    rows.append(row); print(json.dumps(row), flush=True)
    del a, b, out, ref
os.makedirs('gpurun_out', exist_ok=True)
json.dump({'clocks': sampler.stop(), 'timing': 'median of per-call CUDA-event times, L2 flushed before every call, 3 warm-up calls',
           'rows': rows}, open('gpurun_out/gemm_bench.json', 'w'), indent=1)
