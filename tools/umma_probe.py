"""Descriptor-encoding probe: run the tcgen05 GEMM under every operand-major combination and with
LBO/SBO swapped (debug bits), print max|err| vs fp32 torch.  Run on a B200 via gpurun."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributeddeeplearningspark_b200.ops.gemm import gemm_bf16

torch.manual_seed(0)
dev = torch.device('cuda')
res = {}
for (M, N, K) in ((128, 32, 64), (256, 96, 192), (384, 225, 4608)):
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    ref = A.float() @ B.float().t()
    for a_mn in (0, 1):
        for b_mn in (0, 1):
            for dbg in (0, 1, 2, 3):
                a = A.t().contiguous() if a_mn else A
                b = B.t().contiguous() if b_mn else B
                out = gemm_bf16(a, b, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), debug=dbg)
                torch.cuda.synchronize()
                err = (out - ref).abs().max().item()
                res[f'{M}x{N}x{K} a_mn={a_mn} b_mn={b_mn} dbg={dbg}'] = err
                print(f'{M}x{N}x{K} a_mn={a_mn} b_mn={b_mn} dbg={dbg}: max err {err:.4f} (ref scale {ref.abs().max().item():.1f})', flush=True)
os.makedirs('gpurun_out', exist_ok=True)
json.dump(res, open('gpurun_out/umma_probe.json', 'w'), indent=1)
