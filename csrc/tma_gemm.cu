// Persistent, warp-specialised tcgen05 GEMM fed by tensor-map TMA (csrc/tma_pipe.cuh): C[M,N] = op(A)·op(B)ᵀ.
//
//   warp 0   : TMA producer (one thread): box loads of the A/B K-blocks into a 4- or 6-deep ring
//   warp 1   : MMA issuer (one thread) + TMEM owner; accumulators are double-buffered in TMEM (2 x BN columns)
//   warps 2-9: epilogue (two warps per TMEM lane quadrant, half of the columns each) — tcgen05.ld of a finished
//              accumulator, bias (staged once per tile in shared memory) / ReLU / cast / accumulate, 16-byte
//              stores — overlapped with the next tile's MMAs
// Tiles are 128 x BN (BN = 256 when N >= 256), visited n-fastest by a static persistent schedule so the CTAs
// running at the same time share A tiles in L2.  Used for the big GEMMs of the scaled stacked LSTM (hoisted input
// projection, dX, the K = B·T weight gradients) where the cp.async kernel of umma_gemm.cu reached 13% of peak.
// (Keras Dense / LSTM kernels: /root/reference/ddl_nyiso_aztk.py:249-251.)
#include "tma_pipe.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

struct TmaGemmParams {
  void* C;
  const float* bias;
  int M, N, K, ldc;
  unsigned flags;
  int tiles_m, tiles_n;
};

constexpr int TG_THREADS = 320;

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(TG_THREADS, 1)
tma_gemm_kernel(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tb, TmaGemmParams p) {
  using Pipe = TmaPipe<BN, A_MN, B_MN, (BN == 256 ? 4 : 6)>;
  constexpr int S = Pipe::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t full[S], empty[S], tfull[2], tempty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float sbias[2][BN];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 8); }
    mbar_fence_init();
    tma_prefetch_desc(&ta);
    tma_prefetch_desc(&tb);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const int kblocks = (p.K + Pipe::BK - 1) / Pipe::BK;
  const int ntiles = p.tiles_m * p.tiles_n;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int m0 = (tile / p.tiles_n) * Pipe::BM, n0 = (tile % p.tiles_n) * BN;
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], Pipe::STAGE_BYTES);
          uint8_t* st = smem + s * Pipe::STAGE_BYTES;
          Pipe::load_a(st, &full[s], &ta, m0, kb * Pipe::BK, 0);
          Pipe::load_b(st, &full[s], &tb, n0, kb * Pipe::BK, 0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
        const int acc = lt & 1, aph = (lt >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % S, ph = (it / S) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          Pipe::mma_stage(smem + s * Pipe::STAGE_BYTES, tmem + acc * BN, kb == 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull[acc]);
      }
    }
  } else {
    const int quad = warp & 3;                               // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;                        // which half of the tile's columns
    const int et = threadIdx.x - 64;                         // 0..255 among the epilogue threads
    const bool relu = p.flags & DDLS_GEMM_RELU, has_bias = (p.flags & DDLS_GEMM_BIAS) && p.bias != nullptr;
    const bool out_bf16 = p.flags & DDLS_GEMM_OUT_BF16, accumulate = p.flags & DDLS_GEMM_ACCUMULATE;
    int lt = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
      const int m0 = (tile / p.tiles_n) * Pipe::BM, n0 = (tile % p.tiles_n) * BN;
      const int acc = lt & 1, aph = (lt >> 1) & 1;
      if (has_bias) {                                        // sbias[acc] was last read two tiles ago
        if (et < BN) sbias[acc][et] = (n0 + et < p.N) ? __ldg(p.bias + n0 + et) : 0.f;
        named_bar_sync(1, 256);
      }
      mbar_wait(&tfull[acc], aph);
      tc_fence_after();
      const int row = m0 + quad * 32 + lane;
#pragma unroll 1
      for (int c = half * (BN / 2); c < (half + 1) * (BN / 2); c += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN + c, v);
        tmem_ld_wait();
        const int col = n0 + c;
        if (row < p.M && col < p.N) {
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (has_bias) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 bq = *reinterpret_cast<const float4*>(&sbias[acc][c + j]);
              f[j] += bq.x; f[j + 1] += bq.y; f[j + 2] += bq.z; f[j + 3] += bq.w;
            }
          }
          if (relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          if (out_bf16) {
            __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (col + j < p.N) {
                uint4 o;
                o.x = pack_bf16x2(f[j], f[j + 1]); o.y = pack_bf16x2(f[j + 2], f[j + 3]);
                o.z = pack_bf16x2(f[j + 4], f[j + 5]); o.w = pack_bf16x2(f[j + 6], f[j + 7]);
                *reinterpret_cast<uint4*>(dst + j) = o;
              }
            }
          } else {
            float* dst = static_cast<float*>(p.C) + (size_t)row * p.ldc + col;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (col + j < p.N) {
                float4 o = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
                if (accumulate) {
                  const float4 old = *reinterpret_cast<const float4*>(dst + j);
                  o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                }
                *reinterpret_cast<float4*>(dst + j) = o;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 2 * BN);
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

template <int BN, bool A_MN, bool B_MN>
static int launch_tma_gemm(const void* A, const void* B, TmaGemmParams p, int lda, int ldb, cudaStream_t st) {
  using Pipe = TmaPipe<BN, A_MN, B_MN, (BN == 256 ? 4 : 6)>;
  CUtensorMap ta, tb;
  {
    const uint64_t dims[3] = {A_MN ? (uint64_t)p.M : (uint64_t)p.K, A_MN ? (uint64_t)p.K : (uint64_t)p.M, 1};
    const uint64_t strides[2] = {(uint64_t)lda * 2, (uint64_t)lda * 2 * dims[1]};
    const uint32_t box[3] = {64, A_MN ? 64u : 128u, 1};
    if (tmap_bf16_sw128(&ta, A, dims, strides, box)) return 1;
  }
  {
    const uint64_t dims[3] = {B_MN ? (uint64_t)p.N : (uint64_t)p.K, B_MN ? (uint64_t)p.K : (uint64_t)p.N, 1};
    const uint64_t strides[2] = {(uint64_t)ldb * 2, (uint64_t)ldb * 2 * dims[1]};
    const uint32_t box[3] = {64, B_MN ? 64u : (uint32_t)BN, 1};
    if (tmap_bf16_sw128(&tb, B, dims, strides, box)) return 1;
  }
  p.tiles_m = (p.M + 127) / 128;
  p.tiles_n = (p.N + BN - 1) / BN;
  const int ntiles = p.tiles_m * p.tiles_n;
  const int grid = ntiles < sm_count() ? ntiles : sm_count();
  auto kern = tma_gemm_kernel<BN, A_MN, B_MN>;
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Pipe::SMEM_BYTES));
  kern<<<grid, TG_THREADS, Pipe::SMEM_BYTES, st>>>(ta, tb, p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "tma_gemm_kernel launch");
}

template <bool A_MN, bool B_MN>
static int dispatch_tma_bn(const void* A, const void* B, const TmaGemmParams& p, int lda, int ldb, cudaStream_t st) {
  if (p.N > 128) return launch_tma_gemm<256, A_MN, B_MN>(A, B, p, lda, ldb, st);
  return launch_tma_gemm<128, A_MN, B_MN>(A, B, p, lda, ldb, st);
}

bool tma_gemm_applicable(const void* A, const void* B, const void* C, int M, int N, int K, int lda, int ldb, int ldc,
                         unsigned flags) {
  if (flags & (DDLS_GEMM_A_F32 | DDLS_GEMM_B_F32 | DDLS_GEMM_FORCE_V1 | DDLS_GEMM_NO_TMA)) return false;
  if ((flags & DDLS_GEMM_ACCUMULATE) && (flags & DDLS_GEMM_OUT_BF16)) return false;
  const int cal = (flags & DDLS_GEMM_OUT_BF16) ? 8 : 4;
  return lda % 8 == 0 && ldb % 8 == 0 && ldc % cal == 0 && N % 8 == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0 &&
         (reinterpret_cast<uintptr_t>(B) & 15) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 && M > 0 && N > 0 && K > 0;
}

}  // namespace ddls

extern "C" int ddls_gemm_tma(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb,
                             int ldc, int a_mn_major, int b_mn_major, unsigned flags, void* stream) {
  using namespace ddls;
  if (!tma_gemm_applicable(A, B, C, M, N, K, lda, ldb, ldc, flags & ~DDLS_GEMM_NO_TMA)) {
    set_last_error("ddls_gemm_tma: needs bf16 operands, 16-byte aligned bases, lda/ldb/N multiples of 8 "
                   "(M=%d N=%d K=%d lda=%d ldb=%d ldc=%d flags=%u)", M, N, K, lda, ldb, ldc, flags);
    return 1;
  }
  TmaGemmParams p{C, bias, M, N, K, ldc, flags, 0, 0};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a_mn_major) return b_mn_major ? dispatch_tma_bn<true, true>(A, B, p, lda, ldb, st) : dispatch_tma_bn<true, false>(A, B, p, lda, ldb, st);
  return b_mn_major ? dispatch_tma_bn<false, true>(A, B, p, lda, ldb, st) : dispatch_tma_bn<false, false>(A, B, p, lda, ldb, st);
}
