// General bf16 GEMM on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM).
//
//   C[M,N] (+)= A · Bᵀ-style product with selectable operand majors:
//     A_MN = false: A is [M,K] row-major (K contiguous)      A_MN = true: A is [K,M] row-major (M contiguous)
//     B_MN = false: B is [N,K] row-major (K contiguous)      B_MN = true: B is [K,N] row-major (N contiguous)
//
// Operands are staged global -> shared by the CTA's threads straight into the UMMA SWIZZLE_NONE
// canonical layouts (no swizzle ⇒ a 16-byte vector per thread lands conflict-free), double buffered
// against the asynchronous MMAs with mbarriers; one elected thread issues tcgen05.mma; the
// epilogue reads TMEM with tcgen05.ld and fuses bias / ReLU / bf16 cast / split-K reduction.
//
// Used for: Dense layers of arbitrary size, the stacked-LSTM input projection, weight-gradient GEMMs
// (MN-major operands), and as the validation vehicle for the descriptor encodings every other
// tcgen05 kernel in this library relies on.  The reference has no GEMM of its own (Keras→TF Eigen on
// CPU).
#include <cstdio>

#include "common.cuh"
#include "gemm_pipe.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int GEMM_THREADS = 128;

struct GemmParams {
  const void* A;      // bf16, or fp32 when DDLS_GEMM_A_F32 (converted while staging)
  const void* B;      // bf16, or fp32 when DDLS_GEMM_B_F32
  void* C;
  const float* bias;
  int M, N, K;
  int lda, ldb, ldc;
  int k_per_split;
  uint32_t flags;     // DDLS_GEMM_* bits
  float* workspace;   // split-K partials [splits][M][N] fp32 (DDLS_GEMM_SPLITK_WS)
  uint32_t epi;       // epilogue flags applied by the in-kernel split-K reduction (SPLITK_FUSED)
};

// In-kernel split-K reduction (v2 kernel): when every CTA of the launch is resident at once (<= one CTA per SM), the CTAs
// of an output tile write their partial tiles, meet at a counter and each reduces 1/splits of the tile's rows with the
// fused epilogue - no second kernel (the MNIST Dense(225) forward GEMM: 8.1 + 7.4 us as two launches).  The counters are
// self-resetting: {arrivals, departures} per output tile; the last CTA to leave zeroes both.  Like the split-K workspace
// they serialise split-K GEMMs of one device (launch them on one stream).
constexpr uint32_t SPLITK_FUSED = 1u << 20;
constexpr int SPLITK_MAX_TILES = 64;
static __device__ unsigned g_splitk_sync[2 * SPLITK_MAX_TILES];

// 8 consecutive elements starting at `src` (bf16 or fp32 in memory) -> one 16-byte bf16 vector.
// `n_valid` (1..8) elements are in range, the rest are zero-filled; `vec` says 16-byte vector loads are legal.
DDLS_DEVINL uint4 load8_as_bf16(const void* base, size_t elem_off, int n_valid, bool is_f32, bool vec) {
  if (!is_f32) {
    const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(base) + elem_off;
    if (vec && n_valid == 8) return *reinterpret_cast<const uint4*>(src);
    __nv_bfloat16 t[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) t[e] = (e < n_valid) ? src[e] : __float2bfloat16(0.f);
    return *reinterpret_cast<uint4*>(t);
  }
  const float* src = static_cast<const float*>(base) + elem_off;
  float f[8];
  if (vec && n_valid == 8) {
    const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = (e < n_valid) ? src[e] : 0.f;
  }
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}

// Stage a [rows x GEMM_KC] K-major tile: smem layout [KC/8][rows][8].
template <int ROWS>
DDLS_DEVINL void stage_kmajor(uint8_t* smem, const void* g, int ld, int row0, int nrows, int k0, int kend,
                              bool vec_ok, bool is_f32) {
  // warp-level mapping: lane = (kc_local = lane/8, r_local = lane%8) ⇒ 8 rows x 64 B per warp access
  constexpr int KCH = GEMM_KC / 8;
  for (int idx = threadIdx.x; idx < ROWS * KCH; idx += GEMM_THREADS) {
    int blk = idx >> 5, lane = idx & 31;
    int rblk = blk % (ROWS / 8), kblk = blk / (ROWS / 8);
    int r = rblk * 8 + (lane & 7);
    int kc = kblk * 4 + (lane >> 3);
    int gr = row0 + r, gk = k0 + kc * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (gr < nrows && gk < kend) v = load8_as_bf16(g, (size_t)gr * ld + gk, min(8, kend - gk), is_f32, vec_ok);
    *reinterpret_cast<uint4*>(smem + kc * (ROWS * 16) + r * 16) = v;
  }
}

// Stage a [GEMM_KC x COLS] tile given MN-contiguous global data: smem layout [KC/8][COLS/8][8 k][8 mn].
template <int COLS>
DDLS_DEVINL void stage_mnmajor(uint8_t* smem, const void* g, int ld, int col0, int ncols, int k0, int kend,
                               bool vec_ok, bool is_f32) {
  // lane = (k_local = lane%8, mng_local = lane/8): quarter-warps write 128 contiguous bytes
  constexpr int MNG = COLS / 8;
  for (int idx = threadIdx.x; idx < GEMM_KC * MNG; idx += GEMM_THREADS) {
    int blk = idx >> 5, lane = idx & 31;
    int gblk = blk % (MNG / 4), kblk = blk / (MNG / 4);
    int mng = gblk * 4 + (lane >> 3);
    int k = kblk * 8 + (lane & 7);
    int gk = k0 + k, gc = col0 + mng * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (gk < kend && gc < ncols) v = load8_as_bf16(g, (size_t)gk * ld + gc, min(8, ncols - gc), is_f32, vec_ok);
    *reinterpret_cast<uint4*>(smem + (k >> 3) * (COLS * 16) + mng * 128 + (k & 7) * 16) = v;
  }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS) umma_gemm_kernel(GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int A_BYTES = GEMM_BM * GEMM_KC * 2;
  constexpr int B_BYTES = BN * GEMM_KC * 2;
  uint8_t* sA[2] = {smem, smem + A_BYTES};
  uint8_t* sB[2] = {smem + 2 * A_BYTES, smem + 2 * A_BYTES + B_BYTES};
  __shared__ uint64_t mma_done[2];
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t TM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * GEMM_BM, n0 = blockIdx.y * BN;
  const int kbeg = blockIdx.z * p.k_per_split;
  const int kend = min(p.K, kbeg + p.k_per_split);

  if (warp == 0) tmem_alloc(&tmem_base_s, TM_COLS);
  if (threadIdx.x == 32) {
    mbar_init(&mma_done[0], 1);
    mbar_init(&mma_done[1], 1);
    mbar_fence_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  const bool a_f32 = p.flags & DDLS_GEMM_A_F32, b_f32 = p.flags & DDLS_GEMM_B_F32;
  const bool a_vec = ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0) && (p.lda % (a_f32 ? 4 : 8) == 0);
  const bool b_vec = ((reinterpret_cast<uintptr_t>(p.B) & 15) == 0) && (p.ldb % (b_f32 ? 4 : 8) == 0);
  constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN, B_MN);

  const int nkb = (kend - kbeg + GEMM_KC - 1) / GEMM_KC;
  uint32_t waits[2] = {0, 0};  // completed-phase counters per stage (parity = count & 1)
  for (int kb = 0; kb < nkb; ++kb) {
    const int s = kb & 1;
    if (kb >= 2) {  // the MMAs that read this stage two iterations ago must have drained
      mbar_wait(&mma_done[s], waits[s] & 1);
      ++waits[s];
    }
    const int k0 = kbeg + kb * GEMM_KC;
    if (A_MN) stage_mnmajor<GEMM_BM>(sA[s], p.A, p.lda, m0, p.M, k0, kend, a_vec, a_f32);
    else      stage_kmajor<GEMM_BM>(sA[s], p.A, p.lda, m0, p.M, k0, kend, a_vec, a_f32);
    if (B_MN) stage_mnmajor<BN>(sB[s], p.B, p.ldb, n0, p.N, k0, kend, b_vec, b_f32);
    else      stage_kmajor<BN>(sB[s], p.B, p.ldb, n0, p.N, k0, kend, b_vec, b_f32);
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      const uint32_t a_base = smem_u32(sA[s]), b_base = smem_u32(sB[s]);
      // both staged layouts: K-group stride (LBO) = extent*16 B, 8-row / 8-element group stride (SBO) = 128 B
      const uint32_t a_lbo = GEMM_BM * 16u, a_sbo = 128u;
      const uint32_t b_lbo = BN * 16u, b_sbo = 128u;
#pragma unroll
      for (int ks = 0; ks < GEMM_KC / 16; ++ks) {
        // one UMMA consumes K=16 = two 8-wide K groups; both layouts advance by 2 K-groups per step
        const uint32_t a_adv = ks * 2 * (GEMM_BM * 16u);
        const uint32_t b_adv = ks * 2 * (BN * 16u);
        umma_bf16(tmem_d, umma_desc(a_base + a_adv, a_lbo, a_sbo), umma_desc(b_base + b_adv, b_lbo, b_sbo), idesc,
                  (kb | ks) != 0);
      }
      umma_commit(&mma_done[s]);
    }
  }
  // drain every outstanding commit (stage s received ceil((nkb - s) / 2) commits)
  for (int s = 0; s < 2; ++s) {
    const uint32_t commits = (nkb + 1 - s) / 2;
    while (waits[s] < commits) {
      mbar_wait(&mma_done[s], waits[s] & 1);
      ++waits[s];
    }
  }
  tc_fence_after();

  // ---- epilogue: thread (warp w, lane l) owns output row m0 + 32*w + l ---------------------
  const int row = m0 + warp * 32 + lane;
  const uint32_t t_row = tmem_d + (static_cast<uint32_t>(warp * 32) << 16);
  const bool relu = p.flags & DDLS_GEMM_RELU, has_bias = (p.flags & DDLS_GEMM_BIAS) && blockIdx.z == 0;
  const bool out_bf16 = p.flags & DDLS_GEMM_OUT_BF16, atomic = p.flags & DDLS_GEMM_ATOMIC_ADD;
  if (p.flags & DDLS_GEMM_SPLITK_WS) {
    // raw fp32 partial tile -> workspace[split][row][col]; bias/ReLU/cast happen in splitk_finish_kernel
    float* ws = p.workspace + (size_t)blockIdx.z * p.M * p.N;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(t_row + c0, v);
      tmem_ld_wait();
      if (row < p.M) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (n0 + c0 + j < p.N) ws[(size_t)row * p.N + n0 + c0 + j] = __uint_as_float(v[j]);
      }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_d, TM_COLS);
    return;
  }
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(t_row + c0, v);
    tmem_ld_wait();
    if (row < p.M) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int col = n0 + c0 + j;
        if (col < p.N) {
          float x = __uint_as_float(v[j]);
          if (has_bias) x += p.bias[col];
          if (relu) x = fmaxf(x, 0.f);
          const size_t o = (size_t)row * p.ldc + col;
          if (atomic) reinterpret_cast<float*>(p.C)[o] += x;   // one CTA owns each output when splits == 1
          else if (out_bf16) reinterpret_cast<__nv_bfloat16*>(p.C)[o] = __float2bfloat16(x);
          else reinterpret_cast<float*>(p.C)[o] = x;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, TM_COLS);
}

// Sum split-K partials and apply the fused epilogue.  One thread per 4 consecutive outputs of a row.
__global__ void __launch_bounds__(256) splitk_finish_kernel(const float* __restrict__ ws, void* __restrict__ C,
                                                            const float* __restrict__ bias, int M, int N, int ldc,
                                                            int splits, uint32_t flags) {
  pdl_wait();
  const size_t total = (size_t)M * N;
  const bool relu = flags & DDLS_GEMM_RELU, has_bias = flags & DDLS_GEMM_BIAS, out_bf16 = flags & DDLS_GEMM_OUT_BF16;
  const bool accum = flags & DDLS_GEMM_ACCUMULATE;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
#pragma unroll 8
    for (int s = 0; s < splits; ++s) acc += ws[(size_t)s * total + i];
    const int r = (int)(i / N), c = (int)(i - (size_t)r * N);
    if (has_bias) acc += bias[c];
    if (relu) acc = fmaxf(acc, 0.f);
    const size_t o = (size_t)r * ldc + c;
    if (out_bf16) static_cast<__nv_bfloat16*>(C)[o] = __float2bfloat16(acc);
    else if (accum) static_cast<float*>(C)[o] += acc;
    else static_cast<float*>(C)[o] = acc;
  }
}

// =================================================================================================
// v2 mainloop: bf16 operands with 16-byte aligned rows are streamed with cp.async (LDGSTS) through a
// 4-stage ring straight into the canonical layouts — three k-blocks of loads are always in flight, so the
// loop is throughput- rather than latency-bound — and the epilogue transposes through shared memory so
// global stores are coalesced along the output row.
// =================================================================================================
template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(V2_THREADS, 1) umma_gemm_v2_kernel(GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t mma_done[V2_STAGES];
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t TM_COLS = gemm_pipe_tmem_cols<BN>();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * GEMM_BM, n0 = blockIdx.y * BN;
  const int kbeg = blockIdx.z * p.k_per_split, kend = min(p.K, kbeg + p.k_per_split);
  const uint32_t tmem_d = gemm_pipe_run<BN, A_MN, B_MN>(smem, mma_done, &tmem_base_s, static_cast<const __nv_bfloat16*>(p.A),
                                                         p.lda, p.M, m0, static_cast<const __nv_bfloat16*>(p.B), p.ldb, p.N,
                                                         n0, kbeg, kend);

  // ---- epilogue: TMEM -> registers -> smem [128 rows][BN+1] fp32 chunks of 32 columns -> coalesced rows
  const bool relu = p.flags & DDLS_GEMM_RELU, has_bias = (p.flags & DDLS_GEMM_BIAS);
  const bool out_bf16 = p.flags & DDLS_GEMM_OUT_BF16, accum = p.flags & DDLS_GEMM_ATOMIC_ADD;
  const bool to_ws = p.flags & DDLS_GEMM_SPLITK_WS;
  float* sT = reinterpret_cast<float*>(smem);                  // [128][33] per 32-column chunk
  const int q4 = warp & 3, half = warp >> 2;
  for (int c0 = 0; c0 < BN; c0 += 32) {
    if (n0 + c0 >= p.N) break;
    uint32_t v[16];
    tmem_ld16(tmem_d + (static_cast<uint32_t>(q4 * 32) << 16) + c0 + half * 16, v);
    tmem_ld_wait();
    const int r = q4 * 32 + lane;
#pragma unroll
    for (int j = 0; j < 16; ++j) sT[r * 33 + half * 16 + j] = __uint_as_float(v[j]);
    __syncthreads();
    // 8 warps x 16 rows each; lane = column
    for (int rr = warp * 16; rr < warp * 16 + 16; ++rr) {
      const int row = m0 + rr, col = n0 + c0 + lane;
      if (row < p.M && col < p.N) {
        float x = sT[rr * 33 + lane];
        if (to_ws) {
          p.workspace[((size_t)blockIdx.z * p.M + row) * p.N + col] = x;
        } else {
          if (has_bias) x += p.bias[col];
          if (relu) x = fmaxf(x, 0.f);
          const size_t o = (size_t)row * p.ldc + col;
          if (out_bf16) reinterpret_cast<__nv_bfloat16*>(p.C)[o] = __float2bfloat16(x);
          else if (accum) reinterpret_cast<float*>(p.C)[o] += x;
          else reinterpret_cast<float*>(p.C)[o] = x;
        }
      }
    }
    __syncthreads();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, TM_COLS);
  if (p.flags & SPLITK_FUSED) {
    unsigned* arrive = g_splitk_sync + 2 * (blockIdx.x * gridDim.y + blockIdx.y);
    unsigned* depart = arrive + 1;
    __threadfence();                                   // my partial tile is visible device-wide
    __syncthreads();
    if (threadIdx.x == 0) {
      atomicAdd(arrive, 1u);
      unsigned long long t0 = 0, now = 0;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      unsigned spins = 0;
      while (*reinterpret_cast<volatile unsigned*>(arrive) < gridDim.z) {
        if ((++spins & 1023u) == 0) {
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
          if (now - t0 > 2000000000ull) {      // 2 s: the CTAs of this launch are not all resident (never expected)
            printf("ddls gemm: split-K rendezvous timed out (tile %d,%d split %d saw %u of %u)\n", blockIdx.x, blockIdx.y,
                   blockIdx.z, *reinterpret_cast<volatile unsigned*>(arrive), gridDim.z);
            __trap();
          }
        }
      }
      __threadfence();
    }
    __syncthreads();
    const bool relu2 = p.epi & DDLS_GEMM_RELU, bias2 = p.epi & DDLS_GEMM_BIAS, bf2 = p.epi & DDLS_GEMM_OUT_BF16;
    const bool acc2 = p.epi & DDLS_GEMM_ACCUMULATE;
    const int rows_per = (GEMM_BM + gridDim.z - 1) / gridDim.z;
    const int r0 = blockIdx.z * rows_per, r1 = min(GEMM_BM, r0 + rows_per);
    const int splits = gridDim.z;
    for (int idx = threadIdx.x; idx < (r1 - r0) * BN; idx += blockDim.x) {
      const int row = m0 + r0 + idx / BN, col = n0 + idx % BN;
      if (row < p.M && col < p.N) {
        float acc = 0.f;
#pragma unroll 6
        for (int sidx = 0; sidx < splits; ++sidx) acc += __ldcg(p.workspace + ((size_t)sidx * p.M + row) * p.N + col);
        if (bias2) acc += p.bias[col];
        if (relu2) acc = fmaxf(acc, 0.f);
        const size_t o = (size_t)row * p.ldc + col;
        if (bf2) reinterpret_cast<__nv_bfloat16*>(p.C)[o] = __float2bfloat16(acc);
        else if (acc2) reinterpret_cast<float*>(p.C)[o] += acc;
        else reinterpret_cast<float*>(p.C)[o] = acc;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      if (atomicAdd(depart, 1u) == gridDim.z - 1) { *arrive = 0; *depart = 0; __threadfence(); }
    }
  }
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm_v2(const GemmParams& p, int splits, cudaStream_t st) {
  auto kern = umma_gemm_v2_kernel<BN, A_MN, B_MN>;
  const int smem = V2_STAGES * (GEMM_BM * GEMM_KC * 2 + BN * GEMM_KC * 2);
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid((p.M + GEMM_BM - 1) / GEMM_BM, (p.N + BN - 1) / BN, splits);
  GemmParams q = p;
  if (splits > 1 && (int)(grid.x * grid.y) <= SPLITK_MAX_TILES && (int)(grid.x * grid.y * grid.z) <= sm_count())
    q.flags |= SPLITK_FUSED;                  // every CTA resident at once: reduce in the kernel, no finish launch
  DDLS_CUDA_OK(launch_k(kern, grid, dim3(V2_THREADS), smem, st, q));
  count_launch(1);
  int rc = check_cuda(cudaGetLastError(), "umma_gemm_v2 launch");
  return rc ? rc : ((q.flags & SPLITK_FUSED) ? -1 : 0);     // -1: done, including the split-K reduction
}

template <bool A_MN, bool B_MN>
static int dispatch_bn_v2(const GemmParams& p, int splits, cudaStream_t st) {
  if (p.N <= 32) return launch_gemm_v2<32, A_MN, B_MN>(p, splits, st);
  if (p.N <= 64) return launch_gemm_v2<64, A_MN, B_MN>(p, splits, st);
  // These problems are small (the MNIST Dense(225) layer: 0.27 GFLOP per GEMM) and latency-bound: the widest tile that
  // still spreads the work over ~100 CTAs wins, not the one with the best operand reuse.
  const int tm = (p.M + GEMM_BM - 1) / GEMM_BM;
  auto ctas = [&](int bn) { return tm * ((p.N + bn - 1) / bn) * splits; };
  if (p.N > 128 && ctas(256) >= 96) return launch_gemm_v2<256, A_MN, B_MN>(p, splits, st);
  if (p.N > 64 && ctas(128) >= 96) return launch_gemm_v2<128, A_MN, B_MN>(p, splits, st);
  return launch_gemm_v2<64, A_MN, B_MN>(p, splits, st);
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const GemmParams& p, int splits, cudaStream_t st) {
  auto kern = umma_gemm_kernel<BN, A_MN, B_MN>;
  const int smem = 2 * (GEMM_BM * GEMM_KC * 2) + 2 * (BN * GEMM_KC * 2);
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid((p.M + GEMM_BM - 1) / GEMM_BM, (p.N + BN - 1) / BN, splits);
  kern<<<grid, GEMM_THREADS, smem, st>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "umma_gemm launch");
}

template <bool A_MN, bool B_MN>
static int dispatch_bn(const GemmParams& p, int splits, cudaStream_t st) {
  if (p.N <= 32) return launch_gemm<32, A_MN, B_MN>(p, splits, st);
  if (p.N <= 64) return launch_gemm<64, A_MN, B_MN>(p, splits, st);
  if (p.N <= 128 || (p.N % 256 != 0 && p.N % 128 == 0)) return launch_gemm<128, A_MN, B_MN>(p, splits, st);
  return launch_gemm<256, A_MN, B_MN>(p, splits, st);
}

}  // namespace ddls

namespace ddls {
bool tma_gemm_applicable(const void* A, const void* B, const void* C, int M, int N, int K, int lda, int ldb, int ldc,
                         unsigned flags);
}
extern "C" int ddls_gemm_bf16(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda,
                              int ldb, int ldc, int a_mn_major, int b_mn_major, int split_k, unsigned flags,
                              void* workspace, size_t workspace_bytes, void* stream) {
  using namespace ddls;
  if (M <= 0 || N <= 0 || K <= 0) { set_last_error("ddls_gemm_bf16: empty problem"); return 1; }
  GemmParams p;
  p.A = A; p.B = B; p.C = C; p.bias = bias;
  p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldc = ldc; p.workspace = static_cast<float*>(workspace);
  p.flags = flags & ~(DDLS_GEMM_ATOMIC_ADD | DDLS_GEMM_SPLITK_WS | SPLITK_FUSED);
  p.epi = 0;
  int splits = split_k < 1 ? 1 : split_k;
  int kps = (K + splits - 1) / splits;
  kps = (kps + GEMM_KC - 1) / GEMM_KC * GEMM_KC;
  splits = (K + kps - 1) / kps;
  p.k_per_split = kps;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (splits == 1 && M >= 256 && tma_gemm_applicable(A, B, C, M, N, K, lda, ldb, ldc, flags))
    return ddls_gemm_tma(A, B, C, bias, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, flags, stream);
  const unsigned epi = flags & (DDLS_GEMM_RELU | DDLS_GEMM_BIAS | DDLS_GEMM_OUT_BF16 | DDLS_GEMM_ACCUMULATE);
  if (splits > 1) {
    const size_t need = (size_t)splits * M * N * sizeof(float);
    if (workspace == nullptr || workspace_bytes < need) {
      set_last_error("ddls_gemm_bf16: split_k=%d needs a %zu-byte workspace (got %zu)", splits, need, workspace_bytes);
      return 1;
    }
    p.flags = (p.flags & ~epi) | DDLS_GEMM_SPLITK_WS;
    p.epi = epi;
  } else if (flags & DDLS_GEMM_ACCUMULATE) {
    p.flags |= DDLS_GEMM_ATOMIC_ADD;   // single split accumulating into C
  }
  // v2 (cp.async pipeline) needs bf16 operands whose rows are 16-byte aligned
  const bool v2 = !(flags & (DDLS_GEMM_A_F32 | DDLS_GEMM_B_F32)) && lda % 8 == 0 && ldb % 8 == 0 &&
                  (reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 &&
                  !(flags & DDLS_GEMM_FORCE_V1);
  int rc;
  if (v2) {
    if (a_mn_major) rc = b_mn_major ? dispatch_bn_v2<true, true>(p, splits, st) : dispatch_bn_v2<true, false>(p, splits, st);
    else rc = b_mn_major ? dispatch_bn_v2<false, true>(p, splits, st) : dispatch_bn_v2<false, false>(p, splits, st);
  } else {
    if (a_mn_major) rc = b_mn_major ? dispatch_bn<true, true>(p, splits, st) : dispatch_bn<true, false>(p, splits, st);
    else rc = b_mn_major ? dispatch_bn<false, true>(p, splits, st) : dispatch_bn<false, false>(p, splits, st);
  }
  if (rc == -1) return 0;                     // the kernel reduced its own split-K partials
  if (rc || splits == 1) return rc;
  const size_t total = (size_t)M * N;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  DDLS_CUDA_OK(launch_k(splitk_finish_kernel, dim3(blocks), dim3(256), 0, st, p.workspace, C, bias, M, N, ldc, splits, epi));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "splitk_finish_kernel launch");
}
