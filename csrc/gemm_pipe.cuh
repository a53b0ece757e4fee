// cp.async-pipelined tcgen05 GEMM mainloop shared by the generic GEMM (csrc/umma_gemm.cu) and the fused
// LSTM step kernels (csrc/lstm_big.cu): 256 threads stream bf16 operand tiles with LDGSTS through a 4-stage ring
// straight into the SWIZZLE_NONE UMMA layouts, one thread issues the MMAs, accumulators end up in TMEM.
#pragma once
#include "common.cuh"

namespace ddls {

constexpr int GEMM_BM = 128;
constexpr int GEMM_KC = 64;       // K elements per pipeline stage
constexpr int V2_THREADS = 256;
constexpr int V2_STAGES = 4;

template <int ROWS>
DDLS_DEVINL void cpasync_kmajor(uint8_t* smem, const __nv_bfloat16* g, int ld, int row0, int nrows, int k0, int kend) {
  constexpr int KCH = GEMM_KC / 8;
  static_assert((ROWS * KCH) % V2_THREADS == 0, "tile must divide evenly over the CTA");
#pragma unroll
  for (int it = 0; it < ROWS * KCH / V2_THREADS; ++it) {
    const int idx = threadIdx.x + it * V2_THREADS;
    const int blk = idx >> 5, lane = idx & 31;
    const int rblk = blk % (ROWS / 8), kblk = blk / (ROWS / 8);
    const int r = rblk * 8 + (lane & 7), kc = kblk * 4 + (lane >> 3);
    const int gr = row0 + r, gk = k0 + kc * 8;
    const bool ok = gr < nrows && gk < kend;
    const int valid = ok ? min(8, kend - gk) : 0;
    cp_async16(smem + kc * (ROWS * 16) + r * 16, ok ? (const void*)(g + (size_t)gr * ld + gk) : (const void*)g, valid * 2);
  }
}
template <int COLS>
DDLS_DEVINL void cpasync_mnmajor(uint8_t* smem, const __nv_bfloat16* g, int ld, int col0, int ncols, int k0, int kend) {
  constexpr int MNG = COLS / 8;
  static_assert((GEMM_KC * MNG) % V2_THREADS == 0, "tile must divide evenly over the CTA");
#pragma unroll
  for (int it = 0; it < GEMM_KC * MNG / V2_THREADS; ++it) {
    const int idx = threadIdx.x + it * V2_THREADS;
    const int blk = idx >> 5, lane = idx & 31;
    const int gblk = blk % (MNG / 4), kblk = blk / (MNG / 4);
    const int mng = gblk * 4 + (lane >> 3), k = kblk * 8 + (lane & 7);
    const int gk = k0 + k, gc = col0 + mng * 8;
    const bool ok = gk < kend && gc < ncols;
    const int valid = ok ? min(8, ncols - gc) : 0;
    cp_async16(smem + (k >> 3) * (COLS * 16) + mng * 128 + (k & 7) * 16, ok ? (const void*)(g + (size_t)gk * ld + gc) : (const void*)g,
               valid * 2);
  }
}


// Runs the whole K loop for one [128 x BN] output tile.  On return every thread may read the accumulator at
// the returned TMEM address (columns [0, BN)); all pipeline shared memory is idle and may be reused.
// The caller deallocates TMEM (warp 0, `gemm_pipe_tmem_cols<BN>()` columns).
template <int BN>
__host__ __device__ constexpr uint32_t gemm_pipe_tmem_cols() { return BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256; }
template <int BN>
__host__ __device__ constexpr int gemm_pipe_smem_bytes() { return V2_STAGES * (GEMM_BM * GEMM_KC * 2 + BN * GEMM_KC * 2); }

template <int BN, bool A_MN, bool B_MN>
DDLS_DEVINL uint32_t gemm_pipe_run(uint8_t* smem, uint64_t* mma_done, uint32_t* tmem_slot, const __nv_bfloat16* A, int lda,
                                   int M, int m0, const __nv_bfloat16* B, int ldb, int N, int n0, int kbeg, int kend) {
  constexpr int A_BYTES = GEMM_BM * GEMM_KC * 2, B_BYTES = BN * GEMM_KC * 2, STAGE = A_BYTES + B_BYTES;
  constexpr uint32_t TM_COLS = gemm_pipe_tmem_cols<BN>();
  const int warp = threadIdx.x >> 5;
  uint32_t& tmem_base_s = *tmem_slot;
  if (warp == 0) tmem_alloc(&tmem_base_s, TM_COLS);
  if (threadIdx.x == 32) {
    for (int s = 0; s < V2_STAGES; ++s) mbar_init(&mma_done[s], 1);
    mbar_fence_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;
  constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN, B_MN);
  const int nkb = (kend - kbeg + GEMM_KC - 1) / GEMM_KC;
  pdl_wait();                             // TMEM / barrier set-up above overlapped the previous kernel's tail

  auto issue_loads = [&](int kb) {
    uint8_t* sa = smem + (kb % V2_STAGES) * STAGE;
    uint8_t* sb = sa + A_BYTES;
    const int k0 = kbeg + kb * GEMM_KC;
    if (A_MN) cpasync_mnmajor<GEMM_BM>(sa, A, lda, m0, M, k0, kend);
    else      cpasync_kmajor<GEMM_BM>(sa, A, lda, m0, M, k0, kend);
    if (B_MN) cpasync_mnmajor<BN>(sb, B, ldb, n0, N, k0, kend);
    else      cpasync_kmajor<BN>(sb, B, ldb, n0, N, k0, kend);
  };
  // prologue: fill STAGES-1 slots (always commit so group accounting stays uniform)
  for (int kb = 0; kb < V2_STAGES - 1; ++kb) {
    if (kb < nkb) issue_loads(kb);
    cp_async_commit();
  }
  uint32_t waits[V2_STAGES] = {};
  for (int kb = 0; kb < nkb; ++kb) {
    const int s = kb % V2_STAGES;
    // refill the slot that k-block kb-1 used, once its MMAs have drained
    const int nxt = kb + V2_STAGES - 1;
    if (nxt < nkb) {
      const int ns = nxt % V2_STAGES;
      if (kb >= 1) { mbar_wait(&mma_done[ns], waits[ns] & 1); ++waits[ns]; }
      issue_loads(nxt);
    }
    cp_async_commit();
    cp_async_wait<V2_STAGES - 1>();       // k-block kb has landed (for this thread)
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      const uint32_t a_base = smem_u32(smem + s * STAGE), b_base = a_base + A_BYTES;
#pragma unroll
      for (int ks = 0; ks < GEMM_KC / 16; ++ks)
        umma_bf16(tmem_d, umma_desc(a_base + ks * 2 * (GEMM_BM * 16u), GEMM_BM * 16u, 128u),
                  umma_desc(b_base + ks * 2 * (BN * 16u), BN * 16u, 128u), idesc, (kb | ks) != 0);
      umma_commit(&mma_done[s]);
    }
  }
  cp_async_wait<0>();
  for (int s = 0; s < V2_STAGES; ++s) {
    const uint32_t commits = nkb > s ? (nkb - s + V2_STAGES - 1) / V2_STAGES : 0;
    while (waits[s] < commits) { mbar_wait(&mma_done[s], waits[s] & 1); ++waits[s]; }
  }
  tc_fence_after();
  __syncthreads();    // every pipeline slot is idle: reuse it as the epilogue transpose buffer

  return tmem_d;
}

}  // namespace ddls
