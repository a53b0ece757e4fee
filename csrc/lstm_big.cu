// Large-hidden-size LSTM (the scaled stacked forecaster of BASELINE.json config 4: hidden 1024, seq 168,
// batch 512 per GPU; K10 of SURVEY.md §2.5).  Same Keras LSTM math as csrc/lstm.cu (gates i,f,c,o, hard_sigmoid,
// /root/reference/ddl_nyiso_aztk.py:249-251), but H x 4H no longer fits one SM, so the recurrence is one fused
// tcgen05 GEMM launch per time step over the whole GPU (4 batch tiles x 4H/128 column tiles = 128 CTAs), all
// steps of a training iteration captured in one CUDA graph:
//
//   forward step  : Z = h_{t-1} · U  (cp.async-pipelined UMMA mainloop, csrc/gemm_pipe.cuh) with the gate
//                   math in the epilogue — adds the hoisted input projection (or the rank-1 input term) and the
//                   bias, applies the activations, updates c, writes h_t (bf16, next step's A operand) and the
//                   post-activation gates for BPTT.  Nothing of Z ever reaches memory.
//   backward step : dh_{t-1} = dZ_t · Uᵀ with the gate-derivative math of step t-1 in the epilogue, which
//                   writes dZ_{t-1} (bf16, next launch's A operand and the weight-gradient GEMM's operand).
//
// To let one epilogue thread see all four gates of a unit inside a 128-column tile, every 4H-wide tensor lives
// in a permuted column order ("tile-gate-unit"): column n = (unit/32)*128 + gate*32 + unit%32.  Weights are
// re-laid-out (and cast to bf16) once per optimizer step by perm_cast; gradients are un-permuted on the way out.
// The input projection of layer >= 2 over all T, the data gradient towards the layer below and the weight
// gradients are plain large GEMMs (csrc/umma_gemm.cu).
#include "gemm_pipe.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

DDLS_DEVINL int perm_col(int keras_col, int H) {           // Keras column g*H + unit -> permuted column
  const int g = keras_col / H, unit = keras_col - g * H;
  return (unit >> 5) * 128 + g * 32 + (unit & 31);
}

// src fp32 [rows][4H] (Keras) -> dst bf16 [rows][4H] (permuted columns)
__global__ void __launch_bounds__(256) perm_cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                        int rows, int H) {
  const int G = 4 * H;
  const size_t total = (size_t)rows * G;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / G), n = (int)(i - (size_t)r * G);          // n = permuted column (coalesced writes)
    const int ub = n >> 7, g = (n >> 5) & 3, u = n & 31;
    dst[i] = __float2bfloat16(src[(size_t)r * G + g * H + ub * 32 + u]);
  }
}

// vector version: fp32 [4H] Keras -> fp32 [4H] permuted
__global__ void __launch_bounds__(256) perm_vec_kernel(const float* __restrict__ src, float* __restrict__ dst, int H) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < 4 * H) { const int ub = n >> 7, g = (n >> 5) & 3, u = n & 31; dst[n] = src[g * H + ub * 32 + u]; }
}

// dst fp32 [rows][4H] (Keras) (+)= src fp32 [rows][4H] (permuted)
__global__ void __launch_bounds__(256) unperm_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int H,
                                                     int accumulate) {
  const int G = 4 * H;
  const size_t total = (size_t)rows * G;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / G), n = (int)(i - (size_t)r * G);
    const int ub = n >> 7, g = (n >> 5) & 3, u = n & 31;
    float* d = dst + (size_t)r * G + g * H + ub * 32 + u;
    *d = accumulate ? (*d + src[i]) : src[i];
  }
}

// out_plain[keras col] += sum_r X[r][perm col];  out_w[keras col] += sum_r w[r] * X[r][perm col]   (either may be null)
// X bf16 [R][4H] permuted.  One pass over X computes both: a CTA owns 512 columns x 256 rows, a thread 8 columns
// (16-byte loads) of every 4th row; partials meet in shared memory and leave with one atomic per column.
constexpr int CS_ROWS = 256;
__global__ void __launch_bounds__(256) colsum_perm_kernel(const __nv_bfloat16* __restrict__ X, const float* __restrict__ w,
                                                          float* __restrict__ out_plain, float* __restrict__ out_w, int R,
                                                          int H) {
  __shared__ float red[2][4][512];
  const int G = 4 * H;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int col = blockIdx.x * 512 + tx * 8;
  const int r0 = blockIdx.y * CS_ROWS, r1 = min(R, r0 + CS_ROWS);
  float a[8], aw[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = 0.f; aw[j] = 0.f; }
  if (col < G) {
#pragma unroll 4
    for (int r = r0 + ty; r < r1; r += 4) {
      const uint4 q = *reinterpret_cast<const uint4*>(X + (size_t)r * G + col);
      const float wv = (w != nullptr) ? __ldg(w + r) : 0.f;
      const uint32_t u[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float lo = __uint_as_float(u[i] << 16), hi = __uint_as_float(u[i] & 0xFFFF0000u);
        a[2 * i] += lo; a[2 * i + 1] += hi;
        aw[2 * i] = fmaf(wv, lo, aw[2 * i]); aw[2 * i + 1] = fmaf(wv, hi, aw[2 * i + 1]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { red[0][ty][tx * 8 + j] = a[j]; red[1][ty][tx * 8 + j] = aw[j]; }
  __syncthreads();
  for (int c = threadIdx.x; c < 512; c += 256) {
    const int n = blockIdx.x * 512 + c;
    if (n >= G) continue;
    const int ub = n >> 7, g = (n >> 5) & 3, u = n & 31;
    const int k = g * H + ub * 32 + u;
    if (out_plain != nullptr) atomicAdd(out_plain + k, red[0][0][c] + red[0][1][c] + red[0][2][c] + red[0][3][c]);
    if (out_w != nullptr) atomicAdd(out_w + k, red[1][0][c] + red[1][1][c] + red[1][2][c] + red[1][3][c]);
  }
}

struct BigFwd {
  const __nv_bfloat16* h_prev;   // [B][H]   A operand
  const __nv_bfloat16* u_perm;   // [H][4H]  B operand (n contiguous)
  const __nv_bfloat16* zx;       // [B][4H] permuted input projection of this step, or nullptr
  const float* x;                // [B] input scalars of this step (input dim 1), used when zx == nullptr
  int x_stride;                  // stride between batch rows of x (= T for a [B,T] tensor)
  const float* wih_perm;         // [4H] permuted input kernel (with x)
  const float* bias_perm;        // [4H]
  const float* c_prev;           // [B][H]
  float* c_next;                 // [B][H]
  __nv_bfloat16* h_next;         // [B][H]
  float* h_out_f32;              // [B][H] or nullptr
  __nv_bfloat16* gates;          // [B][4H] permuted post-activation gates, or nullptr
  int B, H;
};

__global__ void __launch_bounds__(V2_THREADS, 1) lstm_big_fwd_kernel(BigFwd p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t mma_done[V2_STAGES];
  __shared__ uint32_t tmem_slot;
  constexpr int BN = 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * GEMM_BM, n0 = blockIdx.y * BN;
  const int G = 4 * p.H;
  const uint32_t tmem = gemm_pipe_run<BN, false, true>(smem, mma_done, &tmem_slot, p.h_prev, p.H, p.B, m0, p.u_perm, G, G, n0,
                                                        0, p.H);
  // accumulator -> shared memory, one [128][33] fp32 plane per gate
  float* sT = reinterpret_cast<float*>(smem);
  const int q4 = warp & 3, half = warp >> 2;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t v[16];
    tmem_ld16(tmem + (static_cast<uint32_t>(q4 * 32) << 16) + g * 32 + half * 16, v);
    tmem_ld_wait();
    const int r = q4 * 32 + lane;
#pragma unroll
    for (int j = 0; j < 16; ++j) sT[(g * 128 + r) * 33 + half * 16 + j] = __uint_as_float(v[j]);
  }
  __syncthreads();
  const int unit = (n0 >> 7) * 32 + lane;
  float wi[4], bi[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    bi[g] = p.bias_perm[n0 + g * 32 + lane];
    wi[g] = (p.zx == nullptr) ? p.wih_perm[n0 + g * 32 + lane] : 0.f;
  }
  for (int rr = warp * 16; rr < warp * 16 + 16; ++rr) {
    const int b = m0 + rr;
    if (b >= p.B) break;
    float z[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) z[g] = sT[(g * 128 + rr) * 33 + lane] + bi[g];
    if (p.zx != nullptr) {
      const __nv_bfloat16* zr = p.zx + (size_t)b * G + n0 + lane;
#pragma unroll
      for (int g = 0; g < 4; ++g) z[g] += __bfloat162float(zr[g * 32]);
    } else {
      const float xv = p.x[(size_t)b * p.x_stride];
#pragma unroll
      for (int g = 0; g < 4; ++g) z[g] = fmaf(xv, wi[g], z[g]);
    }
    const float gi = __saturatef(fmaf(0.2f, z[0], 0.5f)), gf = __saturatef(fmaf(0.2f, z[1], 0.5f));
    const float gg = fast_tanh(z[2]), go = __saturatef(fmaf(0.2f, z[3], 0.5f));
    const size_t o = (size_t)b * p.H + unit;
    const float cn = fmaf(gf, p.c_prev[o], gi * gg);
    const float hv = go * fast_tanh(cn);
    p.c_next[o] = cn;
    p.h_next[o] = __float2bfloat16(hv);
    if (p.h_out_f32 != nullptr) p.h_out_f32[o] = hv;
    if (p.gates != nullptr) {
      __nv_bfloat16* gd = p.gates + (size_t)b * G + n0 + lane;
      gd[0] = __float2bfloat16(gi); gd[32] = __float2bfloat16(gf); gd[64] = __float2bfloat16(gg); gd[96] = __float2bfloat16(go);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, gemm_pipe_tmem_cols<BN>());
}

}  // namespace ddls

namespace ddls {

struct BigBwd {
  const __nv_bfloat16* dz_next;  // [B][4H] permuted dZ_{t+1} (A operand); unused when use_acc == 0
  const __nv_bfloat16* u_perm;   // [H][4H] (B operand, K = 4H contiguous)
  const __nv_bfloat16* dout_bf16;// [B][H] gradient arriving from the layer above at this step, or nullptr
  const float* dout_f32;         // [B][H] (last layer, last step), or nullptr
  const __nv_bfloat16* gates;    // [B][4H] permuted gates of step t
  const float* c_t;              // [B][H]
  const float* c_prev;           // [B][H]
  float* dc;                     // [B][H] running cell-state gradient (read-modify-write)
  __nv_bfloat16* dz;             // [B][4H] permuted dZ_t (output)
  int B, H;
  int use_acc;                   // 0 at t = T-1 (no dh from the future)
  int first;                     // 1 at t = T-1: dc starts from zero
};

// Tile = 128 batch rows x 32 hidden units: K = 4H is long, so narrow column tiles (4H/32... H/32 of them) keep
// all SMs busy (4 x 32 = 128 CTAs at B = 512, H = 1024) instead of 32 fat ones.
__global__ void __launch_bounds__(V2_THREADS, 1) lstm_big_bwd_kernel(BigBwd p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t mma_done[V2_STAGES];
  __shared__ uint32_t tmem_slot;
  constexpr int BN = 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * GEMM_BM, n0 = blockIdx.y * BN;      // n0: first hidden unit of this tile
  const int G = 4 * p.H;
  const uint32_t tmem = gemm_pipe_run<BN, false, false>(smem, mma_done, &tmem_slot, p.dz_next, G, p.B, m0, p.u_perm, G, p.H, n0,
                                                         0, p.use_acc ? G : 0);
  float* sT = reinterpret_cast<float*>(smem);                      // [128][33]
  const int q4 = warp & 3, half = warp >> 2;
  {
    uint32_t v[16];
    tmem_ld16(tmem + (static_cast<uint32_t>(q4 * 32) << 16) + half * 16, v);
    tmem_ld_wait();
    const int r = q4 * 32 + lane;
#pragma unroll
    for (int j = 0; j < 16; ++j) sT[r * 33 + half * 16 + j] = p.use_acc ? __uint_as_float(v[j]) : 0.f;
  }
  __syncthreads();
  const int unit = n0 + lane;
  const int pc = (unit >> 5) * 128 + (unit & 31);                  // permuted column of gate 0 of this unit
  for (int rr = warp * 16; rr < warp * 16 + 16; ++rr) {
    const int b = m0 + rr;
    if (b >= p.B) break;
    const size_t o = (size_t)b * p.H + unit;
    float dh = sT[rr * 33 + lane];
    if (p.dout_bf16 != nullptr) dh += __bfloat162float(p.dout_bf16[o]);
    if (p.dout_f32 != nullptr) dh += p.dout_f32[o];
    const __nv_bfloat16* gs = p.gates + (size_t)b * G + pc;
    const float gi = __bfloat162float(gs[0]), gf = __bfloat162float(gs[32]);
    const float gg = __bfloat162float(gs[64]), go = __bfloat162float(gs[96]);
    const float tc = fast_tanh(p.c_t[o]);
    const float dcv = fmaf(dh * go, 1.f - tc * tc, p.first ? 0.f : p.dc[o]);
    const float zo = dh * tc * ((go > 0.f && go < 1.f) ? 0.2f : 0.f);
    const float zi = dcv * gg * ((gi > 0.f && gi < 1.f) ? 0.2f : 0.f);
    const float zf = dcv * p.c_prev[o] * ((gf > 0.f && gf < 1.f) ? 0.2f : 0.f);
    const float zg = dcv * gi * (1.f - gg * gg);
    p.dc[o] = dcv * gf;
    __nv_bfloat16* zd = p.dz + (size_t)b * G + pc;
    zd[0] = __float2bfloat16(zi); zd[32] = __float2bfloat16(zf); zd[64] = __float2bfloat16(zg); zd[96] = __float2bfloat16(zo);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, gemm_pipe_tmem_cols<BN>());
}

}  // namespace ddls

using namespace ddls;
#define ST(s) static_cast<cudaStream_t>(s)

static int check_big(int B, int H, const char* who) {
  if (H % 128 != 0 || H < 128) { set_last_error("%s: hidden size must be a multiple of 128 (got %d)", who, H); return 1; }
  if (B <= 0) { set_last_error("%s: empty batch", who); return 1; }
  return 0;
}

extern "C" int ddls_lstm_big_perm_cast(const float* src, void* dst, int rows, int H, void* stream) {
  if (check_big(1, H, "ddls_lstm_big_perm_cast")) return 1;
  const size_t total = (size_t)rows * 4 * H;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  perm_cast_kernel<<<blocks, 256, 0, ST(stream)>>>(src, static_cast<__nv_bfloat16*>(dst), rows, H);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "perm_cast_kernel launch");
}

extern "C" int ddls_lstm_big_perm_vec(const float* src, float* dst, int H, void* stream) {
  perm_vec_kernel<<<(4 * H + 255) / 256, 256, 0, ST(stream)>>>(src, dst, H);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "perm_vec_kernel launch");
}

extern "C" int ddls_lstm_big_unperm(const float* src, float* dst, int rows, int H, int accumulate, void* stream) {
  const size_t total = (size_t)rows * 4 * H;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  unperm_kernel<<<blocks, 256, 0, ST(stream)>>>(src, dst, rows, H, accumulate);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "unperm_kernel launch");
}

extern "C" int ddls_lstm_big_colsum(const void* X, const float* w, float* out_plain, float* out_weighted, int R, int H,
                                    void* stream) {
  if (check_big(1, H, "ddls_lstm_big_colsum")) return 1;
  dim3 grid((4 * H + 511) / 512, (R + CS_ROWS - 1) / CS_ROWS);
  colsum_perm_kernel<<<grid, 256, 0, ST(stream)>>>(static_cast<const __nv_bfloat16*>(X), w, out_plain, out_weighted, R, H);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "colsum_perm_kernel launch");
}

extern "C" int ddls_lstm_big_step_fwd(const void* h_prev, const void* u_perm, const void* zx, const float* x, int x_stride,
                                      const float* wih_perm, const float* bias_perm, const float* c_prev, float* c_next,
                                      void* h_next, float* h_out_f32, void* gates, int B, int H, void* stream) {
  if (check_big(B, H, "ddls_lstm_big_step_fwd")) return 1;
  BigFwd p{static_cast<const __nv_bfloat16*>(h_prev), static_cast<const __nv_bfloat16*>(u_perm),
           static_cast<const __nv_bfloat16*>(zx), x, x_stride, wih_perm, bias_perm, c_prev, c_next,
           static_cast<__nv_bfloat16*>(h_next), h_out_f32, static_cast<__nv_bfloat16*>(gates), B, H};
  const int smem = gemm_pipe_smem_bytes<128>();
  DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_big_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid((B + 127) / 128, 4 * H / 128);
  lstm_big_fwd_kernel<<<grid, V2_THREADS, smem, ST(stream)>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_big_fwd_kernel launch");
}

extern "C" int ddls_lstm_big_step_bwd(const void* dz_next, const void* u_perm, const void* dout_bf16, const float* dout_f32,
                                      const void* gates, const float* c_t, const float* c_prev, float* dc, void* dz, int B,
                                      int H, int use_acc, int first, void* stream) {
  if (check_big(B, H, "ddls_lstm_big_step_bwd")) return 1;
  BigBwd p{static_cast<const __nv_bfloat16*>(dz_next), static_cast<const __nv_bfloat16*>(u_perm),
           static_cast<const __nv_bfloat16*>(dout_bf16), dout_f32, static_cast<const __nv_bfloat16*>(gates), c_t, c_prev,
           dc, static_cast<__nv_bfloat16*>(dz), B, H, use_acc, first};
  const int smem = gemm_pipe_smem_bytes<32>();
  DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_big_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid((B + 127) / 128, H / 32);
  lstm_big_bwd_kernel<<<grid, V2_THREADS, smem, ST(stream)>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_big_bwd_kernel launch");
}
