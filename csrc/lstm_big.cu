// Large-hidden-size LSTM (the scaled stacked forecaster of BASELINE.json config 4: hidden 1024, seq 168,
// batch 512 per GPU; K10 of SURVEY.md §2.5): layout helpers shared by the step kernels (csrc/lstm_big_tma.cu) and the
// whole-sequence persistent kernels (csrc/lstm_big_seq.cu).  Same Keras LSTM math as csrc/lstm.cu (gates i,f,c,o,
// hard_sigmoid, /root/reference/ddl_nyiso_aztk.py:249-251).
//
// To let one epilogue thread see all four gates of a unit inside a 128-column tile, every 4H-wide tensor lives
// in a permuted column order ("tile-gate-unit"): column n = (unit/32)*128 + gate*32 + unit%32.  Weights are
// re-laid-out (and cast to bf16) once per optimizer step by perm_cast; gradients are un-permuted on the way out;
// bias / input-kernel gradients are column sums of dZ.  The input projection of layer >= 2 over all T, the data
// gradient towards the layer below and the weight gradients are plain large GEMMs (csrc/tma_gemm.cu).
#include "common.cuh"
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
