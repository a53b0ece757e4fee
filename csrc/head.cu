// Fused output heads: forward AND backward of the terminal layers in one launch, since the loss is
// the end of the graph (K5 / K11 of SURVEY.md §2.5).
//   * Dense(H -> O, linear) + mean_squared_error   (/root/reference/ddl_nyiso_aztk.py:203,207,251,255)
//   * Dense(F -> C) + softmax + categorical_crossentropy (/root/reference/ddl_mnist_aztk.py:191-196)
// Each warp walks batch rows; per-lane weight-gradient partials stay in registers and are reduced with
// one atomic per element at the end.  Scratch: none (outputs are accumulated into caller-zeroed buffers).
#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int HD_THREADS = 256;
constexpr int HD_MAX_O = 4;

DDLS_DEVINL float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// h [B,H] fp32, wd [H,O], bd [O], y [B,O] (may be null => inference: pred only)
// zeroed by caller: loss[1], dwd [H,O], dbd [O].  H <= 32*4 per lane stride loop (any H).
template <int O>
__global__ void __launch_bounds__(HD_THREADS) dense_mse_kernel(const float* __restrict__ h, const float* __restrict__ wd,
                                                               const float* __restrict__ bd, const float* __restrict__ y,
                                                               float* __restrict__ pred, float* __restrict__ loss,
                                                               float* __restrict__ dh, float* __restrict__ dwd,
                                                               float* __restrict__ dbd, int B, int H) {
  const int lane = threadIdx.x & 31;
  const int warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  constexpr int MAXJ = 8;                 // supports H <= 256 per lane-strided loop
  float w[MAXJ][O], gw[MAXJ][O];
#pragma unroll
  for (int u = 0; u < MAXJ; ++u) {
    const int j = lane + 32 * u;
#pragma unroll
    for (int o = 0; o < O; ++o) { w[u][o] = (j < H) ? wd[j * O + o] : 0.f; gw[u][o] = 0.f; }
  }
  float gb[O], lsum = 0.f;
#pragma unroll
  for (int o = 0; o < O; ++o) gb[o] = 0.f;
  const float inv = 1.0f / ((float)B * (float)O);
  for (int b = warp_g; b < B; b += nwarps) {
    float hv[MAXJ], acc[O];
#pragma unroll
    for (int o = 0; o < O; ++o) acc[o] = 0.f;
#pragma unroll
    for (int u = 0; u < MAXJ; ++u) {
      const int j = lane + 32 * u;
      hv[u] = (j < H) ? h[(size_t)b * H + j] : 0.f;
#pragma unroll
      for (int o = 0; o < O; ++o) acc[o] = fmaf(hv[u], w[u][o], acc[o]);
    }
    float d[O];
#pragma unroll
    for (int o = 0; o < O; ++o) {
      const float pr = warp_sum(acc[o]) + bd[o];
      if (lane == 0) pred[(size_t)b * O + o] = pr;
      if (y != nullptr) {
        const float e = pr - y[(size_t)b * O + o];
        d[o] = 2.f * e * inv;
        if (lane == 0) { lsum += e * e * inv; gb[o] += d[o]; }
      } else {
        d[o] = 0.f;
      }
    }
    if (y != nullptr) {
#pragma unroll
      for (int u = 0; u < MAXJ; ++u) {
        const int j = lane + 32 * u;
        float dhv = 0.f;
#pragma unroll
        for (int o = 0; o < O; ++o) { dhv = fmaf(d[o], w[u][o], dhv); gw[u][o] = fmaf(hv[u], d[o], gw[u][o]); }
        if (j < H) dh[(size_t)b * H + j] = dhv;
      }
    }
  }
  if (y != nullptr) {
#pragma unroll
    for (int u = 0; u < MAXJ; ++u) {
      const int j = lane + 32 * u;
      if (j < H) {
#pragma unroll
        for (int o = 0; o < O; ++o) atomicAdd(dwd + j * O + o, gw[u][o]);
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int o = 0; o < O; ++o) atomicAdd(dbd + o, gb[o]);
      atomicAdd(loss, lsum);
    }
  }
}

// Dense(F -> C) + softmax + categorical cross-entropy, forward and backward in one launch.
// a [B,F] fp32, w [F,C], bias [C], onehot [B,C] (null => probabilities only).
// zeroed by caller: loss[1], dw [F,C], db [C].  C <= 32; lane c owns class c; F walked in chunks.
__global__ void __launch_bounds__(HD_THREADS) dense_softmax_xent_kernel(
    const float* __restrict__ a, const float* __restrict__ w, const float* __restrict__ bias,
    const float* __restrict__ onehot, float* __restrict__ probs, float* __restrict__ loss, float* __restrict__ da,
    float* __restrict__ dw, float* __restrict__ db, int B, int F, int C) {
  extern __shared__ float sm[];
  float* sw = sm;                         // [F][C] weights
  float* sdw = sm + (size_t)F * C;        // [F][C] block-local weight-gradient accumulator
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int i = threadIdx.x; i < F * C; i += blockDim.x) { sw[i] = w[i]; sdw[i] = 0.f; }
  __syncthreads();
  const float bc = (lane < C) ? bias[lane] : 0.f;
  float gb = 0.f, lsum = 0.f;
  const float invB = 1.0f / (float)B;
  for (int b = blockIdx.x * nw + warp; b < B; b += gridDim.x * nw) {
    const float* arow = a + (size_t)b * F;
    // logits: lane c accumulates over f (weights from smem: bank = (f*C + c) % 32, conflict-free for odd strides)
    float z = bc;
    if (lane < C) {
      for (int f = 0; f < F; ++f) z = fmaf(arow[f], sw[f * C + lane], z);
    } else {
      z = -INFINITY;
    }
    float mx = z;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float e = (lane < C) ? __expf(z - mx) : 0.f;
    const float s = warp_sum(e);
    const float pr = e / s;
    if (lane < C) probs[(size_t)b * C + lane] = pr;
    if (onehot != nullptr) {
      const float yv = (lane < C) ? onehot[(size_t)b * C + lane] : 0.f;
      // Keras clips probabilities to [1e-7, 1-1e-7] inside the log
      const float pl = fminf(fmaxf(pr, 1e-7f), 1.f - 1e-7f);
      lsum += (lane < C) ? -yv * __logf(pl) * invB : 0.f;
      const float d = (lane < C) ? (pr - yv) * invB : 0.f;     // dL/dz for softmax + xent
      gb += d;
      // da[f] = sum_c d_c * w[f][c]  (lanes stride over f; d broadcast by shuffle)
      for (int f0 = 0; f0 < F; f0 += 32) {
        const int f = f0 + lane;
        float acc = 0.f;
        for (int c = 0; c < C; ++c) {
          const float dc = __shfl_sync(0xffffffffu, d, c);
          if (f < F) acc = fmaf(dc, sw[f * C + c], acc);
        }
        if (f < F) da[(size_t)b * F + f] = acc;
      }
      // dw[f][c] += a[f] * d_c : lane c walks f
      if (lane < C) {
        for (int f = 0; f < F; ++f) atomicAdd(&sdw[f * C + lane], arow[f] * d);
      }
    }
  }
  if (onehot != nullptr) {
    __syncthreads();
    for (int i = threadIdx.x; i < F * C; i += blockDim.x) atomicAdd(dw + i, sdw[i]);
    if (lane < C) atomicAdd(db + lane, gb);
    lsum = warp_sum(lsum);
    if (lane == 0) atomicAdd(loss, lsum);
  }
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_dense_mse(const float* h, const float* wd, const float* bd, const float* y, float* pred,
                              float* loss, float* dh, float* dwd, float* dbd, int B, int H, int O, void* stream) {
  if (B <= 0) return 0;
  if (H > 256 || O > HD_MAX_O || O < 1) { set_last_error("ddls_dense_mse: supports H <= 256 and 1 <= O <= 4 (got H=%d O=%d)", H, O); return 1; }
  const int warps_per_block = HD_THREADS / 32;
  int blocks = (B + warps_per_block - 1) / warps_per_block;
  if (blocks > 148 * 2) blocks = 148 * 2;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (O) {
    case 1: dense_mse_kernel<1><<<blocks, HD_THREADS, 0, st>>>(h, wd, bd, y, pred, loss, dh, dwd, dbd, B, H); break;
    case 2: dense_mse_kernel<2><<<blocks, HD_THREADS, 0, st>>>(h, wd, bd, y, pred, loss, dh, dwd, dbd, B, H); break;
    case 3: dense_mse_kernel<3><<<blocks, HD_THREADS, 0, st>>>(h, wd, bd, y, pred, loss, dh, dwd, dbd, B, H); break;
    default: dense_mse_kernel<4><<<blocks, HD_THREADS, 0, st>>>(h, wd, bd, y, pred, loss, dh, dwd, dbd, B, H); break;
  }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "dense_mse_kernel launch");
}

extern "C" int ddls_dense_softmax_xent(const float* a, const float* w, const float* bias, const float* onehot,
                                       float* probs, float* loss, float* da, float* dw, float* db, int B, int F, int C,
                                       void* stream) {
  if (B <= 0) return 0;
  if (C > 32 || C < 1) { set_last_error("ddls_dense_softmax_xent: supports 1 <= C <= 32 (got %d)", C); return 1; }
  const size_t smem = (size_t)F * C * 2 * sizeof(float);
  if (smem > 200 * 1024) { set_last_error("ddls_dense_softmax_xent: F*C too large for shared memory"); return 1; }
  DDLS_CUDA_OK(cudaFuncSetAttribute(dense_softmax_xent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int nw = HD_THREADS / 32;
  int blocks = (B + nw - 1) / nw;
  if (blocks > 148) blocks = 148;
  dense_softmax_xent_kernel<<<blocks, HD_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(
      a, w, bias, onehot, probs, loss, da, dw, db, B, F, C);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "dense_softmax_xent_kernel launch");
}
