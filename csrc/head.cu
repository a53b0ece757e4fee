// Fused output heads: forward AND backward of the terminal layers in one launch, since the loss is
// the end of the graph (K5 / K11 of SURVEY.md §2.5).
//   * Dense(H -> O, linear) + mean_squared_error   (/root/reference/ddl_nyiso_aztk.py:203,207,251,255)
//   * Dense(F -> C) + softmax + categorical_crossentropy (/root/reference/ddl_mnist_aztk.py:191-196)
// Both heads run as ONE CTA (the batches here are 32-1024 rows) and overwrite every output, so nothing has
// to be zeroed beforehand and no atomics are used.
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

// Dense(F -> C) + softmax + categorical cross-entropy, forward AND backward in one launch.
//   a [B,F] fp32 (the previous layer's output), w [F,C], bias [C], onehot [B,C] (null => probabilities only)
// Each CTA owns XH_ROWS batch rows: phase 1 (one warp per row) computes logits / softmax / d = (p - y)/B;
// phase 2 (one thread per input feature, w[f][:] in registers) produces the gradient w.r.t. `a` and the
// CTA's partial weight gradients, which are added to dw / db / dprev_bias with one atomic per element per
// CTA (B/8 CTAs, so 16 adds per address at B = 128).  dw, db, dprev_bias must be zero on entry (they are
// slices of the flat gradient buffer the optimizer kernel clears).  The loss is reduced without a
// pre-zeroed output: CTAs write partial sums to `scratch` and the last CTA to arrive (self-resetting ticket
// in scratch[0]) adds them up.  With relu_input the previous layer's ReLU is folded in (da *= a > 0).
constexpr int XH_THREADS = 256;
constexpr int XH_MAXC = 16;
constexpr int XH_ROWS = 8;
constexpr int XH_MAXF = 256;
template <int CC>
__global__ void __launch_bounds__(XH_THREADS) head_softmax_xent_kernel(
    const float* __restrict__ a, const float* __restrict__ w, const float* __restrict__ bias,
    const float* __restrict__ onehot, float* __restrict__ probs, float* __restrict__ loss, float* __restrict__ da_f32,
    __nv_bfloat16* __restrict__ da_bf16, int ldz, float* __restrict__ dw, float* __restrict__ db,
    float* __restrict__ dprev_bias, float* __restrict__ scratch, int B, int F, int C, int relu_input) {
  pdl_wait();
  extern __shared__ float sm[];
  float* sw = sm;                               // [F][C]
  float* sa = sw + F * C;                       // [XH_ROWS][F]
  float* sd = sa + XH_ROWS * F;                 // [XH_ROWS][CC]
  __shared__ float sloss[XH_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r0 = blockIdx.x * XH_ROWS;
  const int rows = min(XH_ROWS, B - r0);
  {  // stage weights and this CTA's rows: loads batched ahead of the stores
    const int nW = F * C, nA = rows * F;
    const float* asrc = a + (size_t)r0 * F;
    constexpr int IT = (XH_MAXF * XH_MAXC + XH_ROWS * XH_MAXF + XH_THREADS - 1) / XH_THREADS;
    float v[IT];
#pragma unroll
    for (int it = 0; it < IT; ++it) {
      const int i = tid + it * XH_THREADS;
      v[it] = i < nW ? w[i] : (i - nW < nA ? asrc[i - nW] : 0.f);
    }
#pragma unroll
    for (int it = 0; it < IT; ++it) {
      const int i = tid + it * XH_THREADS;
      if (i < nW + nA) sm[i < nW ? i : (F * C + (i - nW))] = v[it];
    }
  }
  __syncthreads();
  const float invB = 1.0f / (float)B;
  float lsum = 0.f;
  if (warp < rows) {
    const int b = r0 + warp;
    float z[CC];
#pragma unroll
    for (int c = 0; c < CC; ++c) z[c] = 0.f;
    for (int ff = lane; ff < F; ff += 32) {
      const float av = sa[warp * F + ff];
#pragma unroll
      for (int c = 0; c < CC; ++c)
        if (CC == 10 || c < C) z[c] = fmaf(av, sw[ff * C + c], z[c]);
    }
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < CC; ++c)
      if (CC == 10 || c < C) { z[c] = warp_sum(z[c]) + bias[c]; mx = fmaxf(mx, z[c]); }
    float ssum = 0.f;
#pragma unroll
    for (int c = 0; c < CC; ++c)
      if (CC == 10 || c < C) { z[c] = __expf(z[c] - mx); ssum += z[c]; }
    const float inv = 1.0f / ssum;
#pragma unroll
    for (int c = 0; c < CC; ++c) {
      if ((CC == 10 || c < C) && lane == c) {
        const float pr = z[c] * inv;
        if (probs != nullptr) probs[(size_t)b * C + c] = pr;
        if (onehot != nullptr) {
          const float yv = onehot[(size_t)b * C + c];
          sd[warp * CC + c] = (pr - yv) * invB;
          // Keras clips probabilities to [1e-7, 1-1e-7] inside the log
          lsum = -yv * __logf(fminf(fmaxf(pr, 1e-7f), 1.f - 1e-7f)) * invB;
        }
      }
    }
  }
  if (onehot == nullptr) return;
  lsum = warp_sum(lsum);
  if (lane == 0) sloss[warp] = lsum;
  __syncthreads();
  // ---- phase 2: thread = input feature ----
  const int f = tid;
  if (f < F) {
    float wr[CC], gw[CC], gprev = 0.f;
#pragma unroll
    for (int c = 0; c < CC; ++c) { wr[c] = (CC == 10 || c < C) ? sw[f * C + c] : 0.f; gw[c] = 0.f; }
    for (int rb = 0; rb < rows; ++rb) {
      const float av = sa[rb * F + f];
      float g = 0.f;
#pragma unroll
      for (int c = 0; c < CC; ++c)
        if (CC == 10 || c < C) { const float d = sd[rb * CC + c]; g = fmaf(d, wr[c], g); gw[c] = fmaf(av, d, gw[c]); }
      if (relu_input && !(av > 0.f)) g = 0.f;
      gprev += g;
      if (da_f32 != nullptr) da_f32[(size_t)(r0 + rb) * F + f] = g;
      if (da_bf16 != nullptr) da_bf16[(size_t)(r0 + rb) * ldz + f] = __float2bfloat16(g);
    }
#pragma unroll
    for (int c = 0; c < CC; ++c)
      if (CC == 10 || c < C) atomicAdd(dw + f * C + c, gw[c]);
    if (dprev_bias != nullptr) atomicAdd(dprev_bias + f, gprev);
  } else if (da_bf16 != nullptr && f < ldz) {
    for (int rb = 0; rb < rows; ++rb) da_bf16[(size_t)(r0 + rb) * ldz + f] = __float2bfloat16(0.f);   // padding columns
  }
  if (tid < C) {
    float v = 0.f;
    for (int rb = 0; rb < rows; ++rb) v += sd[rb * CC + tid];
    atomicAdd(db + tid, v);
  }
  // loss: per-CTA partial -> scratch[1 + block]; last CTA (ticket in scratch[0]) sums in block order
  if (tid == 0) {
    float v = 0.f;
    for (int i = 0; i < XH_THREADS / 32; ++i) v += sloss[i];
    scratch[1 + blockIdx.x] = v;
    __threadfence();
    const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(scratch), 1u);
    if (ticket == gridDim.x - 1) {
      __threadfence();
      float t = 0.f;
      for (unsigned i = 0; i < gridDim.x; ++i) t += reinterpret_cast<volatile float*>(scratch)[1 + i];
      loss[0] = t;
      *reinterpret_cast<unsigned*>(scratch) = 0u;
    }
  }
}

// Dense(H -> O, linear) + MSE, forward AND backward, ONE CTA; every output is overwritten (no zeroing).
//   h [B,H], wd [H,O], bd [O], y [B,O] (null => pred only); outputs pred [B,O], loss[1], dh [B,H], dwd [H,O], dbd [O]
constexpr int MH_THREADS = 512;
template <int O>
__global__ void __launch_bounds__(MH_THREADS, 1) head_mse_kernel(const float* __restrict__ h, const float* __restrict__ wd,
                                                                 const float* __restrict__ bd, const float* __restrict__ y,
                                                                 float* __restrict__ pred, float* __restrict__ loss,
                                                                 float* __restrict__ dh, float* __restrict__ dwd,
                                                                 float* __restrict__ dbd, int B, int H) {
  pdl_wait();
  extern __shared__ float sm[];
  float* sd = sm;                    // [B][O]  dL/dpred
  __shared__ float sloss[MH_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float inv = 1.0f / ((float)B * (float)O);
  float lsum = 0.f;
  for (int b = warp; b < B; b += MH_THREADS / 32) {
    float acc[O];
#pragma unroll
    for (int o = 0; o < O; ++o) acc[o] = 0.f;
    for (int j = lane; j < H; j += 32) {
      const float hv = h[(size_t)b * H + j];
#pragma unroll
      for (int o = 0; o < O; ++o) acc[o] = fmaf(hv, wd[j * O + o], acc[o]);
    }
#pragma unroll
    for (int o = 0; o < O; ++o) {
      const float pr = warp_sum(acc[o]) + bd[o];
      if (lane == 0) {
        pred[(size_t)b * O + o] = pr;
        if (y != nullptr) {
          const float e = pr - y[(size_t)b * O + o];
          sd[b * O + o] = 2.f * e * inv;
          lsum += e * e * inv;
        }
      }
    }
  }
  if (y == nullptr) return;
  if (lane == 0) sloss[warp] = lsum;
  __syncthreads();
  if (warp == 0) {
    float v = (lane < MH_THREADS / 32) ? sloss[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) loss[0] = v;
  }
  // dh[b][j] = sum_o d[b][o] wd[j][o]
  for (int i = tid; i < B * H; i += MH_THREADS) {
    const int b = i / H, j = i - b * H;
    float g = 0.f;
#pragma unroll
    for (int o = 0; o < O; ++o) g = fmaf(sd[b * O + o], wd[j * O + o], g);
    dh[i] = g;
  }
  // dwd[j][o] = sum_b h[b][j] d[b][o]
  for (int i = tid; i < H * O; i += MH_THREADS) {
    const int j = i / O, o = i - j * O;
    float g = 0.f;
    for (int b = 0; b < B; ++b) g = fmaf(h[(size_t)b * H + j], sd[b * O + o], g);
    dwd[i] = g;
  }
  if (tid < O) {
    float g = 0.f;
    for (int b = 0; b < B; ++b) g += sd[b * O + tid];
    dbd[tid] = g;
  }
}

// Large-batch variant (B*H beyond what one CTA does in a few microseconds, e.g. the scaled stacked LSTM's
// [512 x 1024] head): two launches, every output still overwritten.
//   rows kernel: one warp per batch row -> pred, dh (= dL/dpred · wdᵀ)
//   cols kernel: 32 hidden columns per CTA over all rows -> dwd; CTA 0 also reduces the loss and dbd
template <int O>
__global__ void __launch_bounds__(256) head_mse_rows_kernel(const float* __restrict__ h, const float* __restrict__ wd,
                                                            const float* __restrict__ bd, const float* __restrict__ y,
                                                            float* __restrict__ pred, float* __restrict__ dh, int B, int H) {
  pdl_wait();
  const int lane = threadIdx.x & 31, b = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (b >= B) return;
  const float inv = 1.0f / ((float)B * (float)O);
  float acc[O];
#pragma unroll
  for (int o = 0; o < O; ++o) acc[o] = 0.f;
  for (int j = lane; j < H; j += 32) {
    const float hv = h[(size_t)b * H + j];
#pragma unroll
    for (int o = 0; o < O; ++o) acc[o] = fmaf(hv, wd[j * O + o], acc[o]);
  }
  float d[O];
#pragma unroll
  for (int o = 0; o < O; ++o) {
    const float pr = warp_sum(acc[o]) + bd[o];
    if (lane == 0) pred[(size_t)b * O + o] = pr;
    d[o] = (y != nullptr) ? 2.f * (pr - y[(size_t)b * O + o]) * inv : 0.f;
  }
  if (y == nullptr) return;
  for (int j = lane; j < H; j += 32) {
    float g = 0.f;
#pragma unroll
    for (int o = 0; o < O; ++o) g = fmaf(d[o], wd[j * O + o], g);
    dh[(size_t)b * H + j] = g;
  }
}

template <int O>
__global__ void __launch_bounds__(256) head_mse_cols_kernel(const float* __restrict__ h, const float* __restrict__ y,
                                                            const float* __restrict__ pred, float* __restrict__ loss,
                                                            float* __restrict__ dwd, float* __restrict__ dbd, int B, int H) {
  pdl_wait();
  __shared__ float red[8][32][O];
  __shared__ float red2[8][O + 1];
  const int lane = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  const float inv = 1.0f / ((float)B * (float)O);
  float acc[O], dsum[O], lsum = 0.f;
#pragma unroll
  for (int o = 0; o < O; ++o) { acc[o] = 0.f; dsum[o] = 0.f; }
  for (int b = ry; b < B; b += 8) {
    const float hv = (j < H) ? h[(size_t)b * H + j] : 0.f;
#pragma unroll
    for (int o = 0; o < O; ++o) {
      const float e = pred[(size_t)b * O + o] - y[(size_t)b * O + o];
      acc[o] = fmaf(hv, 2.f * e * inv, acc[o]);
      dsum[o] += 2.f * e * inv;
      lsum += e * e * inv;
    }
  }
#pragma unroll
  for (int o = 0; o < O; ++o) red[ry][lane][o] = acc[o];
  if (lane == 0) {
#pragma unroll
    for (int o = 0; o < O; ++o) red2[ry][o] = dsum[o];
    red2[ry][O] = lsum;
  }
  __syncthreads();
  if (ry == 0 && j < H) {
#pragma unroll
    for (int o = 0; o < O; ++o) {
      float g = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) g += red[r][lane][o];
      dwd[j * O + o] = g;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x <= O) {
    float g = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) g += red2[r][threadIdx.x];
    if (threadIdx.x < O) dbd[threadIdx.x] = g; else loss[0] = g;
  }
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_dense_mse(const float* h, const float* wd, const float* bd, const float* y, float* pred,
                              float* loss, float* dh, float* dwd, float* dbd, int B, int H, int O, void* stream) {
  if (B <= 0) return 0;
  if (O > HD_MAX_O || O < 1) { set_last_error("ddls_dense_mse: supports 1 <= O <= 4 (got O=%d)", O); return 1; }
  cudaStream_t st0 = static_cast<cudaStream_t>(stream);
  if ((size_t)B * H > 32768) {
#define DDLS_LAUNCH_MSE_BIG(OO)                                                                                    \
  {                                                                                                                 \
    launch_k(head_mse_rows_kernel<OO>, dim3((B + 7) / 8), dim3(256), 0, st0, h, wd, bd, y, pred, dh, B, H);                          \
    if (y != nullptr) launch_k(head_mse_cols_kernel<OO>, dim3((H + 31) / 32), dim3(256), 0, st0, h, y, pred, loss, dwd, dbd, B, H);   \
  }
    switch (O) {
      case 1: DDLS_LAUNCH_MSE_BIG(1) break;
      case 2: DDLS_LAUNCH_MSE_BIG(2) break;
      case 3: DDLS_LAUNCH_MSE_BIG(3) break;
      default: DDLS_LAUNCH_MSE_BIG(4) break;
    }
#undef DDLS_LAUNCH_MSE_BIG
    count_launch(y != nullptr ? 2 : 1);
    return check_cuda(cudaGetLastError(), "head_mse rows/cols launch");
  }
  const size_t smem = (size_t)B * O * sizeof(float);
  if (smem > 160 * 1024) { set_last_error("ddls_dense_mse: batch %d too large for the single-CTA head", B); return 1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define DDLS_LAUNCH_MSE(OO)                                                                                         \
  {                                                                                                                 \
    if (smem > 48 * 1024)                                                                                           \
      DDLS_CUDA_OK(cudaFuncSetAttribute(head_mse_kernel<OO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    launch_k(head_mse_kernel<OO>, dim3(1), dim3(MH_THREADS), smem, st, h, wd, bd, y, pred, loss, dh, dwd, dbd, B, H);                 \
  }
  switch (O) {
    case 1: DDLS_LAUNCH_MSE(1) break;
    case 2: DDLS_LAUNCH_MSE(2) break;
    case 3: DDLS_LAUNCH_MSE(3) break;
    default: DDLS_LAUNCH_MSE(4) break;
  }
#undef DDLS_LAUNCH_MSE
  count_launch(1);
  return check_cuda(cudaGetLastError(), "head_mse_kernel launch");
}

extern "C" int ddls_dense_softmax_xent(const float* a, const float* w, const float* bias, const float* onehot,
                                       float* probs, float* loss, float* da_f32, void* da_bf16, int ldz, float* dw,
                                       float* db, float* dprev_bias, float* scratch, int B, int F, int C,
                                       int relu_input, void* stream) {
  if (B <= 0) return 0;
  if (C > XH_MAXC || C < 1 || F > XH_MAXF) {
    set_last_error("ddls_dense_softmax_xent: supports C <= 16 and F <= 256 (got C=%d F=%d)", C, F);
    return 1;
  }
  if (da_bf16 != nullptr && ldz > XH_MAXF) { set_last_error("ddls_dense_softmax_xent: ldz must be <= 256"); return 1; }
  const int blocks = (B + XH_ROWS - 1) / XH_ROWS;
  if (onehot != nullptr && scratch == nullptr) {
    set_last_error("ddls_dense_softmax_xent: training needs %d floats of zero-initialised scratch", blocks + 1);
    return 1;
  }
  const int CCsel = (C == 10) ? 10 : XH_MAXC;
  const size_t smem = ((size_t)F * C + (size_t)XH_ROWS * F + (size_t)XH_ROWS * CCsel) * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (C == 10)
    launch_k(head_softmax_xent_kernel<10>, dim3(blocks), dim3(XH_THREADS), smem, st, a, w, bias, onehot, probs, loss, da_f32,
        static_cast<__nv_bfloat16*>(da_bf16), ldz, dw, db, dprev_bias, scratch, B, F, C, relu_input);
  else
    launch_k(head_softmax_xent_kernel<XH_MAXC>, dim3(blocks), dim3(XH_THREADS), smem, st, a, w, bias, onehot, probs, loss, da_f32,
        static_cast<__nv_bfloat16*>(da_bf16), ldz, dw, db, dprev_bias, scratch, B, F, C, relu_input);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "head_softmax_xent_kernel launch");
}
