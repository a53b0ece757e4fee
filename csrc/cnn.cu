// MNIST CNN trunk kernels (K1-K3, K6 of SURVEY.md §2.5) for the network of
// /root/reference/ddl_mnist_aztk.py:180-190:
//   Conv2D(32,3x3,valid)+ReLU -> Conv2D(32,3x3)+ReLU -> MaxPool 2x2 -> Flatten -> Dense(225)+ReLU.
// Activations are NHWC (Keras), conv kernels (kh,kw,cin,cout); the Dense GEMMs run through
// csrc/umma_gemm.cu.  This file holds:
//   conv1_relu_fwd        direct FMA conv (K = 9 is too small for tensor cores), bf16 NHWC output
//   relu_bwd_bias         dz = da * (a > 0) (+ bf16 copy) and the Dense bias gradient
//   unpool_relu_bwd       max-pool / ReLU backward to a dense dz2 map + conv2 bias gradient
//   (conv2 forward+pool, data gradient and weight gradient: csrc/cnn_tma.cu, TMA-fed tcgen05 implicit GEMMs)
//   conv1_wgrad           FMA reduction for the 3x3x1x32 kernel and its bias
#include <atomic>

#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int IMG = 28, C1 = 26, C2 = 24, PL = 12, CH = 32;   // image, conv1 out, conv2 out, pooled, channels
constexpr int K2 = 9 * CH;                                      // conv2 reduction length (tap, cin)
constexpr int CT = 256;                                         // threads per CTA in the tcgen05 kernels

// ------------------------------------------------------------------------------------------------
// conv1 + bias + ReLU : x [B,28,28,1] fp32 -> a1 [B,26,26,32] bf16.  Two CTAs per image (13 output rows each): a
// batch of 128 then fills every SM instead of 128 of 148, and the per-CTA dependent chain is half as long.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) conv1_relu_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                             const float* __restrict__ bias,
                                                             __nv_bfloat16* __restrict__ a1, int B) {
  pdl_wait();
  __shared__ float sx[IMG * IMG];
  __shared__ float sw[9 * CH];
  __shared__ float sb[CH];
  const int b = blockIdx.x >> 1, half = blockIdx.x & 1;
  for (int i = threadIdx.x; i < IMG * IMG; i += blockDim.x) sx[i] = x[(size_t)b * IMG * IMG + i];
  for (int i = threadIdx.x; i < 9 * CH; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < CH) sb[threadIdx.x] = bias[threadIdx.x];
  __syncthreads();
  // item = (pixel, group of 8 channels); 4 consecutive threads write one pixel's 64 bytes
  constexpr int HALF_ITEMS = C1 * C1 * 2;            // 13 rows x 26 pixels x 4 channel groups
  for (int it = half * HALF_ITEMS + threadIdx.x; it < (half + 1) * HALF_ITEMS; it += blockDim.x) {
    const int pix = it >> 2, cg = it & 3;
    const int oh = pix / C1, ow = pix - oh * C1;
    float acc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[c] = sb[cg * 8 + c];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const float xv = sx[(oh + kh) * IMG + ow + kw];
        const float* wr = sw + (kh * 3 + kw) * CH + cg * 8;
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] = fmaf(xv, wr[c], acc[c]);
      }
    uint4 v = make_uint4(pack_bf16x2(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f)), pack_bf16x2(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f)),
                         pack_bf16x2(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f)), pack_bf16x2(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f)));
    *reinterpret_cast<uint4*>(a1 + ((size_t)b * C1 * C1 + pix) * CH + cg * 8) = v;
  }
}

// ------------------------------------------------------------------------------------------------
// dz = da * (a > 0): fp32 [B,F] in -> bf16 [B,ldz] out (zero padded columns) + db[f] += sum_b dz
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(const float* __restrict__ da, const float* __restrict__ a,
                                                            __nv_bfloat16* __restrict__ dz, float* __restrict__ db,
                                                            int B, int F, int ldz) {
  pdl_wait();
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= ldz) return;
  float s = 0.f;
  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    float v = 0.f;
    if (f < F) { v = (a[(size_t)b * F + f] > 0.f) ? da[(size_t)b * F + f] : 0.f; s += v; }
    dz[(size_t)b * ldz + f] = __float2bfloat16(v);
  }
  if (f < F) atomicAdd(db + f, s);
}

// ------------------------------------------------------------------------------------------------
// max-pool + ReLU backward: dflat [B,12,12,32] bf16, pooled, idx -> dz2 [B,24,24,32] bf16 ; db2[c] += sum
// One thread per (pooled pixel, 8 channels).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) unpool_relu_bwd_kernel(const __nv_bfloat16* __restrict__ dflat,
                                                              const __nv_bfloat16* __restrict__ pooled,
                                                              const uint8_t* __restrict__ pidx,
                                                              __nv_bfloat16* __restrict__ dz2, float* __restrict__ db2,
                                                              int B) {
  pdl_wait();
  __shared__ float sdb[CH];
  if (threadIdx.x < CH) sdb[threadIdx.x] = 0.f;
  __syncthreads();
  const int total = B * PL * PL * 4;
  const int it = blockIdx.x * blockDim.x + threadIdx.x;
  float part[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int cg = it & 3;
  if (it < total) {
    const int P = it >> 2;
    const int b = P / (PL * PL), r = P - b * PL * PL, ph = r / PL, pw = r - ph * PL;
    const uint4 g = *reinterpret_cast<const uint4*>(dflat + (size_t)P * CH + cg * 8);
    const uint4 pv = *reinterpret_cast<const uint4*>(pooled + (size_t)P * CH + cg * 8);
    const uint2 ix = *reinterpret_cast<const uint2*>(pidx + (size_t)P * CH + cg * 8);
    const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(&g);
    const __nv_bfloat16* pb = reinterpret_cast<const __nv_bfloat16*>(&pv);
    const uint8_t* ib = reinterpret_cast<const uint8_t*>(&ix);
    __nv_bfloat16 outv[4][8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const bool live = __bfloat162float(pb[c]) > 0.f;
      const __nv_bfloat16 gv = live ? gb[c] : __float2bfloat16(0.f);
      part[c] = __bfloat162float(gv);
#pragma unroll
      for (int q = 0; q < 4; ++q) outv[q][c] = (ib[c] == q) ? gv : __float2bfloat16(0.f);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int oh = 2 * ph + (q >> 1), ow = 2 * pw + (q & 1);
      *reinterpret_cast<uint4*>(dz2 + (((size_t)b * C2 + oh) * C2 + ow) * CH + cg * 8) = *reinterpret_cast<uint4*>(outv[q]);
    }
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    // lanes with equal (lane & 3) hold the same channel group: reduce over the other lane bits
    float v = part[c];
    v += __shfl_xor_sync(0xffffffffu, v, 4); v += __shfl_xor_sync(0xffffffffu, v, 8); v += __shfl_xor_sync(0xffffffffu, v, 16);
    if ((threadIdx.x & 31) < 4) atomicAdd(&sdb[cg * 8 + c], v);
  }
  __syncthreads();
  if (threadIdx.x < CH) atomicAdd(db2 + threadIdx.x, sdb[threadIdx.x]);
}

// out[i] = sum_s partial[s][i] : 8 lanes cooperate on one output
__global__ void __launch_bounds__(256) sum_partials_kernel(const float* __restrict__ partial, float* __restrict__ out,
                                                           int n, int splits) {
  pdl_wait();
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = gid >> 3, sub = gid & 7;
  float acc = 0.f;
  if (i < n)
    for (int s = sub; s < splits; s += 8) acc += partial[(size_t)s * n + i];
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (i < n && sub == 0) out[i] = acc;
}

// ------------------------------------------------------------------------------------------------
// conv1 weight + bias gradient: dW1[tap][co] = sum x[b,oh+kh,ow+kw] * dz1[b,oh,ow,co]; db1[co] = sum dz1
// One CTA per image; thread = (tap or bias slot, co); per-image partials written to scratch
// [B][10][32] and reduced by sum_partials_kernel (no atomics).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(640) conv1_wgrad_kernel(const float* __restrict__ x,
                                                          const __nv_bfloat16* __restrict__ dz1,
                                                          float* __restrict__ partial, int B) {
  pdl_wait();
  extern __shared__ __align__(16) uint8_t smem[];
  __nv_bfloat16* sdz = reinterpret_cast<__nv_bfloat16*>(smem);                       // [676][32]
  float* sx = reinterpret_cast<float*>(smem + C1 * C1 * CH * 2);                     // [784]
  const int b = blockIdx.x;
  const uint4* src = reinterpret_cast<const uint4*>(dz1 + (size_t)b * C1 * C1 * CH);
  for (int i = threadIdx.x; i < C1 * C1 * CH / 8; i += blockDim.x) reinterpret_cast<uint4*>(sdz)[i] = src[i];
  for (int i = threadIdx.x; i < IMG * IMG; i += blockDim.x) sx[i] = x[(size_t)b * IMG * IMG + i];
  __syncthreads();
  // 640 threads: two groups of 320 = (slot, co) split the 26 output rows; four independent accumulators per thread
  // break the 676-long dependent FMA chain of the first version (13.5 us -> see profiles/kernel_times_mnist_cnn.json)
  __shared__ float red[320];
  const int grp = threadIdx.x / 320, t = threadIdx.x - grp * 320;
  const int slot = t >> 5, co = t & 31;                          // slot 0..8 = tap, 9 = bias
  const int kh = slot / 3, kw = slot - kh * 3;
  const int ngrp = blockDim.x / 320;
  const int oh0 = grp * (C1 / 2 + (ngrp == 1 ? C1 / 2 : 0)), oh1 = (ngrp == 1 || grp == 1) ? C1 : C1 / 2;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  if (slot < 9) {
    for (int oh = oh0; oh < oh1; ++oh) {
      const float* xr = sx + (oh + kh) * IMG + kw;
      const __nv_bfloat16* zr = sdz + (size_t)oh * C1 * CH + co;
#pragma unroll
      for (int ow = 0; ow < 24; ow += 4) {
        a0 = fmaf(xr[ow], __bfloat162float(zr[ow * CH]), a0);
        a1 = fmaf(xr[ow + 1], __bfloat162float(zr[(ow + 1) * CH]), a1);
        a2 = fmaf(xr[ow + 2], __bfloat162float(zr[(ow + 2) * CH]), a2);
        a3 = fmaf(xr[ow + 3], __bfloat162float(zr[(ow + 3) * CH]), a3);
      }
      a0 = fmaf(xr[24], __bfloat162float(zr[24 * CH]), a0);
      a1 = fmaf(xr[25], __bfloat162float(zr[25 * CH]), a1);
    }
  } else {
    for (int i = oh0 * C1; i < oh1 * C1; i += 2) {
      a0 += __bfloat162float(sdz[i * CH + co]);
      a1 += __bfloat162float(sdz[(i + 1) * CH + co]);
    }
  }
  float acc = (a0 + a1) + (a2 + a3);
  if (grp == 1) red[t] = acc;
  __syncthreads();
  if (grp == 0) {
    if (ngrp == 2) acc += red[t];
    partial[((size_t)b * 10 + slot) * CH + co] = acc;
  }
}

}  // namespace ddls

using namespace ddls;

#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int ddls_cnn_conv1_fwd(const float* x, const float* w1, const float* b1, void* a1, int B, void* stream) {
  if (B <= 0) return 0;
  DDLS_CUDA_OK(launch_k(conv1_relu_fwd_kernel, dim3(2 * B), dim3(256), 0, ST(stream), x, w1, b1, static_cast<__nv_bfloat16*>(a1), B));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "conv1_relu_fwd_kernel launch");
}

extern "C" int ddls_relu_bwd_bias(const float* da, const float* a, void* dz, float* db, int B, int F, int ldz,
                                  void* stream) {
  if (B <= 0) return 0;
  dim3 grid((ldz + 255) / 256, B < 8 ? B : 8);
  DDLS_CUDA_OK(launch_k(relu_bwd_bias_kernel, grid, dim3(256), 0, ST(stream), da, a, static_cast<__nv_bfloat16*>(dz), db, B, F, ldz));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "relu_bwd_bias_kernel launch");
}

extern "C" int ddls_cnn_unpool_bwd(const void* dflat, const void* pooled, const void* pidx, void* dz2, float* db2, int B,
                                   void* stream) {
  if (B <= 0) return 0;
  const int total = B * PL * PL * 4;
  DDLS_CUDA_OK(launch_k(unpool_relu_bwd_kernel, dim3((total + 255) / 256), dim3(256), 0, ST(stream),
      static_cast<const __nv_bfloat16*>(dflat), static_cast<const __nv_bfloat16*>(pooled),
      static_cast<const uint8_t*>(pidx), static_cast<__nv_bfloat16*>(dz2), db2, B));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "unpool_relu_bwd_kernel launch");
}

extern "C" int ddls_cnn_conv1_wgrad(const float* x, const void* dz1, float* dw1_db1, void* scratch, size_t scratch_bytes,
                                    int B, void* stream) {
  if (B <= 0) return 0;
  const size_t need = (size_t)B * 10 * CH * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need) {
    set_last_error("ddls_cnn_conv1_wgrad: needs %zu bytes of scratch (got %zu)", need, scratch_bytes);
    return 1;
  }
  const int smem = C1 * C1 * CH * 2 + IMG * IMG * 4;
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv1_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  float* part = static_cast<float*>(scratch);
  DDLS_CUDA_OK(launch_k(conv1_wgrad_kernel, dim3(B), dim3(640), smem, ST(stream), x, static_cast<const __nv_bfloat16*>(dz1), part, B));
  // per-image partial rows are [9*32 dW1 | 32 db1]; the output buffer has the same 320-float layout
  DDLS_CUDA_OK(launch_k(sum_partials_kernel, dim3((320 * 8 + 255) / 256), dim3(256), 0, ST(stream), part, dw1_db1, 320, B));
  count_launch(2);
  return check_cuda(cudaGetLastError(), "conv1_wgrad launch");
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of ddls_cnn_conv1_wgrad.
//
// ddls_cnn_conv1_wgrad needs B*320 floats of scratch that the caller must provide, which in practice
// means the caller (or PyTorch's caching allocator on its behalf) has called cudaMalloc at some point.
// Some users of this C ABI cannot do that: a host program that was handed a fixed slice of GPU memory by
// a resource manager (several Spark tasks sharing one B200) and must not grow its footprint at run time;
// a latency-critical loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply
// while copy and collective streams are in flight; a plain C caller without any allocator of its own.
// For them the scratch lives in the file-scope device array below, reserved when the library is loaded:
// every GPU that loads the library holds these 1.25 MiB until the process exits, whether or not the
// entry point is ever called.  The price is a batch limit (DDLS_NOALLOC_CONV1_MAX_BATCH images).
//
// The array name is only a host-side stand-in for the symbol, not a device address, so the entry point
// resolves the address once per device with cudaGetSymbolAddress (with that device current) on its first
// call there and hands the resolved pointer to the kernels.  Make that first call on each device before
// starting any stream capture on it.  The entry point itself calls no cudaMalloc/cudaFree or any other
// allocating routine.  All calls on a device share the one array and it keeps whatever the last call left
// in it: every element the reduction reads is rewritten by the producer kernel of the same call, so nothing
// has to be cleared, but two calls whose kernels overlap on the GPU (different streams, no dependency)
// race on it — order such calls yourself.
// ------------------------------------------------------------------------------------------------
namespace ddls {
constexpr int NOALLOC_MAX_B = DDLS_NOALLOC_CONV1_MAX_BATCH;
static __device__ __align__(256) float g_conv1_wgrad_scratch[(size_t)NOALLOC_MAX_B * 10 * CH];
constexpr int NOALLOC_MAX_DEVICES = 64;
// resolved once per device; atomic because two host threads may make the first call on a device at the same time (both
// would store the same address, but a plain static is still a data race)
static std::atomic<float*> g_conv1_scratch_addr[NOALLOC_MAX_DEVICES];
}  // namespace ddls

extern "C" int ddls_cnn_conv1_wgrad_noalloc(const float* x, const void* dz1, float* dw1_db1, int B, void* stream) {
  if (B <= 0) return 0;
  if (B > NOALLOC_MAX_B) {
    set_last_error("ddls_cnn_conv1_wgrad_noalloc: batch %d exceeds the built-in scratch (max %d images); "
                   "use ddls_cnn_conv1_wgrad with caller-provided scratch", B, NOALLOC_MAX_B);
    return 1;
  }
  int dev = -1;
  DDLS_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
    set_last_error("ddls_cnn_conv1_wgrad_noalloc: device ordinal %d out of range", dev);
    return 1;
  }
  if (g_conv1_scratch_addr[dev].load(std::memory_order_acquire) == nullptr) {   // first call on this device: resolve the symbol's device address
    float* p = nullptr;
    DDLS_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_conv1_wgrad_scratch));
    const int smem0 = C1 * C1 * CH * 2 + IMG * IMG * 4;
    DDLS_CUDA_OK(cudaFuncSetAttribute(conv1_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem0));
    g_conv1_scratch_addr[dev].store(p, std::memory_order_release);
  }
  float* part = g_conv1_scratch_addr[dev].load(std::memory_order_acquire);
  const int smem = C1 * C1 * CH * 2 + IMG * IMG * 4;
  DDLS_CUDA_OK(launch_k(conv1_wgrad_kernel, dim3(B), dim3(640), smem, ST(stream), x, static_cast<const __nv_bfloat16*>(dz1), part, B));
  DDLS_CUDA_OK(launch_k(sum_partials_kernel, dim3((320 * 8 + 255) / 256), dim3(256), 0, ST(stream), part, dw1_db1, 320, B));
  count_launch(2);
  return check_cuda(cudaGetLastError(), "conv1_wgrad (no-alloc) launch");
}
