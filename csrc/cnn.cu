// MNIST CNN trunk kernels (K1-K3, K6 of SURVEY.md §2.5) for the network of
// /root/reference/ddl_mnist_aztk.py:180-190:
//   Conv2D(32,3x3,valid)+ReLU -> Conv2D(32,3x3)+ReLU -> MaxPool 2x2 -> Flatten -> Dense(225)+ReLU.
// Activations are NHWC (Keras), conv kernels (kh,kw,cin,cout); the Dense GEMMs run through
// csrc/umma_gemm.cu.  This file holds:
//   conv1_relu_fwd        direct FMA conv (K = 9 is too small for tensor cores), bf16 NHWC output
//   conv2_relu_pool_fwd   tcgen05 implicit GEMM, M tile = 32 pooled pixels x 4 window positions so the
//                         bias + ReLU + 2x2 max-pool + arg-max run in the epilogue via lane shuffles
//   relu_bwd_bias         dz = da * (a > 0) (+ bf16 copy) and the Dense bias gradient
//   unpool_relu_bwd       max-pool / ReLU backward to a dense dz2 map + conv2 bias gradient
//   conv2_wgrad           tcgen05 GEMM  dW[(tap,ci), co] = sum_pixels a1_shift * dz2, split over pixels,
//                         partial tiles to scratch + reduce kernel (no atomics)
//   conv2_dgrad_relu      tcgen05 implicit GEMM over input pixels, ReLU mask of conv1 in the epilogue
//   conv1_wgrad           FMA reduction for the 3x3x1x32 kernel and its bias
#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int IMG = 28, C1 = 26, C2 = 24, PL = 12, CH = 32;   // image, conv1 out, conv2 out, pooled, channels
constexpr int K2 = 9 * CH;                                      // conv2 reduction length (tap, cin)
constexpr int CT = 256;                                         // threads per CTA in the tcgen05 kernels

// ------------------------------------------------------------------------------------------------
// conv1 + bias + ReLU : x [B,28,28,1] fp32 -> a1 [B,26,26,32] bf16.  One CTA per image.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) conv1_relu_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                             const float* __restrict__ bias,
                                                             __nv_bfloat16* __restrict__ a1, int B) {
  __shared__ float sx[IMG * IMG];
  __shared__ float sw[9 * CH];
  __shared__ float sb[CH];
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < IMG * IMG; i += blockDim.x) sx[i] = x[(size_t)b * IMG * IMG + i];
  for (int i = threadIdx.x; i < 9 * CH; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < CH) sb[threadIdx.x] = bias[threadIdx.x];
  __syncthreads();
  // item = (pixel, group of 8 channels); 4 consecutive threads write one pixel's 64 bytes
  for (int it = threadIdx.x; it < C1 * C1 * 4; it += blockDim.x) {
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

// stage the conv2 kernel (fp32 [tap][ci][co]) as an MN-major B image: k = tap*32+ci, n = co
//   [K/8=36][N/8=4][8 k][8 n] bf16, LBO (k-group stride) = 512, SBO = 128
DDLS_DEVINL void stage_w2_mnmajor(uint8_t* sB, const float* __restrict__ w2) {
  constexpr int ITER = (K2 * 4 + CT - 1) / CT;   // CT threads; loads batched ahead of the converts
  float4 va[ITER], vb[ITER];
#pragma unroll
  for (int it = 0; it < ITER; ++it) {
    const int idx = threadIdx.x + it * CT;
    if (idx < K2 * 4) {
      const int blk = idx >> 5, ln = idx & 31;
      const int k = blk * 8 + (ln & 7), ng = ln >> 3;
      const float4* src = reinterpret_cast<const float4*>(w2 + (size_t)k * CH + ng * 8);
      va[it] = src[0]; vb[it] = src[1];
    }
  }
#pragma unroll
  for (int it = 0; it < ITER; ++it) {
    const int idx = threadIdx.x + it * CT;
    if (idx < K2 * 4) {
      const int blk = idx >> 5, ln = idx & 31;
      const int k = blk * 8 + (ln & 7), ng = ln >> 3;
      *reinterpret_cast<uint4*>(sB + (k >> 3) * 512 + ng * 128 + (k & 7) * 16) =
          make_uint4(pack_bf16x2(va[it].x, va[it].y), pack_bf16x2(va[it].z, va[it].w), pack_bf16x2(vb[it].x, vb[it].y),
                     pack_bf16x2(vb[it].z, vb[it].w));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// conv2 + bias + ReLU + 2x2 max-pool : a1 [B,26,26,32] bf16 -> pooled [B,12,12,32] bf16, idx u8 (0..3)
// M tile = 128 rows = 32 pooled pixels x 4 window positions (row m = p*4 + q), N = 32, K = 288.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CT, 2) conv2_relu_pool_fwd_kernel(const __nv_bfloat16* __restrict__ a1,
                                                                    const float* __restrict__ w2,
                                                                    const float* __restrict__ bias,
                                                                    __nv_bfloat16* __restrict__ pooled,
                                                                    uint8_t* __restrict__ pidx, int B) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                    // K-major A image [K/8=36][128][16 B] = 73,728 B
  uint8_t* sB = smem + 36 * 2048;        // 18,432 B
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total_pooled = B * PL * PL;
  const int p0 = blockIdx.x * 32;

  if (warp == 0) tmem_alloc(&tmem_s, 32);
  if (tid == 32) { mbar_init(&bar, 1); mbar_fence_init(); }
  stage_w2_mnmajor(sB, w2);
  // im2col gather: item = (k-chunk kc = tap*4 + cg, row m); 32 consecutive lanes write 512 contiguous bytes
  // (constant trip count + full unroll: every LDGSTS gets its own address registers, so the copies are
  //  issued back to back instead of each waiting for the previous one to release its operands)
#pragma unroll
  for (int it = 0; it < 36 * 128 / CT; ++it) {
    const int idx = tid + it * CT;
    const int m = idx & 127, kc = idx >> 7;
    const int tap = kc >> 2, cg = kc & 3, kh = tap / 3, kw = tap - kh * 3;
    const int P = p0 + (m >> 2), q = m & 3;
    const __nv_bfloat16* src = a1;
    uint32_t nbytes = 0;
    if (P < total_pooled) {
      const int b = P / (PL * PL), r = P - b * PL * PL, ph = r / PL, pw = r - ph * PL;
      const int ih = 2 * ph + (q >> 1) + kh, iw = 2 * pw + (q & 1) + kw;
      src = a1 + (((size_t)b * C1 + ih) * C1 + iw) * CH + cg * 8;
      nbytes = 16;
    }
    cp_async16(sA + kc * 2048 + m * 16, src, nbytes);       // LDGSTS: no register staging, 18 copies in flight
  }
  cp_async_commit();
  cp_async_wait<0>();
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (tid == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, 32, false, true);
    const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
#pragma unroll
    for (int ks = 0; ks < K2 / 16; ++ks)
      umma_bf16(tmem, umma_desc(a_base + ks * 2 * 2048, 2048, 128), umma_desc(b_base + ks * 2 * 512, 512, 128), idesc,
                ks > 0);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  // epilogue: warps w and w+4 share TMEM lanes 32*(w%4).. and split the 32 channels in halves
  const int q4 = warp & 3, half = warp >> 2;
  uint32_t v[16];
  tmem_ld16(tmem + (static_cast<uint32_t>(q4 * 32) << 16) + half * 16, v);
  tmem_ld_wait();
  const int m = q4 * 32 + lane, P = p0 + (m >> 2);
  float best[16];
  uint32_t arg[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) { best[c] = __uint_as_float(v[c]); arg[c] = m & 3; }
#pragma unroll
  for (int step = 1; step <= 2; step <<= 1) {
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      const float ov = __shfl_xor_sync(0xffffffffu, best[c], step);
      const uint32_t oa = __shfl_xor_sync(0xffffffffu, arg[c], step);
      const bool take = (ov > best[c]) || (ov == best[c] && oa < arg[c]);   // first maximum wins
      if (take) { best[c] = ov; arg[c] = oa; }
    }
  }
  if ((m & 3) == 0 && P < total_pooled) {
    float r[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) r[c] = fmaxf(best[c] + bias[half * 16 + c], 0.f);
    uint4* dst = reinterpret_cast<uint4*>(pooled + (size_t)P * CH + half * 16);
    dst[0] = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    dst[1] = make_uint4(pack_bf16x2(r[8], r[9]), pack_bf16x2(r[10], r[11]), pack_bf16x2(r[12], r[13]), pack_bf16x2(r[14], r[15]));
    uint32_t pk[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) pk[w] = arg[4 * w] | (arg[4 * w + 1] << 8) | (arg[4 * w + 2] << 16) | (arg[4 * w + 3] << 24);
    *reinterpret_cast<uint4*>(pidx + (size_t)P * CH + half * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 32);
}

// ------------------------------------------------------------------------------------------------
// dz = da * (a > 0): fp32 [B,F] in -> bf16 [B,ldz] out (zero padded columns) + db[f] += sum_b dz
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(const float* __restrict__ da, const float* __restrict__ a,
                                                            __nv_bfloat16* __restrict__ dz, float* __restrict__ db,
                                                            int B, int F, int ldz) {
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

// ------------------------------------------------------------------------------------------------
// conv2 weight gradient: dW2[(tap,ci), co] = sum_{b,oh,ow} a1[b,oh+kh,ow+kw,ci] * dz2[b,oh,ow,co]
// GEMM with M = 288 (padded to 3 x 128: tile = 4 taps x 32 ci), N = 32, K = B*576 pixels, split over CTAs.
// A is MN-major (ci contiguous in NHWC), gathered per tap; B = dz2 is MN-major (co contiguous).
// Partial tiles go to scratch [splits][288][32]; conv2_wgrad_reduce sums them.
// ------------------------------------------------------------------------------------------------
constexpr int WG_KB = 64;                 // pixels per pipeline stage
constexpr int WG_A_STAGE = 8 * 48 * 128;  // [8 k-groups][48 m-groups][8 k][8 m] bf16 = 49,152 B
constexpr int WG_B_STAGE = 8 * 4 * 128;   // [8 k-groups][4 n-groups][8 k][8 n]       =  4,096 B
constexpr int WG_STAGE = WG_A_STAGE + WG_B_STAGE;
constexpr int WG_STAGES = 3;

__global__ void __launch_bounds__(CT, 1) conv2_wgrad_kernel(const __nv_bfloat16* __restrict__ a1,
                                                            const __nv_bfloat16* __restrict__ dz2,
                                                            float* __restrict__ partial, int B, int pix_per_cta) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[WG_STAGES];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total_pix = B * C2 * C2;
  const int k_begin = blockIdx.x * pix_per_cta;
  const int k_end = min(total_pix, k_begin + pix_per_cta);

  if (warp == 0) tmem_alloc(&tmem_s, 128);
  if (tid == 32) { for (int s = 0; s < WG_STAGES; ++s) mbar_init(&bars[s], 1); mbar_fence_init(); }
  // zero the padding m-groups (taps 9..11) of every stage once
  for (int idx = tid; idx < WG_STAGES * 8 * 12 * 8; idx += CT) {
    const int s = idx / (8 * 12 * 8), r = idx - s * (8 * 12 * 8);
    const int kg = r / (12 * 8), r2 = r - kg * (12 * 8);
    *reinterpret_cast<uint4*>(smem + s * WG_STAGE + kg * (48 * 128) + (36 + r2 / 8) * 128 + (r2 & 7) * 16) =
        make_uint4(0, 0, 0, 0);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 32, true, true);
  const int nkb = (k_end - k_begin + WG_KB - 1) / WG_KB;

  auto issue = [&](int kb) {
    uint8_t* sA = smem + (kb % WG_STAGES) * WG_STAGE;
    uint8_t* sB = sA + WG_A_STAGE;
    const int kk0 = k_begin + kb * WG_KB;
    // A gather: item = (m-group mg = tap*4 + cg, pixel k); 8 consecutive lanes -> 8 consecutive pixels (128 B)
#pragma unroll
    for (int it = 0; it < 36 * WG_KB / CT; ++it) {
      const int idx = tid + it * CT;
      const int k = idx & (WG_KB - 1), mg = idx >> 6;
      const int tap = mg >> 2, cg = mg & 3, kh = tap / 3, kw = tap - kh * 3;
      const int pix = kk0 + k;
      const __nv_bfloat16* src = a1;
      uint32_t nb = 0;
      if (pix < k_end) {
        const int b = pix / (C2 * C2), r = pix - b * C2 * C2, oh = r / C2, ow = r - oh * C2;
        src = a1 + (((size_t)b * C1 + oh + kh) * C1 + ow + kw) * CH + cg * 8;
        nb = 16;
      }
      cp_async16(sA + (k >> 3) * (48 * 128) + mg * 128 + (k & 7) * 16, src, nb);
    }
    {                                                  // dz2 rows are 64 contiguous bytes per pixel
      const int idx = tid;
      const int k = idx & (WG_KB - 1), ng = idx >> 6;
      const int pix = kk0 + k;
      const bool ok = pix < k_end;
      cp_async16(sB + (k >> 3) * 512 + ng * 128 + (k & 7) * 16, ok ? dz2 + (size_t)pix * CH + ng * 8 : dz2, ok ? 16 : 0);
    }
  };
  for (int kb = 0; kb < WG_STAGES - 1; ++kb) {
    if (kb < nkb) issue(kb);
    cp_async_commit();
  }
  uint32_t waits[WG_STAGES] = {};
  for (int kb = 0; kb < nkb; ++kb) {
    const int s = kb % WG_STAGES;
    const int nxt = kb + WG_STAGES - 1;
    if (nxt < nkb) {
      const int ns = nxt % WG_STAGES;
      if (kb >= 1) { mbar_wait(&bars[ns], waits[ns] & 1); ++waits[ns]; }
      issue(nxt);
    }
    cp_async_commit();
    cp_async_wait<WG_STAGES - 1>();
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t a_base = smem_u32(smem + s * WG_STAGE), b_base = a_base + WG_A_STAGE;
#pragma unroll
      for (int mt = 0; mt < 3; ++mt)
#pragma unroll
        for (int ks = 0; ks < WG_KB / 16; ++ks)
          umma_bf16(tmem + mt * 32, umma_desc(a_base + mt * 16 * 128 + ks * 2 * (48 * 128), 48 * 128, 128),
                    umma_desc(b_base + ks * 2 * 512, 512, 128), idesc, (kb | ks) != 0);
      umma_commit(&bars[s]);
    }
  }
  cp_async_wait<0>();
  for (int s = 0; s < WG_STAGES; ++s) {
    const uint32_t commits = nkb > s ? (nkb - s + WG_STAGES - 1) / WG_STAGES : 0;
    while (waits[s] < commits) { mbar_wait(&bars[s], waits[s] & 1); ++waits[s]; }
  }
  tc_fence_after();
  if (warp < 4) {
    const int mrow = warp * 32 + lane;
    float* dst = partial + (size_t)blockIdx.x * K2 * CH;
#pragma unroll
    for (int mt = 0; mt < 3; ++mt) {
      uint32_t v[32];
      tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + mt * 32, v);
      tmem_ld_wait();
      const int m = mt * 128 + mrow;
      if (m < K2) {
        float4* o = reinterpret_cast<float4*>(dst + (size_t)m * CH);
#pragma unroll
        for (int c = 0; c < 8; ++c)
          o[c] = nkb > 0 ? make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]),
                                       __uint_as_float(v[4 * c + 2]), __uint_as_float(v[4 * c + 3]))
                         : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

// out[i] = sum_s partial[s][i] : 8 lanes cooperate on one output
__global__ void __launch_bounds__(256) sum_partials_kernel(const float* __restrict__ partial, float* __restrict__ out,
                                                           int n, int splits) {
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
// conv2 data gradient + ReLU mask of conv1:
//   dz1[b,ih,iw,ci] = (a1 > 0) * sum_{kh,kw,co} dz2[b,ih-kh,iw-kw,co] * W2[kh,kw,ci,co]
// implicit GEMM: M tile = 128 input pixels, N = 32 (ci), K = 288 (tap, co)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CT, 2) conv2_dgrad_relu_kernel(const __nv_bfloat16* __restrict__ dz2,
                                                                 const float* __restrict__ w2,
                                                                 const __nv_bfloat16* __restrict__ a1,
                                                                 __nv_bfloat16* __restrict__ dz1, int B) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;               // K-major A image [36][128][16 B]
  uint8_t* sB = smem + 36 * 2048;   // K-major B image [K/8=36][N=32][16 B]: LBO = 512, SBO = 128
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total_pix = B * C1 * C1;
  const int m0 = blockIdx.x * 128;

  if (warp == 0) tmem_alloc(&tmem_s, 32);
  if (tid == 32) { mbar_init(&bar, 1); mbar_fence_init(); }
  // B image: element (n = ci, k = tap*32 + co) = W2[tap][ci][co]; vector = 8 consecutive co
  {
    constexpr int ITER = (36 * CH + CT - 1) / CT;
    float4 va[ITER], vb[ITER];
#pragma unroll
    for (int it = 0; it < ITER; ++it) {
      const int idx = tid + it * CT;
      if (idx < 36 * CH) {
        const int n = idx & 31, kc = idx >> 5;          // kc = tap*4 + cog
        const int tap = kc >> 2, cog = kc & 3;
        const float4* src = reinterpret_cast<const float4*>(w2 + ((size_t)tap * CH + n) * CH + cog * 8);
        va[it] = src[0]; vb[it] = src[1];
      }
    }
#pragma unroll
    for (int it = 0; it < ITER; ++it) {
      const int idx = tid + it * CT;
      if (idx < 36 * CH) {
        const int n = idx & 31, kc = idx >> 5;
        *reinterpret_cast<uint4*>(sB + kc * 512 + n * 16) =
            make_uint4(pack_bf16x2(va[it].x, va[it].y), pack_bf16x2(va[it].z, va[it].w), pack_bf16x2(vb[it].x, vb[it].y),
                       pack_bf16x2(vb[it].z, vb[it].w));
      }
    }
  }
#pragma unroll
  for (int it = 0; it < 36 * 128 / CT; ++it) {
    const int idx = tid + it * CT;
    const int m = idx & 127, kc = idx >> 7;
    const int tap = kc >> 2, cog = kc & 3, kh = tap / 3, kw = tap - kh * 3;
    const int pix = m0 + m;
    const __nv_bfloat16* src = dz2;
    uint32_t nbytes = 0;
    if (pix < total_pix) {
      const int b = pix / (C1 * C1), r = pix - b * C1 * C1, ih = r / C1, iw = r - ih * C1;
      const int oh = ih - kh, ow = iw - kw;
      if (oh >= 0 && oh < C2 && ow >= 0 && ow < C2) {
        src = dz2 + (((size_t)b * C2 + oh) * C2 + ow) * CH + cog * 8;
        nbytes = 16;
      }
    }
    cp_async16(sA + kc * 2048 + m * 16, src, nbytes);
  }
  cp_async_commit();
  cp_async_wait<0>();
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (tid == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, 32, false, false);
    const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
#pragma unroll
    for (int ks = 0; ks < K2 / 16; ++ks)
      umma_bf16(tmem, umma_desc(a_base + ks * 2 * 2048, 2048, 128), umma_desc(b_base + ks * 2 * 512, 512, 128), idesc,
                ks > 0);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  const int q4 = warp & 3, half = warp >> 2;
  uint32_t v[16];
  tmem_ld16(tmem + (static_cast<uint32_t>(q4 * 32) << 16) + half * 16, v);
  tmem_ld_wait();
  const int pix = m0 + q4 * 32 + lane;
  if (pix < total_pix) {
    const uint4* av = reinterpret_cast<const uint4*>(a1 + (size_t)pix * CH + half * 16);
    const uint4 a0 = av[0], a1v = av[1];
    const __nv_bfloat16* ab0 = reinterpret_cast<const __nv_bfloat16*>(&a0);
    const __nv_bfloat16* ab1 = reinterpret_cast<const __nv_bfloat16*>(&a1v);
    float r[16];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      r[c] = (__bfloat162float(ab0[c]) > 0.f) ? __uint_as_float(v[c]) : 0.f;
      r[8 + c] = (__bfloat162float(ab1[c]) > 0.f) ? __uint_as_float(v[8 + c]) : 0.f;
    }
    uint4* dst = reinterpret_cast<uint4*>(dz1 + (size_t)pix * CH + half * 16);
    dst[0] = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    dst[1] = make_uint4(pack_bf16x2(r[8], r[9]), pack_bf16x2(r[10], r[11]), pack_bf16x2(r[12], r[13]), pack_bf16x2(r[14], r[15]));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 32);
}

// ------------------------------------------------------------------------------------------------
// conv1 weight + bias gradient: dW1[tap][co] = sum x[b,oh+kh,ow+kw] * dz1[b,oh,ow,co]; db1[co] = sum dz1
// One CTA per image; thread = (tap or bias slot, co); per-image partials written to scratch
// [B][10][32] and reduced by sum_partials_kernel (no atomics).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(640) conv1_wgrad_kernel(const float* __restrict__ x,
                                                          const __nv_bfloat16* __restrict__ dz1,
                                                          float* __restrict__ partial, int B) {
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
  conv1_relu_fwd_kernel<<<B, 256, 0, ST(stream)>>>(x, w1, b1, static_cast<__nv_bfloat16*>(a1), B);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "conv1_relu_fwd_kernel launch");
}

extern "C" int ddls_cnn_conv2_pool_fwd(const void* a1, const float* w2, const float* b2, void* pooled, void* pidx, int B,
                                       void* stream) {
  if (B <= 0) return 0;
  const int smem = 36 * 2048 + 36 * 512;
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv2_relu_pool_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int tiles = (B * PL * PL + 31) / 32;
  conv2_relu_pool_fwd_kernel<<<tiles, CT, smem, ST(stream)>>>(static_cast<const __nv_bfloat16*>(a1), w2, b2,
                                                               static_cast<__nv_bfloat16*>(pooled),
                                                               static_cast<uint8_t*>(pidx), B);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "conv2_relu_pool_fwd_kernel launch");
}

extern "C" int ddls_relu_bwd_bias(const float* da, const float* a, void* dz, float* db, int B, int F, int ldz,
                                  void* stream) {
  if (B <= 0) return 0;
  dim3 grid((ldz + 255) / 256, B < 8 ? B : 8);
  relu_bwd_bias_kernel<<<grid, 256, 0, ST(stream)>>>(da, a, static_cast<__nv_bfloat16*>(dz), db, B, F, ldz);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "relu_bwd_bias_kernel launch");
}

extern "C" int ddls_cnn_unpool_bwd(const void* dflat, const void* pooled, const void* pidx, void* dz2, float* db2, int B,
                                   void* stream) {
  if (B <= 0) return 0;
  const int total = B * PL * PL * 4;
  unpool_relu_bwd_kernel<<<(total + 255) / 256, 256, 0, ST(stream)>>>(
      static_cast<const __nv_bfloat16*>(dflat), static_cast<const __nv_bfloat16*>(pooled),
      static_cast<const uint8_t*>(pidx), static_cast<__nv_bfloat16*>(dz2), db2, B);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "unpool_relu_bwd_kernel launch");
}

extern "C" size_t ddls_cnn_conv2_wgrad_scratch_bytes(int B) {
  const int total_pix = B * C2 * C2;
  int splits = (total_pix + 511) / 512;
  if (splits > 296) splits = 296;
  return (size_t)splits * K2 * CH * sizeof(float);
}

extern "C" int ddls_cnn_conv2_wgrad(const void* a1, const void* dz2, float* dw2, void* scratch, size_t scratch_bytes, int B,
                                    void* stream) {
  if (B <= 0) return 0;
  const int total_pix = B * C2 * C2;
  int splits = (total_pix + 511) / 512;
  if (splits > 296) splits = 296;
  int per = (total_pix + splits - 1) / splits;
  per = (per + WG_KB - 1) / WG_KB * WG_KB;
  splits = (total_pix + per - 1) / per;
  const size_t need = (size_t)splits * K2 * CH * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need) {
    set_last_error("ddls_cnn_conv2_wgrad: needs %zu bytes of scratch (got %zu)", need, scratch_bytes);
    return 1;
  }
  const int smem = WG_STAGES * WG_STAGE;
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  conv2_wgrad_kernel<<<splits, CT, smem, ST(stream)>>>(static_cast<const __nv_bfloat16*>(a1),
                                                        static_cast<const __nv_bfloat16*>(dz2),
                                                        static_cast<float*>(scratch), B, per);
  const int n = K2 * CH;
  sum_partials_kernel<<<(n * 8 + 255) / 256, 256, 0, ST(stream)>>>(static_cast<const float*>(scratch), dw2, n, splits);
  count_launch(2);
  return check_cuda(cudaGetLastError(), "conv2_wgrad launch");
}

extern "C" int ddls_cnn_conv2_dgrad(const void* dz2, const float* w2, const void* a1, void* dz1, int B, void* stream) {
  if (B <= 0) return 0;
  const int smem = 36 * 2048 + 36 * 512;
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv2_dgrad_relu_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int tiles = (B * C1 * C1 + 127) / 128;
  conv2_dgrad_relu_kernel<<<tiles, CT, smem, ST(stream)>>>(static_cast<const __nv_bfloat16*>(dz2), w2,
                                                            static_cast<const __nv_bfloat16*>(a1),
                                                            static_cast<__nv_bfloat16*>(dz1), B);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "conv2_dgrad_relu_kernel launch");
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
  conv1_wgrad_kernel<<<B, 640, smem, ST(stream)>>>(x, static_cast<const __nv_bfloat16*>(dz1), part, B);
  // per-image partial rows are [9*32 dW1 | 32 db1]; the output buffer has the same 320-float layout
  sum_partials_kernel<<<(320 * 8 + 255) / 256, 256, 0, ST(stream)>>>(part, dw1_db1, 320, B);
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
static float* g_conv1_scratch_addr[NOALLOC_MAX_DEVICES] = {};
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
  if (g_conv1_scratch_addr[dev] == nullptr) {   // first call on this device: resolve the symbol's device address
    float* p = nullptr;
    DDLS_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_conv1_wgrad_scratch));
    const int smem0 = C1 * C1 * CH * 2 + IMG * IMG * 4;
    DDLS_CUDA_OK(cudaFuncSetAttribute(conv1_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem0));
    g_conv1_scratch_addr[dev] = p;
  }
  float* part = g_conv1_scratch_addr[dev];
  const int smem = C1 * C1 * CH * 2 + IMG * IMG * 4;
  conv1_wgrad_kernel<<<B, 640, smem, ST(stream)>>>(x, static_cast<const __nv_bfloat16*>(dz1), part, B);
  sum_partials_kernel<<<(320 * 8 + 255) / 256, 256, 0, ST(stream)>>>(part, dw1_db1, 320, B);
  count_launch(2);
  return check_cuda(cudaGetLastError(), "conv1_wgrad (no-alloc) launch");
}
