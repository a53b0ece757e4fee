// Second generation of the two heaviest MNIST-CNN kernels (K2/K3/K6 of SURVEY.md §2.5; network of
// /root/reference/ddl_mnist_aztk.py:180-190): conv2 + bias + ReLU + 2x2 max-pool forward and the conv2 data gradient.
//
// The first-generation kernels (csrc/cnn.cu) gather their im2col tile with one cp.async per 16 bytes: 14.8 M warp
// instructions per launch, one tile per CTA, 22-27 us.  Here the im2col tile is never gathered by threads: an M tile is
// a band of 4 full image rows, and for every filter tap the band shifted by (kh, kw) is ONE 4-D box of the NHWC
// activation tensor — `cp.async.bulk.tensor.4d` with 64-byte swizzle writes it as a K-major UMMA operand block
// (32 channels = 64 bytes per pixel row); image borders of the data gradient are TMA out-of-bounds zero fill.
// CTAs are persistent: warp 0 issues the nine tap boxes of the next band, warp 1 the 18 MMAs (M128 x N32 x K16),
// warps 2-5 drain the previous band's accumulator from TMEM (double buffered) — max-pool / ReLU mask / bf16 stores.
// The filter (18 KB as bf16) is staged once per CTA instead of once per tile.
#include "tma_pipe.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

namespace {
constexpr int C1 = 26, C2 = 24, PL = 12, CH = 32, K2 = 9 * CH;
constexpr int CV_THREADS = 192;
constexpr int TAP_BYTES = 128 * 64;                 // one tap block: 128 pixel rows x 32 channels bf16
constexpr int STAGE_BYTES = 9 * TAP_BYTES;          // 72 KB
constexpr int CV_STAGES = 2;
constexpr int W_BYTES = 36 * 512;                   // filter image, SWIZZLE_NONE canonical layout
constexpr int EPI_LD = 33;
constexpr int CV_SMEM = CV_STAGES * STAGE_BYTES + W_BYTES + 128 * EPI_LD * 4 + 1024;

struct ConvParams {
  const float* w2;              // [9][32 ci][32 co] fp32
  const float* bias;            // forward: [32]
  const __nv_bfloat16* a1;      // dgrad: conv1 activations for the ReLU mask
  __nv_bfloat16* out;           // forward: pooled [B,12,12,32]; dgrad: dz1 [B,26,26,32]
  uint8_t* pidx;                // forward: arg-max (0..3) per pooled element
  int B, tiles;
};

// K-major SWIZZLE_64B descriptor low word (LBO unused for swizzled K-major)
DDLS_DEVINL uint32_t sw64_lo(uint32_t addr) { return umma_desc_lo(addr, 16); }
}  // namespace

// FWD: B operand = filter as MN-major image (k = tap*32 + ci, n = co), exactly the layout of stage_w2_mnmajor in
// csrc/cnn.cu: [K/8 = 36][N/8 = 4][8 k][8 n].  DGRAD: K-major image (n = ci, k = tap*32 + co): [K/8 = 36][N = 32][8 k].
template <bool DGRAD>
DDLS_DEVINL void stage_filter(uint8_t* sB, const float* __restrict__ w2) {
  for (int idx = threadIdx.x; idx < 36 * 32; idx += blockDim.x) {
    if (!DGRAD) {
      const int blk = idx >> 5, ln = idx & 31;
      const int k = blk * 8 + (ln & 7), ng = ln >> 3;
      const float4* src = reinterpret_cast<const float4*>(w2 + (size_t)k * CH + ng * 8);
      const float4 a = src[0], b = src[1];
      *reinterpret_cast<uint4*>(sB + (k >> 3) * 512 + ng * 128 + (k & 7) * 16) =
          make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    } else {
      const int n = idx & 31, kc = idx >> 5;          // kc = tap*4 + group of 8 co
      const int tap = kc >> 2, cog = kc & 3;
      const float4* src = reinterpret_cast<const float4*>(w2 + ((size_t)tap * CH + n) * CH + cog * 8);
      const float4 a = src[0], b = src[1];
      *reinterpret_cast<uint4*>(sB + kc * 512 + n * 16) =
          make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    }
  }
}

// Tiles: forward = (image, band of 4 output rows) -> 6 per image, 96 of the 128 M rows used;
//        dgrad   = (image, band of 4 input rows)  -> 7 per image, 104 rows used (the last band has 2 valid rows).
template <bool DGRAD>
__global__ void __launch_bounds__(CV_THREADS, 1) conv2_tma_kernel(const __grid_constant__ CUtensorMap tmap, ConvParams p) {
  constexpr int BANDS = DGRAD ? 7 : 6;
  constexpr int WID = DGRAD ? C1 : C2;                // pixels per band row
  constexpr int ROWS = 4 * WID;                       // used M rows
  constexpr uint32_t BOX_BYTES = ROWS * 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sB = smem + CV_STAGES * STAGE_BYTES;
  float* epi = reinterpret_cast<float*>(sB + W_BYTES);
  __shared__ uint64_t full[CV_STAGES], empty[CV_STAGES], tfull[2], tempty[2];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < CV_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
    mbar_fence_init();
    tma_prefetch_desc(&tmap);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 64);
  // rows ROWS..127 of every tap block are never written by TMA: clear them once so the unused accumulator rows stay finite
  for (int i = threadIdx.x; i < CV_STAGES * 9 * (128 - ROWS) * 4; i += blockDim.x) {
    const int blk = i / ((128 - ROWS) * 4), r = i - blk * (128 - ROWS) * 4;
    *reinterpret_cast<uint4*>(smem + blk * TAP_BYTES + ROWS * 64 + r * 16) = make_uint4(0u, 0u, 0u, 0u);
  }
  pdl_wait();                                         // everything above overlapped the previous kernel's tail
  stage_filter<DGRAD>(sB, p.w2);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        const int b = tile / BANDS, y0 = (tile - b * BANDS) * 4;
        const int s = it % CV_STAGES, ph = (it / CV_STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], 9 * BOX_BYTES);
        const uint32_t base = smem_u32(smem + s * STAGE_BYTES);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int kh = tap / 3, kw = tap - kh * 3;
          if (DGRAD) tma_load_4d(base + tap * TAP_BYTES, &tmap, &full[s], 0, -kw, y0 - kh, b);
          else tma_load_4d(base + tap * TAP_BYTES, &tmap, &full[s], 0, kw, y0 + kh, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 32, false, !DGRAD);
      const uint32_t b_base = smem_u32(sB);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        const int s = it % CV_STAGES, ph = (it / CV_STAGES) & 1;
        const int acc = it & 1, aph = (it >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * STAGE_BYTES);
#pragma unroll
        for (int ks = 0; ks < K2 / 16; ++ks) {
          const uint32_t a_lo = sw64_lo(a_base + (ks >> 1) * TAP_BYTES + (ks & 1) * 32);
          const uint64_t bd = umma_desc(b_base + ks * 2 * 512, 512, 128);
          umma_bf16_lohi(tmem + acc * 32, a_lo, SW64_DESC_HI, static_cast<uint32_t>(bd), static_cast<uint32_t>(bd >> 32), idesc,
                         ks > 0);
        }
        umma_commit(&empty[s]);
        umma_commit(&tfull[acc]);
      }
    }
  } else {
    const int quad = warp & 3;
    const int m = quad * 32 + lane;                     // M row = pixel (y - y0) * WID + x
    const int et = threadIdx.x - 64, ew = et >> 5;      // 0..127 among the epilogue threads
    int it = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      const int b = tile / BANDS, y0 = (tile - b * BANDS) * 4;
      const int acc = it & 1, aph = (it >> 1) & 1;
      if (DGRAD) {
        const int y = y0 + m / WID, x = m - (m / WID) * WID;
        const bool live = m < ROWS && y < C1;
        uint4 mask[4];
        const size_t o = (((size_t)b * C1 + y) * C1 + x) * CH;
        if (live) {
#pragma unroll
          for (int j = 0; j < 4; ++j) mask[j] = *reinterpret_cast<const uint4*>(p.a1 + o + j * 8);
        }
        mbar_wait(&tfull[acc], aph);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tmem + (static_cast<uint32_t>(quad * 32) << 16) + acc * 32, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);
        if (live) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t mw[4] = {mask[j].x, mask[j].y, mask[j].z, mask[j].w};
            uint32_t ow[4];
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              // a1 is a ReLU output (>= 0): "positive" == non-zero bit pattern apart from the sign of -0
              const float lo = (mw[h] & 0x7FFFu) ? __uint_as_float(v[j * 8 + 2 * h]) : 0.f;
              const float hi = (mw[h] & 0x7FFF0000u) ? __uint_as_float(v[j * 8 + 2 * h + 1]) : 0.f;
              ow[h] = pack_bf16x2(lo, hi);
            }
            *reinterpret_cast<uint4*>(p.out + o + j * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
          }
        }
      } else {
        mbar_wait(&tfull[acc], aph);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tmem + (static_cast<uint32_t>(quad * 32) << 16) + acc * 32, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);
        if (m < ROWS) {
#pragma unroll
          for (int c = 0; c < 32; ++c) epi[m * EPI_LD + c] = __uint_as_float(v[c]);
        }
        named_bar_sync(1, 128);
        // 24 pooled pixels x 32 channels: warp = pooled pixel (6 each), lane = channel
        const float bc = p.bias[lane];
        for (int pp = ew; pp < 24; pp += 4) {
          const int prow = pp / PL, pcol = pp - prow * PL;
          const int m00 = prow * 2 * WID + 2 * pcol;
          const float q[4] = {epi[m00 * EPI_LD + lane], epi[(m00 + 1) * EPI_LD + lane], epi[(m00 + WID) * EPI_LD + lane],
                              epi[(m00 + WID + 1) * EPI_LD + lane]};
          float best = q[0];
          uint32_t arg = 0;
#pragma unroll
          for (int k = 1; k < 4; ++k)
            if (q[k] > best) { best = q[k]; arg = k; }      // first maximum wins
          const size_t P = ((size_t)b * PL + (y0 >> 1) + prow) * PL + pcol;
          p.out[P * CH + lane] = __float2bfloat16(fmaxf(best + bc, 0.f));
          p.pidx[P * CH + lane] = static_cast<uint8_t>(arg);
        }
        named_bar_sync(1, 128);                             // epi is rewritten by the next band
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 64);
}


// ---------------------------------------------------------------------------------------------------------------
// conv2 weight gradient: dW[(tap, ci), co] = sum over pixels of a1[b, y+kh, x+kw, ci] * dz2[b, y, x, co].
// K = pixels.  A band of 4 output rows (96 pixels = six K16 steps) needs the same nine shifted boxes of a1 as the
// forward pass plus ONE box of dz2; here they are MN-major operands (K = pixel rows of 64 bytes, 32 channels each):
// A tile j = taps 4j..4j+3 (M = 128, LBO = one tap block), B = the dz2 band (N = 32).  A persistent CTA accumulates
// all its bands into three TMEM accumulators (M = 288 padded to 384) and writes ONE partial dW at the end; the existing
// sum_partials kernel adds the per-CTA partials (no atomics).
constexpr int WG_TAPS = 12;                              // 9 real tap blocks + 3 zero blocks so that M tiles are whole
constexpr int WG_STAGE_BYTES = (WG_TAPS + 1) * TAP_BYTES;   // + the dz2 band
constexpr int WG_SMEM = 2 * WG_STAGE_BYTES + 1024;

__global__ void __launch_bounds__(CV_THREADS, 1)
conv2_wgrad_tma_kernel(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tz, float* partial,
                       int tiles) {
  constexpr uint32_t BOX_BYTES = 96 * 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t full[2], empty[2], done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&done, 1);
    mbar_fence_init();
    tma_prefetch_desc(&ta); tma_prefetch_desc(&tz);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 128);
  // tap blocks 9..11 (and pixel rows 96..127 of every block, never read: K = 96) — zero the padding taps once
  for (int i = threadIdx.x; i < 2 * 3 * (TAP_BYTES / 16); i += blockDim.x) {
    const int s = i / (3 * (TAP_BYTES / 16)), r = i - s * 3 * (TAP_BYTES / 16);
    *reinterpret_cast<uint4*>(smem + s * WG_STAGE_BYTES + 9 * TAP_BYTES + r * 16) = make_uint4(0u, 0u, 0u, 0u);
  }
  pdl_wait();                                         // everything above overlapped the previous kernel's tail
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const int my_tiles = (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
        const int b = tile / 6, y0 = (tile - b * 6) * 4;
        const int s = it & 1, ph = (it >> 1) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], 10 * BOX_BYTES);
        const uint32_t base = smem_u32(smem + s * WG_STAGE_BYTES);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int kh = tap / 3, kw = tap - kh * 3;
          tma_load_4d(base + tap * TAP_BYTES, &ta, &full[s], 0, kw, y0 + kh, b);
        }
        tma_load_4d(base + WG_TAPS * TAP_BYTES, &tz, &full[s], 0, 0, y0, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 32, true, true);
      int it = 0;
      for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
        const int s = it & 1, ph = (it >> 1) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t base = smem_u32(smem + s * WG_STAGE_BYTES);
        const uint32_t b_base = base + WG_TAPS * TAP_BYTES;
#pragma unroll
        for (int mt = 0; mt < 3; ++mt)
#pragma unroll
          for (int k16 = 0; k16 < 6; ++k16) {
            const uint32_t a_lo = umma_desc_lo(base + mt * 4 * TAP_BYTES + k16 * 1024, TAP_BYTES);
            const uint32_t b_lo = umma_desc_lo(b_base + k16 * 1024, TAP_BYTES);
            umma_bf16_lohi(tmem + mt * 32, a_lo, SW64_DESC_HI, b_lo, SW64_DESC_HI, idesc, it > 0 || k16 > 0);
          }
        umma_commit(&empty[s]);
      }
      umma_commit(&done);
    }
  } else {
    const int quad = warp & 3;
    float* dst = partial + (size_t)blockIdx.x * K2 * CH;
    if (my_tiles > 0) {
      mbar_wait(&done, 0);
      tc_fence_after();
    }
#pragma unroll
    for (int mt = 0; mt < 3; ++mt) {
      uint32_t v[32];
      if (my_tiles > 0) {
        tmem_ld32(tmem + (static_cast<uint32_t>(quad * 32) << 16) + mt * 32, v);
        tmem_ld_wait();
      }
      const int m = mt * 128 + quad * 32 + lane;          // = tap*32 + ci
      if (m < K2) {
        float4* o = reinterpret_cast<float4*>(dst + (size_t)m * CH);
#pragma unroll
        for (int c = 0; c < 8; ++c)
          o[c] = my_tiles > 0 ? make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]),
                                            __uint_as_float(v[4 * c + 2]), __uint_as_float(v[4 * c + 3]))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

}  // namespace ddls

using namespace ddls;

namespace {
int conv_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

int nhwc_map(CUtensorMap* out, const void* base, int B, int HW, int box_w) {
  const uint64_t dims[4] = {(uint64_t)CH, (uint64_t)HW, (uint64_t)HW, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)CH * 2, (uint64_t)CH * 2 * HW, (uint64_t)CH * 2 * HW * HW};
  const uint32_t box[4] = {(uint32_t)CH, (uint32_t)box_w, 4, 1};
  return tmap_bf16(out, base, 4, dims, strides, box, 2);
}
}  // namespace

extern "C" int ddls_cnn_conv2_pool_fwd(const void* a1, const float* w2, const float* b2, void* pooled, void* pidx, int B,
                                           void* stream) {
  if (B <= 0) return 0;
  CUtensorMap tm;
  if (nhwc_map(&tm, a1, B, C1, C2)) return 1;
  ConvParams p{w2, b2, nullptr, static_cast<__nv_bfloat16*>(pooled), static_cast<uint8_t*>(pidx), B, B * 6};
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv2_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CV_SMEM));
  const int grid = p.tiles < conv_sm_count() ? p.tiles : conv_sm_count();
  DDLS_CUDA_OK(launch_k(conv2_tma_kernel<false>, dim3(grid), dim3(CV_THREADS), CV_SMEM, static_cast<cudaStream_t>(stream), tm, p));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "conv2_tma_kernel<fwd> launch");
}

extern "C" int ddls_cnn_conv2_dgrad(const void* dz2, const float* w2, const void* a1, void* dz1, int B, void* stream) {
  if (B <= 0) return 0;
  CUtensorMap tm;
  if (nhwc_map(&tm, dz2, B, C2, C1)) return 1;
  ConvParams p{w2, nullptr, static_cast<const __nv_bfloat16*>(a1), static_cast<__nv_bfloat16*>(dz1), nullptr, B, B * 7};
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv2_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CV_SMEM));
  const int grid = p.tiles < conv_sm_count() ? p.tiles : conv_sm_count();
  DDLS_CUDA_OK(launch_k(conv2_tma_kernel<true>, dim3(grid), dim3(CV_THREADS), CV_SMEM, static_cast<cudaStream_t>(stream), tm, p));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "conv2_tma_kernel<dgrad> launch");
}

namespace ddls {
__global__ void sum_partials_kernel(const float* __restrict__ partial, float* __restrict__ out, int n, int splits);
}

extern "C" size_t ddls_cnn_conv2_wgrad_scratch_bytes(int B) {
  // one partial dW (36,864 bytes) per persistent CTA; fewer bytes are accepted and simply run fewer CTAs
  const int tiles = B * 6;
  const int grid = tiles < conv_sm_count() ? tiles : conv_sm_count();
  return (size_t)(grid < 1 ? 1 : grid) * K2 * CH * sizeof(float);
}

extern "C" int ddls_cnn_conv2_wgrad(const void* a1, const void* dz2, float* dw2, void* scratch, size_t scratch_bytes, int B,
                                        void* stream) {
  if (B <= 0) return 0;
  const int tiles = B * 6;
  const size_t per = (size_t)K2 * CH * sizeof(float);
  int grid = tiles < conv_sm_count() ? tiles : conv_sm_count();
  if (scratch == nullptr || scratch_bytes < per) {
    set_last_error("ddls_cnn_conv2_wgrad: needs at least %zu bytes of scratch (got %zu)", per, scratch_bytes);
    return 1;
  }
  if ((size_t)grid * per > scratch_bytes) grid = (int)(scratch_bytes / per);
  CUtensorMap ta, tz;
  if (nhwc_map(&ta, a1, B, C1, C2) || nhwc_map(&tz, dz2, B, C2, C2)) return 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DDLS_CUDA_OK(cudaFuncSetAttribute(conv2_wgrad_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
  DDLS_CUDA_OK(launch_k(conv2_wgrad_tma_kernel, dim3(grid), dim3(CV_THREADS), WG_SMEM, st, ta, tz, static_cast<float*>(scratch), tiles));
  const int n = K2 * CH;
  DDLS_CUDA_OK(launch_k(sum_partials_kernel, dim3((n * 8 + 255) / 256), dim3(256), 0, st, static_cast<const float*>(scratch), dw2, n, grid));
  count_launch(2);
  return check_cuda(cudaGetLastError(), "conv2_wgrad_tma launch");
}
