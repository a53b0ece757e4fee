// Persistent GRU(128) recurrence on tcgen05 — forward and BPTT (K8/K9 of SURVEY.md §2.5).
//
// Keras-2 GRU with reset_after=False as the reference builds it (`GRU(128, input_shape=(25,1))`,
// /root/reference/ddl_nyiso_aztk.py:201-203): gate order z,r,h, hard_sigmoid gates,
//   z = hs(x·W_z + h·U_z + b_z);  r = hs(x·W_r + h·U_r + b_r);
//   hh = tanh(x·W_h + (r⊙h)·U_h + b_h);  h' = z⊙h + (1−z)⊙hh
// The reset gate multiplies h BEFORE the recurrent matmul, so every step is two dependent tensor-core
// phases: [z|r] from h (M = 2·128, K = 128), then hh from r⊙h (M = 128, K = 128).  Same layout
// strategy as csrc/lstm.cu: weights on the UMMA M side (resident bf16 in shared memory for all T,
// read in place from the Keras (H,3H) layout as an MN-major operand), the batch tile on the N side,
// accumulators in TMEM, TMEM lane = hidden unit so the owning thread does the gate math and keeps
// h in registers.
#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int GH = 128;
constexpr int GG = 3 * GH;
constexpr int G_THREADS = 256;

struct GruFwdParams {
  const float* x;        // [B, T]
  const float* w_ih;     // [3H]
  const float* w_hh;     // [H, 3H]
  const float* bias;     // [3H]
  float* out;            // [B, H] or [B, T, H]
  __nv_bfloat16* hs;     // [T, B, H]  h_t
  __nv_bfloat16* rh;     // [T, B, H]  r_t ⊙ h_{t-1}
  float4* gates;         // [T, B, H]  (z, r, hh, h_{t-1})
  int B, T;
  int return_sequences;
  int save_for_backward;
};

template <int NB>
__global__ void __launch_bounds__(G_THREADS, 1) gru_fwd_kernel(GruFwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sW = smem;                          // Uᵀ MN-major A: [16][48][8 k][8 m] bf16 = 96 KB
  uint8_t* sH = sW + GH * GG * 2;              // h      MN-major B: [16][NB/8][8][8]
  uint8_t* sR = sH + GH * NB * 2;              // r ⊙ h  MN-major B
  float* sX = reinterpret_cast<float*>(sR + GH * NB * 2);   // [NB][T]
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t TM_COLS = NB <= 32 ? 128 : (NB <= 64 ? 256 : 512);
  constexpr int CPT = NB / 2;
  constexpr uint32_t W_LBO = (GG / 8) * 128;   // k-group stride of the weight image

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hc = warp >> 2;
  const int j = q * 32 + lane;
  const int b_tile0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  if (warp == 0) tmem_alloc(&tmem_base_s, TM_COLS);
  if (tid == 32) { mbar_init(&mma_bar, 1); mbar_fence_init(); }

  for (int idx = tid; idx < GH * (GG / 8); idx += G_THREADS) {   // 128 x 48 vectors
    const int blk = idx >> 5, ln = idx & 31;
    const int kg = blk & 15, mg4 = blk >> 4;                      // 16 k-groups x 12 blocks of 4 m-groups
    const int k = kg * 8 + (ln & 7), mg = mg4 * 4 + (ln >> 3);
    const float4* src = reinterpret_cast<const float4*>(p.w_hh + (size_t)k * GG + mg * 8);
    const float4 a = src[0], b = src[1];
    *reinterpret_cast<uint4*>(sW + kg * W_LBO + mg * 128 + (k & 7) * 16) =
        make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
  for (int idx = tid; idx < NB * T; idx += G_THREADS) {
    const int b = idx / T, t = idx - b * T;
    sX[idx] = (b_tile0 + b < B) ? p.x[(size_t)(b_tile0 + b) * T + t] : 0.f;
  }
  float wi[3], bi[3];
#pragma unroll
  for (int g = 0; g < 3; ++g) { wi[g] = p.w_ih[g * GH + j]; bi[g] = p.bias[g * GH + j]; }

  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
  constexpr uint32_t idesc = umma_idesc_bf16(128, NB, true, true);
  const uint32_t w_base = smem_u32(sW), h_base = smem_u32(sH), r_base = smem_u32(sR);

  float hreg[CPT], zreg[CPT];
#pragma unroll
  for (int i = 0; i < CPT; ++i) hreg[i] = 0.f;

  uint32_t phase = 0;
  for (int t = 0; t < T; ++t) {
    const bool last = (t == T - 1);
    // ---------------- phase A: z, r ----------------
    if (t > 0) {
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int g = 0; g < 2; ++g)
#pragma unroll
          for (int ks = 0; ks < GH / 16; ++ks)
            umma_bf16(tmem + g * NB, umma_desc(w_base + g * 16 * 128 + ks * 2 * W_LBO, W_LBO, 128),
                      umma_desc(h_base + ks * 2 * (NB * 16), NB * 16, 128), idesc, ks > 0);
        umma_commit(&mma_bar);
      }
      mbar_wait(&mma_bar, phase);
      phase ^= 1;
      tc_fence_after();
    }
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;
      uint32_t vz[8], vr[8];
      if (t > 0) {
        tmem_ld8(t_lane + 0 * NB + col0, vz);
        tmem_ld8(t_lane + 1 * NB + col0, vr);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) vz[e] = vr[e] = 0u;
      }
      float rhv[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bl = col0 + e;
        const float xv = sX[bl * T + t];
        const float z = hard_sigmoid(__uint_as_float(vz[e]) + fmaf(xv, wi[0], bi[0]));
        const float r = hard_sigmoid(__uint_as_float(vr[e]) + fmaf(xv, wi[1], bi[1]));
        zreg[ch * 8 + e] = z;
        rhv[e] = r * hreg[ch * 8 + e];
        const int bglob = b_tile0 + bl;
        if (p.save_for_backward && bglob < B) {
          const size_t o = ((size_t)t * B + bglob) * GH + j;
          // hh (.z) is filled in by phase B below
          p.gates[o] = make_float4(z, r, 0.f, hreg[ch * 8 + e]);
          p.rh[o] = __float2bfloat16(rhv[e]);
        }
      }
      if (t > 0)
        *reinterpret_cast<uint4*>(sR + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) =
            make_uint4(pack_bf16x2(rhv[0], rhv[1]), pack_bf16x2(rhv[2], rhv[3]), pack_bf16x2(rhv[4], rhv[5]),
                       pack_bf16x2(rhv[6], rhv[7]));
    }
    // ---------------- phase B: candidate hh from r ⊙ h ----------------
    if (t > 0) {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < GH / 16; ++ks)
          umma_bf16(tmem + 2 * NB, umma_desc(w_base + 2 * 16 * 128 + ks * 2 * W_LBO, W_LBO, 128),
                    umma_desc(r_base + ks * 2 * (NB * 16), NB * 16, 128), idesc, ks > 0);
        umma_commit(&mma_bar);
      }
      mbar_wait(&mma_bar, phase);
      phase ^= 1;
      tc_fence_after();
    }
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;
      uint32_t vh[8];
      if (t > 0) { tmem_ld8(t_lane + 2 * NB + col0, vh); tmem_ld_wait(); }
      else {
#pragma unroll
        for (int e = 0; e < 8; ++e) vh[e] = 0u;
      }
      float hv[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bl = col0 + e;
        const float xv = sX[bl * T + t];
        const float hh = fast_tanh(__uint_as_float(vh[e]) + fmaf(xv, wi[2], bi[2]));
        const float z = zreg[ch * 8 + e];
        const float hn = fmaf(z, hreg[ch * 8 + e] - hh, hh);      // z*h + (1-z)*hh
        hreg[ch * 8 + e] = hn;
        hv[e] = hn;
        const int bglob = b_tile0 + bl;
        if (bglob < B) {
          if (p.save_for_backward) {
            const size_t o = ((size_t)t * B + bglob) * GH + j;
            reinterpret_cast<float*>(p.gates + o)[2] = hh;
            p.hs[o] = __float2bfloat16(hn);
          }
          if (p.return_sequences) p.out[((size_t)bglob * T + t) * GH + j] = hn;
          else if (last) p.out[(size_t)bglob * GH + j] = hn;
        }
      }
      if (!last)
        *reinterpret_cast<uint4*>(sH + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) =
            make_uint4(pack_bf16x2(hv[0], hv[1]), pack_bf16x2(hv[2], hv[3]), pack_bf16x2(hv[4], hv[5]),
                       pack_bf16x2(hv[6], hv[7]));
    }
    if (!last) {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, TM_COLS);
}

// ------------------------------------------------------------------------------------------------
struct GruBwdParams {
  const float* dout;       // [B, H] or [B, T, H]
  const float* x;          // [B, T]
  const float* w_hh;       // [H, 3H]
  const float4* gates;     // [T, B, H] (z, r, hh, h_{t-1})
  __nv_bfloat16* da;       // [T, B, 3H]  d(az), d(ar), d(ah)
  float* d_wih;            // [3H] (atomic accumulate)
  float* d_bias;           // [3H] (atomic accumulate)
  int B, T;
  int return_sequences;
};

template <int NB>
__global__ void __launch_bounds__(G_THREADS, 1) gru_bwd_kernel(GruBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sUzr = smem;                         // U[:, 0:2H] K-major A: [K/8=32][M=128][16 B] = 64 KB
  uint8_t* sUh = sUzr + GH * 2 * GH * 2;        // U[:, 2H:3H] K-major A: [16][128][16 B]      = 32 KB
  uint8_t* sB1 = sUh + GH * GH * 2;             // d(ah)ᵀ        MN-major B, K = 128
  uint8_t* sB2 = sB1 + GH * NB * 2;             // [d(az)|d(ar)]ᵀ MN-major B, K = 256
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t TM_COLS = 2 * NB < 32 ? 32 : 2 * NB;
  constexpr int CPT = NB / 2;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hc = warp >> 2;
  const int j = q * 32 + lane;
  const int b_tile0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  if (warp == 0) tmem_alloc(&tmem_base_s, TM_COLS);
  if (tid == 32) { mbar_init(&mma_bar, 1); mbar_fence_init(); }

  // stage U (fp32, row m = h_{t-1} unit, 3H columns) into the two K-major images
  for (int idx = tid; idx < GH * (GG / 8); idx += G_THREADS) {     // 128 rows x 48 chunks of 8 columns
    const int blk = idx >> 5, ln = idx & 31;
    const int mblk = blk & 15, kc4 = blk >> 4;                      // 16 row blocks x 12 blocks of 4 chunks
    const int m = mblk * 8 + (ln & 7), kc = kc4 * 4 + (ln >> 3);   // kc in [0,48)
    const float4* src = reinterpret_cast<const float4*>(p.w_hh + (size_t)m * GG + kc * 8);
    const float4 a = src[0], b = src[1];
    const uint4 v = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    if (kc < 32) *reinterpret_cast<uint4*>(sUzr + kc * (GH * 16) + m * 16) = v;
    else         *reinterpret_cast<uint4*>(sUh + (kc - 32) * (GH * 16) + m * 16) = v;
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
  constexpr uint32_t idesc = umma_idesc_bf16(128, NB, false, true);
  const uint32_t uzr_base = smem_u32(sUzr), uh_base = smem_u32(sUh), b1_base = smem_u32(sB1), b2_base = smem_u32(sB2);

  float dh[CPT];
#pragma unroll
  for (int i = 0; i < CPT; ++i) dh[i] = 0.f;
  float acc_w[3] = {0.f, 0.f, 0.f}, acc_b[3] = {0.f, 0.f, 0.f};
  uint32_t phase = 0;

  for (int t = T - 1; t >= 0; --t) {
    // ---- stage 1: dh_t complete -> d(ah), d(az); MMA1: drh = d(ah)·U_hᵀ -------------------------------
    float daz_r[CPT], pdh[CPT];   // d(az) and the partial dh_{t-1} = dh ⊙ z carried to stage 2
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;
      float dah[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bl = col0 + e, bglob = b_tile0 + bl;
        float v_dah = 0.f, v_daz = 0.f, v_pdh = 0.f;
        if (bglob < B) {
          const float4 g = p.gates[((size_t)t * B + bglob) * GH + j];   // z, r, hh, h_prev
          float d = dh[ch * 8 + e];
          if (p.return_sequences) d += p.dout[((size_t)bglob * T + t) * GH + j];
          else if (t == T - 1) d += p.dout[(size_t)bglob * GH + j];
          v_dah = d * (1.f - g.x) * (1.f - g.z * g.z);
          v_daz = d * (g.w - g.z) * ((g.x > 0.f && g.x < 1.f) ? 0.2f : 0.f);
          v_pdh = d * g.x;
        }
        dah[e] = v_dah; daz_r[ch * 8 + e] = v_daz; pdh[ch * 8 + e] = v_pdh;
      }
      if (t > 0)
        *reinterpret_cast<uint4*>(sB1 + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) =
            make_uint4(pack_bf16x2(dah[0], dah[1]), pack_bf16x2(dah[2], dah[3]), pack_bf16x2(dah[4], dah[5]),
                       pack_bf16x2(dah[6], dah[7]));
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bglob = b_tile0 + col0 + e;
        if (bglob < B) {
          const float xv = p.x[(size_t)bglob * T + t];
          acc_w[0] = fmaf(xv, daz_r[ch * 8 + e], acc_w[0]); acc_b[0] += daz_r[ch * 8 + e];
          acc_w[2] = fmaf(xv, dah[e], acc_w[2]);            acc_b[2] += dah[e];
          __nv_bfloat16* row = p.da + ((size_t)t * B + bglob) * GG + j;
          row[0] = __float2bfloat16(daz_r[ch * 8 + e]);
          row[2 * GH] = __float2bfloat16(dah[e]);
        }
      }
    }
    if (t == 0) {
      // h_{-1} = 0: d(ar) = 0 and nothing propagates further back
#pragma unroll
      for (int ch = 0; ch < CPT / 8; ++ch)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int bglob = b_tile0 + hc * CPT + ch * 8 + e;
          if (bglob < B) p.da[((size_t)t * B + bglob) * GG + GH + j] = __float2bfloat16(0.f);
        }
      break;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < GH / 16; ++ks)
        umma_bf16(tmem, umma_desc(uh_base + ks * 2 * (GH * 16), GH * 16, 128),
                  umma_desc(b1_base + ks * 2 * (NB * 16), NB * 16, 128), idesc, ks > 0);
      umma_commit(&mma_bar);
    }
    mbar_wait(&mma_bar, phase);
    phase ^= 1;
    tc_fence_after();
    // ---- stage 2: drh -> d(ar); MMA2: [d(az)|d(ar)]·U_zrᵀ -------------------------------------------------
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;
      uint32_t vd[8];
      tmem_ld8(t_lane + col0, vd);
      tmem_ld_wait();
      float dar[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bglob = b_tile0 + col0 + e;
        float v = 0.f;
        if (bglob < B) {
          const float4 g = p.gates[((size_t)t * B + bglob) * GH + j];
          const float drh = __uint_as_float(vd[e]);
          v = drh * g.w * ((g.y > 0.f && g.y < 1.f) ? 0.2f : 0.f);
          pdh[ch * 8 + e] = fmaf(drh, g.y, pdh[ch * 8 + e]);
          const float xv = p.x[(size_t)bglob * T + t];
          acc_w[1] = fmaf(xv, v, acc_w[1]); acc_b[1] += v;
          p.da[((size_t)t * B + bglob) * GG + GH + j] = __float2bfloat16(v);
        }
        dar[e] = v;
      }
      const uint32_t off = (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16;
      *reinterpret_cast<uint4*>(sB2 + off) =
          make_uint4(pack_bf16x2(daz_r[ch * 8 + 0], daz_r[ch * 8 + 1]), pack_bf16x2(daz_r[ch * 8 + 2], daz_r[ch * 8 + 3]),
                     pack_bf16x2(daz_r[ch * 8 + 4], daz_r[ch * 8 + 5]), pack_bf16x2(daz_r[ch * 8 + 6], daz_r[ch * 8 + 7]));
      *reinterpret_cast<uint4*>(sB2 + 16 * (NB * 16) + off) =
          make_uint4(pack_bf16x2(dar[0], dar[1]), pack_bf16x2(dar[2], dar[3]), pack_bf16x2(dar[4], dar[5]),
                     pack_bf16x2(dar[6], dar[7]));
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < 2 * GH / 16; ++ks)
        umma_bf16(tmem + NB, umma_desc(uzr_base + ks * 2 * (GH * 16), GH * 16, 128),
                  umma_desc(b2_base + ks * 2 * (NB * 16), NB * 16, 128), idesc, ks > 0);
      umma_commit(&mma_bar);
    }
    mbar_wait(&mma_bar, phase);
    phase ^= 1;
    tc_fence_after();
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;
      uint32_t vd[8];
      tmem_ld8(t_lane + NB + col0, vd);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 8; ++e) dh[ch * 8 + e] = pdh[ch * 8 + e] + __uint_as_float(vd[e]);
    }
    tc_fence_before();
    __syncthreads();   // TMEM reads done before the next iteration's MMA1 overwrites column block 0
  }
#pragma unroll
  for (int g = 0; g < 3; ++g) {
    atomicAdd(p.d_wih + g * GH + j, acc_w[g]);
    atomicAdd(p.d_bias + g * GH + j, acc_b[g]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, TM_COLS);
}

template <int NB>
static int launch_gru_fwd(const GruFwdParams& p, cudaStream_t st) {
  const size_t smem = (size_t)GH * GG * 2 + 2 * (size_t)GH * NB * 2 + (size_t)NB * p.T * 4;
  if (smem > 227 * 1024) { set_last_error("ddls_gru128_forward: T=%d too long for a %d-row tile", p.T, NB); return 1; }
  auto kern = gru_fwd_kernel<NB>;
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(p.B + NB - 1) / NB, G_THREADS, smem, st>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "gru_fwd_kernel launch");
}

template <int NB>
static int launch_gru_bwd(const GruBwdParams& p, cudaStream_t st) {
  const size_t smem = (size_t)GH * GG * 2 + 3 * (size_t)GH * NB * 2;
  auto kern = gru_bwd_kernel<NB>;
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(p.B + NB - 1) / NB, G_THREADS, smem, st>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "gru_bwd_kernel launch");
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_gru128_forward(const float* x, const float* w_ih, const float* w_hh, const float* bias, float* out,
                                   void* hs, void* rh, void* gates, int B, int T, int return_sequences,
                                   int save_for_backward, int tile, void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_gru128_forward: empty problem"); return 1; }
  GruFwdParams p{x, w_ih, w_hh, bias, out, static_cast<__nv_bfloat16*>(hs), static_cast<__nv_bfloat16*>(rh),
                 static_cast<float4*>(gates), B, T, return_sequences, save_for_backward};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tile <= 0) tile = (B <= 32 * 148) ? 32 : (B <= 64 * 148 ? 64 : 128);
  if (tile == 128 && (size_t)128 * T * 4 > 60 * 1024) tile = 64;
  switch (tile) {
    case 32: return launch_gru_fwd<32>(p, st);
    case 64: return launch_gru_fwd<64>(p, st);
    case 128: return launch_gru_fwd<128>(p, st);
  }
  set_last_error("ddls_gru128_forward: tile must be 32, 64 or 128");
  return 1;
}

extern "C" int ddls_gru128_backward(const float* dout, const float* x, const float* w_hh, const void* gates, void* da,
                                    float* d_wih, float* d_bias, int B, int T, int return_sequences, int tile,
                                    void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_gru128_backward: empty problem"); return 1; }
  GruBwdParams p{dout, x, w_hh, static_cast<const float4*>(gates), static_cast<__nv_bfloat16*>(da), d_wih, d_bias,
                 B, T, return_sequences};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tile <= 0) tile = (B <= 32 * 148) ? 32 : 64;
  switch (tile) {
    case 32: return launch_gru_bwd<32>(p, st);
    case 64: return launch_gru_bwd<64>(p, st);
  }
  set_last_error("ddls_gru128_backward: tile must be 32 or 64");
  return 1;
}
