// Persistent recurrent kernels, second generation (LSTM(128) / GRU(128), input dim 1).
//
// Same "weights on M, batch on N" tcgen05 design as csrc/lstm.cu / csrc/gru.cu (which remain as the
// first-generation reference implementation), re-engineered after profiling showed the step loop was
// bound by instruction count and exposed global latency, not by the tensor pipe:
//   * the recurrent weights, the input kernel and the bias are packed ONCE per optimizer step into a bf16
//     UMMA image (ddls_rnn128_pack) that every CTA pulls into shared memory with one bulk TMA copy
//     (cp.async.bulk -> UBLKCP) — the same bytes serve the forward (MN-major A) and the BPTT (K-major A);
//   * the rank-1 input projection and the bias ride in the MMA as one extra K = 16 step: A rows
//     [w_hi, w_lo, b_hi, b_lo, w_hi], B rows [x_hi, x_hi, 1, 1, x_lo] (hi/lo bf16 splits keep ~fp32 accuracy),
//     so the epilogue is only the gate non-linearities (hard_sigmoid = one FFMA.SAT);
//   * 512 threads per CTA (4 TMEM lane quarters x 4 column groups) for latency hiding;
//   * everything saved for BPTT is laid out per thread as 8 consecutive batch rows (16/32-byte vector
//     stores, coalesced across the hidden-unit lanes); h_t and dz_t are stored directly as K-major UMMA
//     images, so the recurrent weight-gradient GEMM (ddls_rnn128_wgrad) streams them with bulk TMA copies
//     through a 4-stage mbarrier pipeline with a dedicated producer thread — no staging pass at all.
// Keras semantics as in the first generation: LSTM gates i,f,c,o, GRU gates z,r,h with reset applied
// before the recurrent matmul (/root/reference/ddl_nyiso_aztk.py:201-203,249-251).
#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int RH = 128;             // hidden units
constexpr int R_KG = 18;            // k-groups in the weight image: 16 (hidden) + 1 (input/bias rows) + 1 (zeros)

// image element (k, m): ((k>>3) * (GH/8) + (m>>3)) * 64 + (k&7)*8 + (m&7)   [bf16 elements]
__global__ void __launch_bounds__(256) rnn128_pack_kernel(const float* __restrict__ rec, const float* __restrict__ wih,
                                                          const float* __restrict__ bias, __nv_bfloat16* __restrict__ img,
                                                          int GH) {
  pdl_wait();
  const int MG = GH / 8;
  const int total = R_KG * MG * 8;                  // one item = (k, m-group): 8 consecutive m
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int kr = idx & 7, mg = (idx >> 3) % MG, kg = idx / (8 * MG);
    const int k = kg * 8 + kr;
    float v[8];
    if (k < RH) {
      const float4* s = reinterpret_cast<const float4*>(rec + (size_t)k * GH + mg * 8);
      const float4 a = s[0], b = s[1];
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
      const int row = k - RH;                       // 0: w_hi 1: w_lo 2: b_hi 3: b_lo 4: w_hi, else 0
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float w = wih[mg * 8 + e], bb = bias[mg * 8 + e];
        const float whi = __bfloat162float(__float2bfloat16(w)), bhi = __bfloat162float(__float2bfloat16(bb));
        v[e] = (row == 0 || row == 4) ? whi : row == 1 ? (w - whi) : row == 2 ? bhi : row == 3 ? (bb - bhi) : 0.f;
      }
    }
    *reinterpret_cast<uint4*>(img + ((size_t)(kg * MG + mg) * 64 + kr * 8)) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}

// x image for the extra K step: per time step one k-group [NB/8][8 k][8 n]: rows x_hi, x_hi, 1, 1, x_lo, 0, 0, 0
template <int NB>
DDLS_DEVINL void build_x_image(uint8_t* sX, uint8_t* sZero, const float* __restrict__ x, int b0, int B, int T) {
  for (int idx = threadIdx.x; idx < T * NB; idx += blockDim.x) {
    const int t = idx / NB, b = idx - t * NB;
    const float xv = (b0 + b < B) ? x[(size_t)(b0 + b) * T + t] : 0.f;
    const __nv_bfloat16 hi = __float2bfloat16(xv);
    const __nv_bfloat16 lo = __float2bfloat16(xv - __bfloat162float(hi));
    const __nv_bfloat16 one = __float2bfloat16(1.f), zero = __float2bfloat16(0.f);
    __nv_bfloat16* g = reinterpret_cast<__nv_bfloat16*>(sX + (size_t)t * (NB * 16) + (b >> 3) * 128) + (b & 7);
    g[0 * 8] = hi; g[1 * 8] = hi; g[2 * 8] = one; g[3 * 8] = one; g[4 * 8] = lo; g[5 * 8] = zero; g[6 * 8] = zero; g[7 * 8] = zero;
  }
  for (int idx = threadIdx.x; idx < NB * 16 / 16; idx += blockDim.x) reinterpret_cast<uint4*>(sZero)[idx] = make_uint4(0, 0, 0, 0);
}

// Thread organisation of the recurrence kernels: warps 0..15 are the epilogue (TMEM lane quarter q = warp & 3,
// column group warp >> 2), warp 16 only issues MMAs.  Per step the epilogue (a) reads its accumulators, does the
// gate math and writes the next N-side operand to shared memory, (b) ARRIVES on a named barrier the MMA warp is
// waiting on, and only then (c) writes the backward stash to global memory — so (c) overlaps the next MMAs.
// Batch tile NB -> column groups: NB >= 32 uses 4 (16 epilogue warps), NB = 16 uses 2 (8 epilogue warps).  A batch of
// 32 therefore runs as two 16-row CTAs on two SMs: half the epilogue work and half the stash traffic per SM.
template <int NB>
struct RCfg {
  static constexpr int NCG = NB >= 32 ? 4 : 2;       // column groups
  static constexpr int EPI = NCG * 4 * 32;           // epilogue threads
  static constexpr int ALL = EPI + 32;               // + MMA warp
  static constexpr int MMA_WARP = NCG * 4;
  static constexpr int CH8 = NB / (8 * NCG);         // 8-column chunks per epilogue thread
};
constexpr uint32_t BAR_H = 1, BAR_R = 2;

DDLS_DEVINL uint4 pack8(const float (&v)[8]) {
  return make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}
DDLS_DEVINL void unpack8(const uint4& p, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&p);
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __low2float(h[i]); v[2 * i + 1] = __high2float(h[i]); }
}
// K-major view of the weight image (M = image row k, K = image column m): copy `nchunk16` x 16 columns of row
// `row`, starting at 8-column chunk `c8_0`, into TMEM columns [tcol, tcol + nchunk16*8) of the calling thread's lane.
template <int MG>
DDLS_DEVINL void image_row_to_tmem(const uint8_t* sW, int row, int c8_0, int nchunk32, uint32_t taddr) {
  const uint8_t* src = sW + (row >> 3) * (MG * 128) + (row & 7) * 16;
  for (int h = 0; h < nchunk32; ++h) {            // 32 TMEM columns = 64 bf16 = 8 chunks of 16 bytes
    uint32_t v[32];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint4 u = *reinterpret_cast<const uint4*>(src + (c8_0 + h * 8 + c) * 128);
      v[4 * c] = u.x; v[4 * c + 1] = u.y; v[4 * c + 2] = u.z; v[4 * c + 3] = u.w;
    }
    tmem_st32(taddr + h * 32, v);
  }
  tmem_st_wait();
}

// MN-major view (M = image column m, K = image row k): copy the 128 K-values of column `m` (gathered 2 bytes
// at a time from the image) into 64 TMEM columns of the calling thread's lane.
template <int MG>
DDLS_DEVINL void image_col_to_tmem(const uint8_t* sW, int m, uint32_t taddr) {
  const uint8_t* src = sW + (m >> 3) * 128 + (m & 7) * 2;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    uint32_t v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int k0 = half * 64 + 2 * i;
      const uint32_t lo = *reinterpret_cast<const uint16_t*>(src + (k0 >> 3) * (MG * 128) + (k0 & 7) * 16);
      const uint32_t hi16 = *reinterpret_cast<const uint16_t*>(src + ((k0 + 1) >> 3) * (MG * 128) + ((k0 + 1) & 7) * 16);
      v[i] = lo | (hi16 << 16);
    }
    tmem_st32(taddr + half * 32, v);
  }
  tmem_st_wait();
}

DDLS_DEVINL float hs_grad(float a) { return (a > 0.f && a < 1.f) ? 0.2f : 0.f; }   // hard_sigmoid' from its output

struct LstmV2Fwd {
  const float* x;                 // [B, T]
  const __nv_bfloat16* img;       // packed weights [18][64][8][8]
  float* out;                     // [B, H] or [B, T, H]
  uint4* hs;                      // [T][B8][128] x 16 B   (8 batch rows of h_t, bf16)  == K-major UMMA image
  uint4* gates;                   // [T][B8][4][128] x 16 B (post-activation i,f,g,o, bf16)
  float4* cs;                     // [T][B8][128] x 32 B   (c_t, fp32)
  int B, T, B8;
  int return_sequences, save;
};

template <int NB>
__global__ void __launch_bounds__(RCfg<NB>::ALL, 1) lstm_fwd_v2_kernel(LstmV2Fwd p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int GH = 4 * RH, MG = GH / 8;
  constexpr uint32_t W_LBO = MG * 128;                       // 8192: k-group stride of the image
  constexpr int W_BYTES = R_KG * MG * 128;                   // 147,456
  constexpr int CH8 = RCfg<NB>::CH8, R_ALL = RCfg<NB>::ALL, R_EPI = RCfg<NB>::EPI, NCG = RCfg<NB>::NCG;
  uint8_t* sW = smem;
  uint8_t* sH = sW + W_BYTES;                                // [16][NB/8][8][8] bf16
  uint8_t* sX = sH + RH * NB * 2;                            // [T][NB*16 B]
  uint8_t* sZero = sX + (size_t)p.T * NB * 16;               // one zero k-group
  __shared__ uint64_t bar_w, bar_mma;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  // TMEM: [0, 4*NB) accumulators, [A_COL0, A_COL0 + 256) the four gate tiles of Uᵀ as the resident A operand
  constexpr uint32_t A_COL0 = 4 * NB;
  constexpr uint32_t TM_COLS = (4 * NB + 256) <= 256 ? 256 : 512;
  if (warp == RCfg<NB>::MMA_WARP) tmem_alloc(&tmem_s, TM_COLS);
  if (tid == 0) { mbar_init(&bar_w, 1); mbar_init(&bar_mma, 1); mbar_fence_init(); }
  __syncthreads();
  pdl_wait();                                               // set-up above overlapped the previous kernel\'s drain
  if (tid == R_EPI) {                                        // bulk TMA copies bring the whole weight image
    mbar_expect_tx(&bar_w, W_BYTES);
#pragma unroll
    for (int c = 0; c < R_KG; ++c)
      tma_bulk_g2s(sW + c * (MG * 128), reinterpret_cast<const uint8_t*>(p.img) + (size_t)c * (MG * 128), MG * 128, &bar_w);
  }
  build_x_image<NB>(sX, sZero, p.x, b0, B, T);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;

  if (warp == RCfg<NB>::MMA_WARP) {
    // ================= MMA warp: the whole warp runs the (warp-uniform) loop, one elected lane issues =================
    constexpr uint32_t idesc = umma_idesc_bf16(128, NB, true, true);      // x step: A from the smem image (MN-major)
    constexpr uint32_t idesc_ts = umma_idesc_bf16(128, NB, false, true);  // h steps: A from TMEM is K-major by construction
    const uint32_t w_base = smem_u32(sW), h_base = smem_u32(sH), x_base = smem_u32(sX), z_base = smem_u32(sZero);
    // descriptor low words never change: build them once, keep them in registers
    constexpr uint32_t hi = umma_desc_hi(128);
    // (start addresses are < 256 KB, so adding a 16-byte-unit offset never carries into the LBO field)
    const uint32_t a_lo0 = umma_desc_lo(w_base, W_LBO), b_lo0 = umma_desc_lo(h_base, NB * 16);
    const bool leader = elect_one();
    mbar_wait(&bar_w, 0);
    named_bar_sync(BAR_R, R_ALL);                            // the epilogue warps have copied Uᵀ into TMEM
    for (int t = 0; t < T; ++t) {
      if (t > 0) named_bar_sync(BAR_H, R_ALL);               // h_{t-1} is in shared memory, TMEM has been drained
      tc_fence_after();
      if (leader) {
        const uint32_t xt = x_base + t * (NB * 16);
        const uint32_t x_lo = umma_desc_lo(xt, z_base - xt);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if (t > 0) {
#pragma unroll
            for (int ks = 0; ks < RH / 16; ++ks)      // A from TMEM: 8 columns (16 bf16) per K step
              umma_bf16_ts(tmem + g * NB, tmem + A_COL0 + g * 64 + ks * 8, b_lo0 + ((ks * 2 * (NB * 16)) >> 4), hi,
                           idesc_ts, ks > 0);
          }
          // input projection + bias: image k-groups 16 (payload) / 17 (zeros) times [x rows of step t | zero group]
          umma_bf16_lohi(tmem + g * NB, a_lo0 + ((g * 16 * 128 + 16 * W_LBO) >> 4), hi, x_lo, hi, idesc, t > 0);
        }
        umma_commit(&bar_mma);
      }
      __syncwarp();
    }
  } else {
    // ================= epilogue warps =================
    const int q = warp & 3, cgp = warp >> 2;
    const int j = q * 32 + lane;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
    {
      // Make the recurrent weights TMEM-resident: warp (q, g = cgp) copies gate tile g, rows 32q..32q+31.
      // Row m = j of tile g is Uᵀ[g*128 + j][k], k = 0..127, read from the shared-memory image.
      mbar_wait(&bar_w, 0);
      for (int g = cgp; g < 4; g += NCG) image_col_to_tmem<MG>(sW, g * RH + j, t_lane + A_COL0 + g * 64);
      tc_fence_before();
      named_bar_arrive(BAR_R, R_ALL);
    }
    float c[CH8 * 8];
#pragma unroll
    for (int i = 0; i < CH8 * 8; ++i) c[i] = 0.f;
    uint32_t phase = 0;
    for (int t = 0; t < T; ++t) {
      const bool last = (t == T - 1);
      mbar_wait(&bar_mma, phase);
      phase ^= 1;
      tc_fence_after();
      float gi[CH8][8], gf[CH8][8], gg[CH8][8], go[CH8][8], hv[CH8][8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        uint32_t vi[8], vf[8], vg[8], vo[8];
        tmem_ld8(t_lane + 0 * NB + col0, vi);
        tmem_ld8(t_lane + 1 * NB + col0, vf);
        tmem_ld8(t_lane + 2 * NB + col0, vg);
        tmem_ld8(t_lane + 3 * NB + col0, vo);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          gi[ch][e] = __saturatef(fmaf(0.2f, __uint_as_float(vi[e]), 0.5f));
          gf[ch][e] = __saturatef(fmaf(0.2f, __uint_as_float(vf[e]), 0.5f));
          gg[ch][e] = fast_tanh(__uint_as_float(vg[e]));
          go[ch][e] = __saturatef(fmaf(0.2f, __uint_as_float(vo[e]), 0.5f));
          const float cn = fmaf(gf[ch][e], c[ch * 8 + e], gi[ch][e] * gg[ch][e]);
          c[ch * 8 + e] = cn;
          hv[ch][e] = go[ch][e] * fast_tanh(cn);
        }
        if (!last) *reinterpret_cast<uint4*>(sH + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) = pack8(hv[ch]);
      }
      if (!last) {                                           // hand h_t to the MMA warp, then do the non-critical stores
        fence_proxy_async_smem();
        tc_fence_before();
        named_bar_arrive(BAR_H, R_ALL);
      }
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        const int bc = (b0 + col0) >> 3;
        if (bc < p.B8) {
          if (p.save) {
            const size_t base = (size_t)t * p.B8 + bc;
            p.hs[base * RH + j] = pack8(hv[ch]);
            uint4* gd = p.gates + base * 4 * RH + j;
            gd[0 * RH] = pack8(gi[ch]); gd[1 * RH] = pack8(gf[ch]); gd[2 * RH] = pack8(gg[ch]); gd[3 * RH] = pack8(go[ch]);
            float4* cd = p.cs + (base * RH + j) * 2;
            cd[0] = make_float4(c[ch * 8 + 0], c[ch * 8 + 1], c[ch * 8 + 2], c[ch * 8 + 3]);
            cd[1] = make_float4(c[ch * 8 + 4], c[ch * 8 + 5], c[ch * 8 + 6], c[ch * 8 + 7]);
          }
          if (p.return_sequences || last) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int bg = b0 + col0 + e;
              if (bg < B) {
                if (p.return_sequences) p.out[((size_t)bg * T + t) * RH + j] = hv[ch][e];
                else p.out[(size_t)bg * RH + j] = hv[ch][e];
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == RCfg<NB>::MMA_WARP) tmem_dealloc(tmem, TM_COLS);
}

struct LstmV2Bwd {
  const float* dout;              // [B, H] or [B, T, H]
  const float* x;                 // [B, T]
  const __nv_bfloat16* img;
  const uint4* gates;             // [T][B8][4][128]
  const float4* cs;               // [T][B8][128] x 2
  uint4* dz;                      // [T][B8][512] x 16 B  == K-major UMMA image (N = gate column)
  float* d_wih;                   // [4H] atomic accumulate
  float* d_bias;                  // [4H]
  int B, T, B8;
  int return_sequences;
};

template <int NB>
__global__ void __launch_bounds__(RCfg<NB>::ALL, 1) lstm_bwd_v2_kernel(LstmV2Bwd p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int GH = 4 * RH, MG = GH / 8;
  constexpr int W_BYTES = 16 * MG * 128;                     // hidden k-groups only: 131,072
  constexpr int CH8 = RCfg<NB>::CH8, R_ALL = RCfg<NB>::ALL, R_EPI = RCfg<NB>::EPI, NCG = RCfg<NB>::NCG;
  uint8_t* sW = smem;                                        // image as K-major A: LBO = 128, SBO = MG*128
  uint8_t* sZ = sW + W_BYTES;                                // dz MN-major B: [64 k-groups][NB/8][8][8]
  float* sX = reinterpret_cast<float*>(sZ + GH * NB * 2);    // [NB][T]
  __shared__ uint64_t bar_w, bar_mma;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;
  // TMEM: [0, NB) the dh accumulator, [A_COL0, A_COL0 + 256) U (128 x 512 bf16) as the resident A operand
  constexpr uint32_t A_COL0 = 64, TM_COLS = 512;

  if (warp == RCfg<NB>::MMA_WARP) tmem_alloc(&tmem_s, TM_COLS);
  if (tid == 0) { mbar_init(&bar_w, 1); mbar_init(&bar_mma, 1); mbar_fence_init(); }
  __syncthreads();
  pdl_wait();                                               // set-up above overlapped the previous kernel\'s drain
  if (tid == R_EPI) {
    mbar_expect_tx(&bar_w, W_BYTES);
#pragma unroll
    for (int c = 0; c < 16; ++c)
      tma_bulk_g2s(sW + c * (MG * 128), reinterpret_cast<const uint8_t*>(p.img) + (size_t)c * (MG * 128), MG * 128, &bar_w);
  }
  for (int idx = tid; idx < NB * T; idx += R_ALL) {
    const int b = idx / T, t = idx - b * T;
    sX[idx] = (b0 + b < B) ? p.x[(size_t)(b0 + b) * T + t] : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;

  if (warp == RCfg<NB>::MMA_WARP) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, NB, false, true);
    const uint32_t w_base = smem_u32(sW), z_base = smem_u32(sZ);
    constexpr uint32_t b_hi = umma_desc_hi(128);
    const uint32_t b_lo0 = umma_desc_lo(z_base, NB * 16);
    (void)w_base;
    const bool leader = elect_one();
    named_bar_sync(BAR_R, R_ALL);                            // the epilogue warps have copied U into TMEM
    for (int t = T - 1; t >= 1; --t) {
      named_bar_sync(BAR_H, R_ALL);                          // dz_t is in shared memory
      tc_fence_after();
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < GH / 16; ++ks)                 // A from TMEM: 8 columns (16 gate columns) per K step
          umma_bf16_ts(tmem, tmem + A_COL0 + ks * 8, b_lo0 + ((ks * 2 * (NB * 16)) >> 4), b_hi, idesc, ks > 0);
        umma_commit(&bar_mma);
      }
      __syncwarp();
    }
  } else {
    const int q = warp & 3, cgp = warp >> 2;
    const int j = q * 32 + lane;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
    // warp (q, cgp) copies gate block cgp (128 columns = 64 TMEM columns) of rows 32q..32q+31 into TMEM
    mbar_wait(&bar_w, 0);
    for (int g = cgp; g < 4; g += NCG) image_row_to_tmem<MG>(sW, j, g * 16, 2, t_lane + A_COL0 + g * 64);
    tc_fence_before();
    named_bar_arrive(BAR_R, R_ALL);
    float dc[CH8 * 8];
#pragma unroll
    for (int i = 0; i < CH8 * 8; ++i) dc[i] = 0.f;
    float acc_w[4] = {0.f, 0.f, 0.f, 0.f}, acc_b[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t phase = 0;
    for (int t = T - 1; t >= 0; --t) {
      const bool have_dh = (t < T - 1);
      // ---- this step's loads do not depend on the tensor pipe: issue them before waiting on it ----
      uint4 g4[CH8][4];
      float4 ct[CH8][2], cp[CH8][2];
      float dext[CH8][8];
      bool live[CH8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        const int bc = (b0 + col0) >> 3;
        live[ch] = bc < p.B8;
        if (live[ch]) {
          const size_t base = (size_t)t * p.B8 + bc;
          const uint4* gs = p.gates + base * 4 * RH + j;
          g4[ch][0] = gs[0]; g4[ch][1] = gs[RH]; g4[ch][2] = gs[2 * RH]; g4[ch][3] = gs[3 * RH];
          const float4* cc = p.cs + (base * RH + j) * 2;
          ct[ch][0] = cc[0]; ct[ch][1] = cc[1];
          if (t > 0) {
            const float4* pc = p.cs + (((size_t)(t - 1) * p.B8 + bc) * RH + j) * 2;
            cp[ch][0] = pc[0]; cp[ch][1] = pc[1];
          } else {
            cp[ch][0] = cp[ch][1] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
        if (p.return_sequences || t == T - 1) {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int bg = b0 + col0 + e;
            float d = 0.f;
            if (bg < B) d = p.return_sequences ? p.dout[((size_t)bg * T + t) * RH + j] : p.dout[(size_t)bg * RH + j];
            dext[ch][e] = d;
          }
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) dext[ch][e] = 0.f;
        }
      }
      if (have_dh) {
        mbar_wait(&bar_mma, phase);
        phase ^= 1;
        tc_fence_after();
      }
      float zi[CH8][8], zf[CH8][8], zg[CH8][8], zo[CH8][8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        uint32_t vd[8];
        if (have_dh) { tmem_ld8(t_lane + col0, vd); tmem_ld_wait(); }
        if (live[ch]) {
          float gi[8], gf[8], gg[8], go[8];
          unpack8(g4[ch][0], gi); unpack8(g4[ch][1], gf); unpack8(g4[ch][2], gg); unpack8(g4[ch][3], go);
          const float* ctv = reinterpret_cast<const float*>(&ct[ch][0]);
          const float* cpv = reinterpret_cast<const float*>(&cp[ch][0]);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float dh = dext[ch][e] + (have_dh ? __uint_as_float(vd[e]) : 0.f);
            const float tc = fast_tanh(ctv[e]);
            const float dcv = fmaf(dh * go[e], 1.f - tc * tc, dc[ch * 8 + e]);
            zo[ch][e] = dh * tc * hs_grad(go[e]);
            zi[ch][e] = dcv * gg[e] * hs_grad(gi[e]);
            zf[ch][e] = dcv * cpv[e] * hs_grad(gf[e]);
            zg[ch][e] = dcv * gi[e] * (1.f - gg[e] * gg[e]);
            dc[ch * 8 + e] = dcv * gf[e];
          }
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) zi[ch][e] = zf[ch][e] = zg[ch][e] = zo[ch][e] = 0.f;
        }
        if (t > 0) {
          const uint32_t off = (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16;
          constexpr uint32_t gstride = 16 * (NB * 16);
          *reinterpret_cast<uint4*>(sZ + 0 * gstride + off) = pack8(zi[ch]);
          *reinterpret_cast<uint4*>(sZ + 1 * gstride + off) = pack8(zf[ch]);
          *reinterpret_cast<uint4*>(sZ + 2 * gstride + off) = pack8(zg[ch]);
          *reinterpret_cast<uint4*>(sZ + 3 * gstride + off) = pack8(zo[ch]);
        }
      }
      if (t > 0) {
        fence_proxy_async_smem();
        tc_fence_before();
        named_bar_arrive(BAR_H, R_ALL);
      }
      // ---- off the critical path: global dz image, input-kernel / bias gradient partials ----
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        if (!live[ch]) continue;
        const int col0 = (cgp * CH8 + ch) * 8;
        uint4* dzp = p.dz + ((size_t)t * p.B8 + ((b0 + col0) >> 3)) * GH + j;
        dzp[0 * RH] = pack8(zi[ch]); dzp[1 * RH] = pack8(zf[ch]); dzp[2 * RH] = pack8(zg[ch]); dzp[3 * RH] = pack8(zo[ch]);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xv = sX[(col0 + e) * T + t];
          acc_w[0] = fmaf(xv, zi[ch][e], acc_w[0]); acc_w[1] = fmaf(xv, zf[ch][e], acc_w[1]);
          acc_w[2] = fmaf(xv, zg[ch][e], acc_w[2]); acc_w[3] = fmaf(xv, zo[ch][e], acc_w[3]);
          acc_b[0] += zi[ch][e]; acc_b[1] += zf[ch][e]; acc_b[2] += zg[ch][e]; acc_b[3] += zo[ch][e];
        }
      }
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      atomicAdd(p.d_wih + g * RH + j, acc_w[g]);
      atomicAdd(p.d_bias + g * RH + j, acc_b[g]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == RCfg<NB>::MMA_WARP) tmem_dealloc(tmem, TM_COLS);
}

// ======================================================================================================
// GRU(128), reset_after = False.  Per step two dependent tensor-core phases:
//   batch A: [z | r] = Uzrᵀ·h + x-terms, plus the x-term of the candidate;   batch B: candidate += Uhᵀ·(r ⊙ h)
// ======================================================================================================
struct GruV2Fwd {
  const float* x;
  const __nv_bfloat16* img;       // [18][48][8][8]
  float* out;
  uint4* hs;                      // [T][B8][128] x 16 B   h_t (bf16 image)
  uint4* rh;                      // [T][B8][128] x 16 B   r_t ⊙ h_{t-1} (bf16 image)
  uint4* zrh;                     // [T][B8][3][128] x 16 B  z, r, hh (bf16)
  float4* hprev;                  // [T][B8][128] x 32 B   h_{t-1} (fp32)
  int B, T, B8;
  int return_sequences, save;
};

template <int NB>
__global__ void __launch_bounds__(RCfg<NB>::ALL, 1) gru_fwd_v2_kernel(GruV2Fwd p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int GH = 3 * RH, MG = GH / 8;                    // 384, 48
  constexpr uint32_t W_LBO = MG * 128;                       // 6144
  constexpr int W_BYTES = R_KG * MG * 128;                   // 110,592
  constexpr int CH8 = RCfg<NB>::CH8, R_ALL = RCfg<NB>::ALL, R_EPI = RCfg<NB>::EPI, NCG = RCfg<NB>::NCG;
  uint8_t* sW = smem;
  uint8_t* sH = sW + W_BYTES;
  uint8_t* sR = sH + RH * NB * 2;
  uint8_t* sX = sR + RH * NB * 2;
  uint8_t* sZero = sX + (size_t)p.T * NB * 16;
  __shared__ uint64_t bar_w, bar_a, bar_b;
  __shared__ uint32_t tmem_s;
  // TMEM: [0, 3*NB) accumulators z | r | candidate, [A_COL0, A_COL0 + 192) the three gate tiles of Uᵀ (resident A)
  constexpr uint32_t A_COL0 = 3 * NB, TM_COLS = 512;
  constexpr uint32_t BAR_W = 3;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  if (warp == RCfg<NB>::MMA_WARP) tmem_alloc(&tmem_s, TM_COLS);
  if (tid == 0) { mbar_init(&bar_w, 1); mbar_init(&bar_a, 1); mbar_init(&bar_b, 1); mbar_fence_init(); }
  __syncthreads();
  pdl_wait();                                               // set-up above overlapped the previous kernel\'s drain
  if (tid == R_EPI) {
    mbar_expect_tx(&bar_w, W_BYTES);
#pragma unroll
    for (int c = 0; c < R_KG; ++c)
      tma_bulk_g2s(sW + c * (MG * 128), reinterpret_cast<const uint8_t*>(p.img) + (size_t)c * (MG * 128), MG * 128, &bar_w);
  }
  build_x_image<NB>(sX, sZero, p.x, b0, B, T);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;

  if (warp == RCfg<NB>::MMA_WARP) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, NB, true, true);
    const uint32_t w_base = smem_u32(sW), h_base = smem_u32(sH), r_base = smem_u32(sR), x_base = smem_u32(sX),
                   z_base = smem_u32(sZero);
    constexpr uint32_t hi = umma_desc_hi(128);
    const uint32_t a_lo0 = umma_desc_lo(w_base, W_LBO), h_lo0 = umma_desc_lo(h_base, NB * 16),
                   r_lo0 = umma_desc_lo(r_base, NB * 16);
    constexpr uint32_t idesc_ts = umma_idesc_bf16(128, NB, false, true);   // A from TMEM (K-major by construction)
    const bool leader = elect_one();
    mbar_wait(&bar_w, 0);
    named_bar_sync(BAR_W, R_ALL);                            // the epilogue warps have copied Uᵀ into TMEM
    for (int t = 0; t < T; ++t) {
      if (t > 0) named_bar_sync(BAR_H, R_ALL);
      tc_fence_after();
      if (leader) {
        const uint32_t xt = x_base + t * (NB * 16);
        const uint32_t x_lo = umma_desc_lo(xt, z_base - xt);
#pragma unroll
        for (int g = 0; g < 3; ++g) {
          if (t > 0 && g < 2) {
#pragma unroll
            for (int ks = 0; ks < RH / 16; ++ks)
              umma_bf16_ts(tmem + g * NB, tmem + A_COL0 + g * 64 + ks * 8, h_lo0 + ((ks * 2 * (NB * 16)) >> 4), hi,
                           idesc_ts, ks > 0);
          }
          umma_bf16_lohi(tmem + g * NB, a_lo0 + ((g * 16 * 128 + 16 * W_LBO) >> 4), hi, x_lo, hi, idesc, t > 0 && g < 2);
        }
        umma_commit(&bar_a);
      }
      __syncwarp();
      if (t > 0) {
        named_bar_sync(BAR_R, R_ALL);                        // r ⊙ h is in shared memory
        tc_fence_after();
        if (leader) {
#pragma unroll
          for (int ks = 0; ks < RH / 16; ++ks)
            umma_bf16_ts(tmem + 2 * NB, tmem + A_COL0 + 2 * 64 + ks * 8, r_lo0 + ((ks * 2 * (NB * 16)) >> 4), hi,
                         idesc_ts, true);
          umma_commit(&bar_b);
        }
        __syncwarp();
      }
    }
  } else {
    const int q = warp & 3, cgp = warp >> 2;
    const int j = q * 32 + lane;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
    mbar_wait(&bar_w, 0);
    for (int g = cgp; g < 3; g += NCG) image_col_to_tmem<MG>(sW, g * RH + j, t_lane + A_COL0 + g * 64);
    tc_fence_before();
    named_bar_arrive(BAR_W, R_ALL);
    float hreg[CH8 * 8];
#pragma unroll
    for (int i = 0; i < CH8 * 8; ++i) hreg[i] = 0.f;
    uint32_t pha = 0, phb = 0;
    for (int t = 0; t < T; ++t) {
      const bool last = (t == T - 1);
      mbar_wait(&bar_a, pha);
      pha ^= 1;
      tc_fence_after();
      float z[CH8][8], r[CH8][8], rhv[CH8][8], hp[CH8][8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        uint32_t vz[8], vr[8];
        tmem_ld8(t_lane + 0 * NB + col0, vz);
        tmem_ld8(t_lane + 1 * NB + col0, vr);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          z[ch][e] = __saturatef(fmaf(0.2f, __uint_as_float(vz[e]), 0.5f));
          r[ch][e] = __saturatef(fmaf(0.2f, __uint_as_float(vr[e]), 0.5f));
          hp[ch][e] = hreg[ch * 8 + e];
          rhv[ch][e] = r[ch][e] * hp[ch][e];
        }
        if (t > 0) *reinterpret_cast<uint4*>(sR + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) = pack8(rhv[ch]);
      }
      if (t > 0) {
        fence_proxy_async_smem();
        tc_fence_before();
        named_bar_arrive(BAR_R, R_ALL);
      }
      if (p.save) {                                          // overlaps the candidate MMAs
#pragma unroll
        for (int ch = 0; ch < CH8; ++ch) {
          const int col0 = (cgp * CH8 + ch) * 8;
          const int bc = (b0 + col0) >> 3;
          if (bc < p.B8) {
            const size_t base = (size_t)t * p.B8 + bc;
            p.rh[base * RH + j] = pack8(rhv[ch]);
            uint4* zd = p.zrh + base * 3 * RH + j;
            zd[0] = pack8(z[ch]); zd[RH] = pack8(r[ch]);
            float4* hd = p.hprev + (base * RH + j) * 2;
            hd[0] = make_float4(hp[ch][0], hp[ch][1], hp[ch][2], hp[ch][3]);
            hd[1] = make_float4(hp[ch][4], hp[ch][5], hp[ch][6], hp[ch][7]);
          }
        }
      }
      if (t > 0) {
        mbar_wait(&bar_b, phb);
        phb ^= 1;
        tc_fence_after();
      }
      float hh[CH8][8], hn[CH8][8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        uint32_t vh[8];
        tmem_ld8(t_lane + 2 * NB + col0, vh);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          hh[ch][e] = fast_tanh(__uint_as_float(vh[e]));
          hn[ch][e] = fmaf(z[ch][e], hp[ch][e] - hh[ch][e], hh[ch][e]);     // z*h + (1-z)*hh
          hreg[ch * 8 + e] = hn[ch][e];
        }
        if (!last) *reinterpret_cast<uint4*>(sH + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) = pack8(hn[ch]);
      }
      if (!last) {
        fence_proxy_async_smem();
        tc_fence_before();
        named_bar_arrive(BAR_H, R_ALL);
      }
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        const int bc = (b0 + col0) >> 3;
        if (bc < p.B8) {
          if (p.save) {
            const size_t base = (size_t)t * p.B8 + bc;
            p.hs[base * RH + j] = pack8(hn[ch]);
            p.zrh[base * 3 * RH + 2 * RH + j] = pack8(hh[ch]);
          }
          if (p.return_sequences || last) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int bg = b0 + col0 + e;
              if (bg < B) {
                if (p.return_sequences) p.out[((size_t)bg * T + t) * RH + j] = hn[ch][e];
                else p.out[(size_t)bg * RH + j] = hn[ch][e];
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == RCfg<NB>::MMA_WARP) tmem_dealloc(tmem, TM_COLS);
}

struct GruV2Bwd {
  const float* dout;
  const float* x;
  const __nv_bfloat16* img;
  const uint4* zrh;               // [T][B8][3][128]
  const float4* hprev;            // [T][B8][128] x 2
  uint4* da;                      // [T][B8][384] x 16 B: d(az), d(ar), d(ah)  == K-major UMMA image
  float* d_wih;                   // [3H]
  float* d_bias;                  // [3H]
  int B, T, B8;
  int return_sequences;
};

template <int NB>
__global__ void __launch_bounds__(RCfg<NB>::ALL, 1) gru_bwd_v2_kernel(GruV2Bwd p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int GH = 3 * RH, MG = GH / 8;
  constexpr int W_BYTES = 16 * MG * 128;                     // 98,304
  constexpr int CH8 = RCfg<NB>::CH8, R_ALL = RCfg<NB>::ALL, R_EPI = RCfg<NB>::EPI, NCG = RCfg<NB>::NCG;
  uint8_t* sW = smem;                                        // image as K-major A (M = h_{t-1} unit, K = gate column)
  uint8_t* sB1 = sW + W_BYTES;                               // [d(az) | d(ah)]ᵀ : k-groups 0..15 = az, 16..31 = ah
  uint8_t* sB2 = sB1 + 2 * RH * NB * 2;                      // d(ar)ᵀ
  float* sX = reinterpret_cast<float*>(sB2 + RH * NB * 2);
  __shared__ uint64_t bar_w, bar_1, bar_2;
  __shared__ uint32_t tmem_s;
  // TMEM: [0, 2*NB) accumulators D1 | D2, [A_COL0, A_COL0 + 192) U (128 x 384 bf16) as the resident A operand
  constexpr uint32_t A_COL0 = 128, TM_COLS = 512;
  constexpr uint32_t BAR_W = 3;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  if (warp == RCfg<NB>::MMA_WARP) tmem_alloc(&tmem_s, TM_COLS);
  if (tid == 0) { mbar_init(&bar_w, 1); mbar_init(&bar_1, 1); mbar_init(&bar_2, 1); mbar_fence_init(); }
  __syncthreads();
  pdl_wait();                                               // set-up above overlapped the previous kernel\'s drain
  if (tid == R_EPI) {
    mbar_expect_tx(&bar_w, W_BYTES);
#pragma unroll
    for (int c = 0; c < 16; ++c)
      tma_bulk_g2s(sW + c * (MG * 128), reinterpret_cast<const uint8_t*>(p.img) + (size_t)c * (MG * 128), MG * 128, &bar_w);
  }
  for (int idx = tid; idx < NB * T; idx += R_ALL) {
    const int b = idx / T, t = idx - b * T;
    sX[idx] = (b0 + b < B) ? p.x[(size_t)(b0 + b) * T + t] : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;

  if (warp == RCfg<NB>::MMA_WARP) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, NB, false, true);
    const uint32_t w_base = smem_u32(sW), b1_base = smem_u32(sB1), b2_base = smem_u32(sB2);
    constexpr uint32_t b_hi = umma_desc_hi(128);
    const uint32_t b1_lo0 = umma_desc_lo(b1_base, NB * 16), b2_lo0 = umma_desc_lo(b2_base, NB * 16);
    (void)w_base;
    constexpr uint32_t B_STEP = (2 * NB * 16) >> 4;          // per K = 16 step, in 16-byte units
    const bool leader = elect_one();
    named_bar_sync(BAR_W, R_ALL);                            // the epilogue warps have copied U into TMEM
    for (int t = T - 1; t >= 1; --t) {
      named_bar_sync(BAR_H, R_ALL);                          // d(az), d(ah) staged
      tc_fence_after();
      if (leader) {
        // D1 (cols 0..NB)   = U_h · d(ah)ᵀ            (image columns 256..383 -> K chunk pairs 16..23)
        // D2 (cols NB..2NB) = U_z · d(az)ᵀ            (image columns 0..127   -> K chunk pairs 0..7)
#pragma unroll
        for (int ks = 0; ks < RH / 16; ++ks)
          umma_bf16_ts(tmem, tmem + A_COL0 + 128 + ks * 8, b1_lo0 + (8 + ks) * B_STEP, b_hi, idesc, ks > 0);
#pragma unroll
        for (int ks = 0; ks < RH / 16; ++ks)
          umma_bf16_ts(tmem + NB, tmem + A_COL0 + ks * 8, b1_lo0 + ks * B_STEP, b_hi, idesc, ks > 0);
        umma_commit(&bar_1);
      }
      __syncwarp();
      named_bar_sync(BAR_R, R_ALL);                          // d(ar) staged
      tc_fence_after();
      if (leader) {
        // D2 += U_r · d(ar)ᵀ                             (image columns 128..255 -> K chunk pairs 8..15)
#pragma unroll
        for (int ks = 0; ks < RH / 16; ++ks)
          umma_bf16_ts(tmem + NB, tmem + A_COL0 + 64 + ks * 8, b2_lo0 + ks * B_STEP, b_hi, idesc, true);
        umma_commit(&bar_2);
      }
      __syncwarp();
    }
  } else {
    const int q = warp & 3, cgp = warp >> 2;
    const int j = q * 32 + lane;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
    mbar_wait(&bar_w, 0);
    for (int g = cgp; g < 3; g += NCG) image_row_to_tmem<MG>(sW, j, g * 16, 2, t_lane + A_COL0 + g * 64);
    tc_fence_before();
    named_bar_arrive(BAR_W, R_ALL);
    float dh[CH8 * 8];
#pragma unroll
    for (int i = 0; i < CH8 * 8; ++i) dh[i] = 0.f;
    float acc_w[3] = {0.f, 0.f, 0.f}, acc_b[3] = {0.f, 0.f, 0.f};
    uint32_t ph1 = 0, ph2 = 0;
    for (int t = T - 1; t >= 0; --t) {
      uint4 g3[CH8][3];
      float4 hpv[CH8][2];
      float dext[CH8][8];
      bool live[CH8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        const int bc = (b0 + col0) >> 3;
        live[ch] = bc < p.B8;
        if (live[ch]) {
          const size_t base = (size_t)t * p.B8 + bc;
          const uint4* gs = p.zrh + base * 3 * RH + j;
          g3[ch][0] = gs[0]; g3[ch][1] = gs[RH]; g3[ch][2] = gs[2 * RH];
          const float4* hh = p.hprev + (base * RH + j) * 2;
          hpv[ch][0] = hh[0]; hpv[ch][1] = hh[1];
        }
        if (p.return_sequences || t == T - 1) {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int bg = b0 + col0 + e;
            float d = 0.f;
            if (bg < B) d = p.return_sequences ? p.dout[((size_t)bg * T + t) * RH + j] : p.dout[(size_t)bg * RH + j];
            dext[ch][e] = d;
          }
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) dext[ch][e] = 0.f;
        }
      }
      // ---- stage 1: d(ah), d(az) ----
      float daz[CH8][8], dah[CH8][8], pdh[CH8][8], zv[CH8][8], rv[CH8][8], hp[CH8][8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        if (live[ch]) {
          float hh[8];
          unpack8(g3[ch][0], zv[ch]); unpack8(g3[ch][1], rv[ch]); unpack8(g3[ch][2], hh);
          const float* hpf = reinterpret_cast<const float*>(&hpv[ch][0]);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            hp[ch][e] = hpf[e];
            const float d = dh[ch * 8 + e] + dext[ch][e];
            dah[ch][e] = d * (1.f - zv[ch][e]) * (1.f - hh[e] * hh[e]);
            daz[ch][e] = d * (hp[ch][e] - hh[e]) * hs_grad(zv[ch][e]);
            pdh[ch][e] = d * zv[ch][e];
          }
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) { dah[ch][e] = daz[ch][e] = pdh[ch][e] = zv[ch][e] = rv[ch][e] = hp[ch][e] = 0.f; }
        }
        if (t > 0) {
          const uint32_t off = (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16;
          *reinterpret_cast<uint4*>(sB1 + off) = pack8(daz[ch]);
          *reinterpret_cast<uint4*>(sB1 + 16 * (NB * 16) + off) = pack8(dah[ch]);
        }
      }
      if (t > 0) {
        fence_proxy_async_smem();
        tc_fence_before();
        named_bar_arrive(BAR_H, R_ALL);
      }
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {                     // off the critical path
        if (!live[ch]) continue;
        const int col0 = (cgp * CH8 + ch) * 8;
        uint4* dap = p.da + ((size_t)t * p.B8 + ((b0 + col0) >> 3)) * GH + j;
        dap[0] = pack8(daz[ch]); dap[2 * RH] = pack8(dah[ch]);
        if (t == 0) { const float zero8[8] = {0, 0, 0, 0, 0, 0, 0, 0}; dap[RH] = pack8(zero8); }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xv = sX[(col0 + e) * T + t];
          acc_w[0] = fmaf(xv, daz[ch][e], acc_w[0]); acc_b[0] += daz[ch][e];
          acc_w[2] = fmaf(xv, dah[ch][e], acc_w[2]); acc_b[2] += dah[ch][e];
        }
      }
      if (t == 0) break;
      // ---- stage 2: drh -> d(ar) ----
      mbar_wait(&bar_1, ph1);
      ph1 ^= 1;
      tc_fence_after();
      float dar[CH8][8];
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        uint32_t vd[8];
        tmem_ld8(t_lane + col0, vd);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float drh = live[ch] ? __uint_as_float(vd[e]) : 0.f;
          dar[ch][e] = drh * hp[ch][e] * hs_grad(rv[ch][e]);
          pdh[ch][e] = fmaf(drh, rv[ch][e], pdh[ch][e]);
        }
        *reinterpret_cast<uint4*>(sB2 + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) = pack8(dar[ch]);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      named_bar_arrive(BAR_R, R_ALL);
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        if (!live[ch]) continue;
        const int col0 = (cgp * CH8 + ch) * 8;
        p.da[((size_t)t * p.B8 + ((b0 + col0) >> 3)) * GH + RH + j] = pack8(dar[ch]);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xv = sX[(col0 + e) * T + t];
          acc_w[1] = fmaf(xv, dar[ch][e], acc_w[1]); acc_b[1] += dar[ch][e];
        }
      }
      mbar_wait(&bar_2, ph2);
      ph2 ^= 1;
      tc_fence_after();
#pragma unroll
      for (int ch = 0; ch < CH8; ++ch) {
        const int col0 = (cgp * CH8 + ch) * 8;
        uint32_t vd[8];
        tmem_ld8(t_lane + NB + col0, vd);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 8; ++e) dh[ch * 8 + e] = pdh[ch][e] + (live[ch] ? __uint_as_float(vd[e]) : 0.f);
      }
      // the next iteration's stage-1 stores and MMAs are ordered behind these TMEM reads by the BAR_H arrive
    }
#pragma unroll
    for (int g = 0; g < 3; ++g) {
      atomicAdd(p.d_wih + g * RH + j, acc_w[g]);
      atomicAdd(p.d_bias + g * RH + j, acc_b[g]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == RCfg<NB>::MMA_WARP) tmem_dealloc(tmem, TM_COLS);
}

// ------------------------------------------------------------------------------------------------------
// Recurrent weight-gradient GEMM on pre-tiled operands:
//   C[128, n0:n0+BN] = sum over k-groups g of  A_img[g]ᵀ(8 x 128) · B_img[g](8 x N)
// A_img: [KG][128] x 16 B, B_img: [KG][b_ntotal] x 16 B — both already K-major UMMA images in global memory,
// so a producer thread streams them with bulk TMA copies through a 4-stage full/empty mbarrier ring, one
// thread issues the MMAs, four warps drain TMEM.  KG must be a multiple of 8 (callers zero-pad).
// ------------------------------------------------------------------------------------------------------
constexpr int WG2_STAGES = 4;
constexpr int WG2_THREADS = 192;   // warp 0: TMA producer, warp 1: MMA issuer, warps 2-5: epilogue

template <int BN>
__global__ void __launch_bounds__(WG2_THREADS, 1) rnn128_wgrad_kernel(const uint8_t* __restrict__ a_img,
                                                                      const uint8_t* __restrict__ b_img, float* __restrict__ C,
                                                                      int kgroups, int b_ntotal, int b_noff, int ldc) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int A_ST = 8 * RH * 16, B_ST = 8 * BN * 16, ST = A_ST + B_ST;
  __shared__ uint64_t full[WG2_STAGES], empty[WG2_STAGES], done;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * BN;
  constexpr uint32_t TM_COLS = BN < 32 ? 32 : BN;
  if (warp == 0) tmem_alloc(&tmem_s, TM_COLS);
  if (tid == 32) {
    for (int s = 0; s < WG2_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&done, 1);
    mbar_fence_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem = tmem_s;
  const int nkb = kgroups / 8;
  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % WG2_STAGES;
        if (kb >= WG2_STAGES) mbar_wait(&empty[s], ((kb / WG2_STAGES) - 1) & 1);
        uint8_t* sa = smem + s * ST;
        uint8_t* sb = sa + A_ST;
        mbar_expect_tx(&full[s], ST);
        tma_bulk_g2s(sa, a_img + (size_t)kb * A_ST, A_ST, &full[s]);       // 8 k-groups of A are contiguous
#pragma unroll
        for (int g = 0; g < 8; ++g)
          tma_bulk_g2s(sb + g * (BN * 16), b_img + ((size_t)(kb * 8 + g) * b_ntotal + b_noff + n0) * 16, BN * 16, &full[s]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, BN, false, false);
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % WG2_STAGES;
        mbar_wait(&full[s], (kb / WG2_STAGES) & 1);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * ST), b_base = a_base + A_ST;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          umma_bf16(tmem, umma_desc(a_base + ks * 2 * (RH * 16), RH * 16, 128),
                    umma_desc(b_base + ks * 2 * (BN * 16), BN * 16, 128), idesc, (kb | ks) != 0);
        umma_commit(&empty[s]);
      }
      umma_commit(&done);
    }
  } else {
    mbar_wait(&done, 0);
    tc_fence_after();
    const int q = warp & 3;                                   // warps 2..5 -> lane quarters 2,3,0,1
    const int row = q * 32 + lane;
    float* dst = C + (size_t)row * ldc + n0;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (nkb > 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          reinterpret_cast<float4*>(dst + c0)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                               __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(dst + c0)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, TM_COLS);
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_rnn128_pack(const float* recurrent, const float* w_ih, const float* bias, void* image, int gates,
                                void* stream) {
  if (gates != 3 && gates != 4) { set_last_error("ddls_rnn128_pack: gates must be 3 (GRU) or 4 (LSTM)"); return 1; }
  const int GH = gates * RH;
  const int total = R_KG * (GH / 8) * 8;
  launch_k(rnn128_pack_kernel, dim3((total + 255) / 256), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      recurrent, w_ih, bias, static_cast<__nv_bfloat16*>(image), GH);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "rnn128_pack_kernel launch");
}

static int pick_tile(int B, int tile) {
  if (tile > 0) return tile;
  if (B <= 16 * 148) return 16;          // small batches: spread over more SMs (B = 32 -> two 16-row CTAs)
  return (B <= 32 * 148) ? 32 : 64;
}

extern "C" int ddls_lstm128_forward_v2(const float* x, const void* image, float* out, void* hs, void* gates, void* cs, int B,
                                       int T, int return_sequences, int save_for_backward, int tile, void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_lstm128_forward_v2: empty problem"); return 1; }
  LstmV2Fwd p{x, static_cast<const __nv_bfloat16*>(image), out, static_cast<uint4*>(hs), static_cast<uint4*>(gates),
              static_cast<float4*>(cs), B, T, (B + 7) / 8, return_sequences, save_for_backward};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = pick_tile(B, tile);
  const size_t smem = (size_t)R_KG * 64 * 128 + (size_t)RH * nb * 2 + (size_t)T * nb * 16 + (size_t)nb * 16;
  if (smem > 227 * 1024) { set_last_error("ddls_lstm128_forward_v2: T=%d too long for a %d-row tile", T, nb); return 1; }
  if (nb == 16) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_fwd_v2_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(lstm_fwd_v2_kernel<16>, dim3((B + 15) / 16), dim3(RCfg<16>::ALL), smem, st, p);
  } else if (nb == 32) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_fwd_v2_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(lstm_fwd_v2_kernel<32>, dim3((B + 31) / 32), dim3(RCfg<32>::ALL), smem, st, p);
  } else if (nb == 64) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_fwd_v2_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(lstm_fwd_v2_kernel<64>, dim3((B + 63) / 64), dim3(RCfg<64>::ALL), smem, st, p);
  } else { set_last_error("ddls_lstm128_forward_v2: tile must be 16, 32 or 64"); return 1; }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_fwd_v2_kernel launch");
}

extern "C" int ddls_lstm128_backward_v2(const float* dout, const float* x, const void* image, const void* gates,
                                        const void* cs, void* dz, float* d_wih, float* d_bias, int B, int T,
                                        int return_sequences, int tile, void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_lstm128_backward_v2: empty problem"); return 1; }
  LstmV2Bwd p{dout, x, static_cast<const __nv_bfloat16*>(image), static_cast<const uint4*>(gates),
              static_cast<const float4*>(cs), static_cast<uint4*>(dz), d_wih, d_bias, B, T, (B + 7) / 8, return_sequences};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = pick_tile(B, tile);
  const size_t smem = (size_t)16 * 64 * 128 + (size_t)4 * RH * nb * 2 + (size_t)nb * T * 4;
  if (smem > 227 * 1024) { set_last_error("ddls_lstm128_backward_v2: T=%d too long for a %d-row tile", T, nb); return 1; }
  if (nb == 16) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_bwd_v2_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(lstm_bwd_v2_kernel<16>, dim3((B + 15) / 16), dim3(RCfg<16>::ALL), smem, st, p);
  } else if (nb == 32) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_bwd_v2_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(lstm_bwd_v2_kernel<32>, dim3((B + 31) / 32), dim3(RCfg<32>::ALL), smem, st, p);
  } else if (nb == 64) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_bwd_v2_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(lstm_bwd_v2_kernel<64>, dim3((B + 63) / 64), dim3(RCfg<64>::ALL), smem, st, p);
  } else { set_last_error("ddls_lstm128_backward_v2: tile must be 16, 32 or 64"); return 1; }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_bwd_v2_kernel launch");
}

extern "C" int ddls_rnn128_wgrad(const void* a_img, const void* b_img, float* C, int kgroups, int n, int b_ntotal,
                                 int b_noff, int ldc, void* stream) {
  if (kgroups % 8 != 0 || n % 64 != 0) {
    set_last_error("ddls_rnn128_wgrad: kgroups must be a multiple of 8 and n a multiple of 64 (got %d, %d)", kgroups, n);
    return 1;
  }
  constexpr int BN = 64;
  const int smem = WG2_STAGES * (8 * RH * 16 + 8 * BN * 16);
  DDLS_CUDA_OK(cudaFuncSetAttribute(rnn128_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  launch_k(rnn128_wgrad_kernel<BN>, dim3(n / BN), dim3(WG2_THREADS), smem, static_cast<cudaStream_t>(stream), 
      static_cast<const uint8_t*>(a_img), static_cast<const uint8_t*>(b_img), C, kgroups, b_ntotal, b_noff, ldc);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "rnn128_wgrad_kernel launch");
}

extern "C" int ddls_gru128_forward_v2(const float* x, const void* image, float* out, void* hs, void* rh, void* zrh,
                                      void* hprev, int B, int T, int return_sequences, int save_for_backward, int tile,
                                      void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_gru128_forward_v2: empty problem"); return 1; }
  GruV2Fwd p{x, static_cast<const __nv_bfloat16*>(image), out, static_cast<uint4*>(hs), static_cast<uint4*>(rh),
             static_cast<uint4*>(zrh), static_cast<float4*>(hprev), B, T, (B + 7) / 8, return_sequences, save_for_backward};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = pick_tile(B, tile);
  const size_t smem = (size_t)R_KG * 48 * 128 + 2 * (size_t)RH * nb * 2 + (size_t)T * nb * 16 + (size_t)nb * 16;
  if (smem > 227 * 1024) { set_last_error("ddls_gru128_forward_v2: T=%d too long for a %d-row tile", T, nb); return 1; }
  if (nb == 16) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(gru_fwd_v2_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(gru_fwd_v2_kernel<16>, dim3((B + 15) / 16), dim3(RCfg<16>::ALL), smem, st, p);
  } else if (nb == 32) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(gru_fwd_v2_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(gru_fwd_v2_kernel<32>, dim3((B + 31) / 32), dim3(RCfg<32>::ALL), smem, st, p);
  } else if (nb == 64) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(gru_fwd_v2_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(gru_fwd_v2_kernel<64>, dim3((B + 63) / 64), dim3(RCfg<64>::ALL), smem, st, p);
  } else { set_last_error("ddls_gru128_forward_v2: tile must be 16, 32 or 64"); return 1; }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "gru_fwd_v2_kernel launch");
}

extern "C" int ddls_gru128_backward_v2(const float* dout, const float* x, const void* image, const void* zrh,
                                       const void* hprev, void* da, float* d_wih, float* d_bias, int B, int T,
                                       int return_sequences, int tile, void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_gru128_backward_v2: empty problem"); return 1; }
  GruV2Bwd p{dout, x, static_cast<const __nv_bfloat16*>(image), static_cast<const uint4*>(zrh),
             static_cast<const float4*>(hprev), static_cast<uint4*>(da), d_wih, d_bias, B, T, (B + 7) / 8, return_sequences};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = pick_tile(B, tile);
  const size_t smem = (size_t)16 * 48 * 128 + 3 * (size_t)RH * nb * 2 + (size_t)nb * T * 4;
  if (smem > 227 * 1024) { set_last_error("ddls_gru128_backward_v2: T=%d too long for a %d-row tile", T, nb); return 1; }
  if (nb == 16) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(gru_bwd_v2_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(gru_bwd_v2_kernel<16>, dim3((B + 15) / 16), dim3(RCfg<16>::ALL), smem, st, p);
  } else if (nb == 32) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(gru_bwd_v2_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(gru_bwd_v2_kernel<32>, dim3((B + 31) / 32), dim3(RCfg<32>::ALL), smem, st, p);
  } else if (nb == 64) {
    DDLS_CUDA_OK(cudaFuncSetAttribute(gru_bwd_v2_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch_k(gru_bwd_v2_kernel<64>, dim3((B + 63) / 64), dim3(RCfg<64>::ALL), smem, st, p);
  } else { set_last_error("ddls_gru128_backward_v2: tile must be 16, 32 or 64"); return 1; }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "gru_bwd_v2_kernel launch");
}
