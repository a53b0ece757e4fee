// Persistent LSTM(128) recurrence on tcgen05 — forward and BPTT (K7/K9 of SURVEY.md §2.5).
//
// Keras-2 LSTM as the reference builds it (`LSTM(128, input_shape=(25,1))`,
// /root/reference/ddl_nyiso_aztk.py:249-251): gate order i,f,c,o, hard_sigmoid gates, tanh cell,
//   z = x_t·W + h_{t-1}·U + b ;  c_t = f⊙c_{t-1} + i⊙tanh(z_c) ;  h_t = o⊙tanh(c_t).
//
// Design ("weights on M, batch on N"): one CTA owns a tile of NB batch rows for ALL T steps; rows are
// independent, so there is no grid-wide sync.  Per step the CTA computes
//        Zᵀ[4H x NB] = Uᵀ[4H x H] · h_{t-1}ᵀ[H x NB]
// as four M=128 UMMA tiles (one per gate), K = 128, N = NB.  Uᵀ is read straight from the Keras (H,4H)
// layout as an MN-major operand and stays resident in shared memory (bf16, 128 KB) for the whole
// sequence; the accumulators live in TMEM (4·NB columns), TMEM lane = hidden unit, column = batch row.
// The epilogue thread that owns unit j reads its four gate pre-activations with tcgen05.ld, adds the
// rank-1 input projection (input dim 1) and bias, applies the activations, keeps c in registers,
// and writes h_t (bf16) back into shared memory as the next step's N-side operand — eight batch
// rows per 16-byte store, conflict-free in the SWIZZLE_NONE MN-major layout.  Small batches therefore
// cost N = NB columns of tensor time instead of a padded M = 128 tile.
//
// BPTT runs the same way in reverse: dh_{t-1}ᵀ[H x NB] = U[H x 4H] · dzᵀ[4H x NB] (U K-major, resident),
// gate derivatives fused in the epilogue, dz stashed (bf16) for one weight-gradient GEMM over all
// (t, b), input-kernel and bias gradients accumulated in registers and reduced with atomics.
#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int LH = 128;          // hidden units (NYISO models)
constexpr int LG = 4 * LH;       // gate columns
constexpr int L_THREADS = 256;   // 8 warps: 4 TMEM lane quarters x 2 column halves

struct LstmFwdParams {
  const float* x;        // [B, T] (input dim 1)
  const float* w_ih;     // [4H]   kernel (1, 4H)
  const float* w_hh;     // [H, 4H] recurrent kernel
  const float* bias;     // [4H]
  float* out;            // [B, H] or [B, T, H]
  __nv_bfloat16* hs;     // [T, B, H]  h_t (bf16) for the weight-gradient GEMM
  __nv_bfloat16* gates;  // [T, B, H, 4] post-activation i,f,g,o
  float* cs;             // [T, B, H]  c_t
  int B, T;
  int return_sequences;
  int save_for_backward;
};

template <int NB>
__global__ void __launch_bounds__(L_THREADS, 1) lstm_fwd_kernel(LstmFwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sW = smem;                        // Uᵀ as MN-major A: [K/8=16][M/8=64][8 k][8 m] bf16 = 128 KB
  uint8_t* sH = smem + LH * LG * 2;          // h as MN-major B:  [K/8=16][NB/8][8 k][8 n] bf16
  float* sX = reinterpret_cast<float*>(sH + LH * NB * 2);  // [NB][T]
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t TM_COLS = 4 * NB;       // 128 / 256 / 512
  constexpr int CPT = NB / 2;                // batch columns per thread

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hc = warp >> 2;
  const int j = q * 32 + lane;               // hidden unit owned by this thread (= TMEM lane)
  const int b_tile0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  if (warp == 0) tmem_alloc(&tmem_base_s, TM_COLS);
  if (tid == 32) { mbar_init(&mma_bar, 1); mbar_fence_init(); }

  // ---- stage Uᵀ: fp32 [k][m] -> bf16 MN-major (quarter-warps store 128 contiguous bytes) ----------
  for (int idx = tid; idx < LH * (LG / 8); idx += L_THREADS) {
    const int blk = idx >> 5, ln = idx & 31;
    const int kg = blk & 15, mg4 = blk >> 4;            // 16 k-groups x 16 blocks of 4 m-groups
    const int k = kg * 8 + (ln & 7), mg = mg4 * 4 + (ln >> 3);
    const float4* src = reinterpret_cast<const float4*>(p.w_hh + (size_t)k * LG + mg * 8);
    const float4 a = src[0], b = src[1];
    uint4 v = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    *reinterpret_cast<uint4*>(sW + kg * 8192 + mg * 128 + (k & 7) * 16) = v;
  }
  for (int idx = tid; idx < NB * T; idx += L_THREADS) {
    const int b = idx / T, t = idx - b * T;
    sX[idx] = (b_tile0 + b < B) ? p.x[(size_t)(b_tile0 + b) * T + t] : 0.f;
  }
  float wi[4], bi[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) { wi[g] = p.w_ih[g * LH + j]; bi[g] = p.bias[g * LH + j]; }

  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
  constexpr uint32_t idesc = umma_idesc_bf16(128, NB, true, true);
  const uint32_t w_base = smem_u32(sW), h_base = smem_u32(sH);

  float c[CPT];
#pragma unroll
  for (int i = 0; i < CPT; ++i) c[i] = 0.f;

  uint32_t phase = 0;
  for (int t = 0; t < T; ++t) {
    if (t > 0) {
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
          for (int ks = 0; ks < LH / 16; ++ks) {
            umma_bf16(tmem + g * NB, umma_desc(w_base + g * 16 * 128 + ks * 2 * 8192, 8192, 128),
                      umma_desc(h_base + ks * 2 * (NB * 16), NB * 16, 128), idesc, ks > 0);
          }
        }
        umma_commit(&mma_bar);
      }
      mbar_wait(&mma_bar, phase);
      phase ^= 1;
      tc_fence_after();
    }
    const bool last = (t == T - 1);
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;    // first batch column of this chunk (within the tile)
      uint32_t vi[8], vf[8], vg[8], vo[8];
      if (t > 0) {
        tmem_ld8(t_lane + 0 * NB + col0, vi);
        tmem_ld8(t_lane + 1 * NB + col0, vf);
        tmem_ld8(t_lane + 2 * NB + col0, vg);
        tmem_ld8(t_lane + 3 * NB + col0, vo);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) vi[e] = vf[e] = vg[e] = vo[e] = 0u;
      }
      float hval[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bl = col0 + e;
        const float xv = sX[bl * T + t];
        const float gi = hard_sigmoid(__uint_as_float(vi[e]) + fmaf(xv, wi[0], bi[0]));
        const float gf = hard_sigmoid(__uint_as_float(vf[e]) + fmaf(xv, wi[1], bi[1]));
        const float gg = fast_tanh(__uint_as_float(vg[e]) + fmaf(xv, wi[2], bi[2]));
        const float go = hard_sigmoid(__uint_as_float(vo[e]) + fmaf(xv, wi[3], bi[3]));
        const float cn = fmaf(gf, c[ch * 8 + e], gi * gg);
        c[ch * 8 + e] = cn;
        const float hv = go * fast_tanh(cn);
        hval[e] = hv;
        const int bglob = b_tile0 + bl;
        if (bglob < B) {
          if (p.save_for_backward) {
            const size_t o = ((size_t)t * B + bglob) * LH + j;
            uint2 pk = make_uint2(pack_bf16x2(gi, gf), pack_bf16x2(gg, go));
            *reinterpret_cast<uint2*>(p.gates + o * 4) = pk;
            p.cs[o] = cn;
            p.hs[o] = __float2bfloat16(hv);
          }
          if (p.return_sequences) p.out[((size_t)bglob * T + t) * LH + j] = hv;
          else if (last) p.out[(size_t)bglob * LH + j] = hv;
        }
      }
      if (!last) {
        uint4 hv4 = make_uint4(pack_bf16x2(hval[0], hval[1]), pack_bf16x2(hval[2], hval[3]),
                               pack_bf16x2(hval[4], hval[5]), pack_bf16x2(hval[6], hval[7]));
        *reinterpret_cast<uint4*>(sH + (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16) = hv4;
      }
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
struct LstmBwdParams {
  const float* dout;           // [B, H] (last step) or [B, T, H]
  const float* x;              // [B, T]
  const float* w_hh;           // [H, 4H]
  const __nv_bfloat16* gates;  // [T, B, H, 4]
  const float* cs;             // [T, B, H]
  __nv_bfloat16* dz;           // [T, B, 4H]  gate pre-activation gradients (bf16) for the wgrad GEMM
  float* d_wih;                // [4H]  (atomic accumulate)
  float* d_bias;               // [4H]  (atomic accumulate)
  int B, T;
  int return_sequences;
};

template <int NB>
__global__ void __launch_bounds__(L_THREADS, 1) lstm_bwd_kernel(LstmBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sW = smem;                    // U as K-major A: [K/8=64][M=128][8] bf16 = 128 KB (m = h_{t-1} unit, k = gate col)
  uint8_t* sZ = smem + LH * LG * 2;      // dz as MN-major B: [K/8=64][NB/8][8 k][8 n] bf16 = NB KB
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t TM_COLS = NB < 32 ? 32 : NB;
  constexpr int CPT = NB / 2;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = warp & 3, hc = warp >> 2;
  const int j = q * 32 + lane;
  const int b_tile0 = blockIdx.x * NB;
  const int T = p.T, B = p.B;

  if (warp == 0) tmem_alloc(&tmem_base_s, TM_COLS);
  if (tid == 32) { mbar_init(&mma_bar, 1); mbar_fence_init(); }

  // ---- stage U: fp32 [m][k] (k contiguous) -> bf16 K-major ------------------------------------------
  for (int idx = tid; idx < LH * (LG / 8); idx += L_THREADS) {
    const int blk = idx >> 5, ln = idx & 31;
    const int mblk = blk & 15, kc4 = blk >> 4;           // 16 blocks of 8 rows x 16 blocks of 4 k-chunks
    const int m = mblk * 8 + (ln & 7), kc = kc4 * 4 + (ln >> 3);
    const float4* src = reinterpret_cast<const float4*>(p.w_hh + (size_t)m * LG + kc * 8);
    const float4 a = src[0], b = src[1];
    uint4 v = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    *reinterpret_cast<uint4*>(sW + kc * (LH * 16) + m * 16) = v;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
  constexpr uint32_t idesc = umma_idesc_bf16(128, NB, false, true);
  const uint32_t w_base = smem_u32(sW), z_base = smem_u32(sZ);

  float dc[CPT];
#pragma unroll
  for (int i = 0; i < CPT; ++i) dc[i] = 0.f;
  float acc_w[4] = {0.f, 0.f, 0.f, 0.f}, acc_b[4] = {0.f, 0.f, 0.f, 0.f};

  uint32_t phase = 0;
  for (int t = T - 1; t >= 0; --t) {
    const bool have_dh = (t < T - 1);        // TMEM holds dh_t from the previous iteration's MMA
    if (have_dh) {
      mbar_wait(&mma_bar, phase);
      phase ^= 1;
      tc_fence_after();
    }
#pragma unroll
    for (int ch = 0; ch < CPT / 8; ++ch) {
      const int col0 = hc * CPT + ch * 8;
      uint32_t vd[8];
      if (have_dh) { tmem_ld8(t_lane + col0, vd); tmem_ld_wait(); }
      float dzi[8], dzf[8], dzg[8], dzo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int bl = col0 + e, bglob = b_tile0 + bl;
        float zi = 0.f, zf = 0.f, zg = 0.f, zo = 0.f;
        if (bglob < B) {
          const size_t o = ((size_t)t * B + bglob) * LH + j;
          const uint2 pk = *reinterpret_cast<const uint2*>(p.gates + o * 4);
          const __nv_bfloat162 p0 = *reinterpret_cast<const __nv_bfloat162*>(&pk.x);
          const __nv_bfloat162 p1 = *reinterpret_cast<const __nv_bfloat162*>(&pk.y);
          const float gi = __low2float(p0), gf = __high2float(p0), gg = __low2float(p1), go = __high2float(p1);
          const float ct = p.cs[o];
          const float cprev = (t > 0) ? p.cs[o - (size_t)B * LH] : 0.f;
          float dh = have_dh ? __uint_as_float(vd[e]) : 0.f;
          if (p.return_sequences) dh += p.dout[((size_t)bglob * T + t) * LH + j];
          else if (t == T - 1) dh += p.dout[(size_t)bglob * LH + j];
          const float tc = fast_tanh(ct);
          const float dcv = fmaf(dh * go, 1.f - tc * tc, dc[ch * 8 + e]);
          zo = dh * tc * ((go > 0.f && go < 1.f) ? 0.2f : 0.f);
          zi = dcv * gg * ((gi > 0.f && gi < 1.f) ? 0.2f : 0.f);
          zf = dcv * cprev * ((gf > 0.f && gf < 1.f) ? 0.2f : 0.f);
          zg = dcv * gi * (1.f - gg * gg);
          dc[ch * 8 + e] = dcv * gf;
          const float xv = p.x[(size_t)bglob * T + t];
          acc_w[0] = fmaf(xv, zi, acc_w[0]); acc_w[1] = fmaf(xv, zf, acc_w[1]);
          acc_w[2] = fmaf(xv, zg, acc_w[2]); acc_w[3] = fmaf(xv, zo, acc_w[3]);
          acc_b[0] += zi; acc_b[1] += zf; acc_b[2] += zg; acc_b[3] += zo;
          __nv_bfloat16* dzrow = p.dz + ((size_t)t * B + bglob) * LG + j;
          dzrow[0 * LH] = __float2bfloat16(zi); dzrow[1 * LH] = __float2bfloat16(zf);
          dzrow[2 * LH] = __float2bfloat16(zg); dzrow[3 * LH] = __float2bfloat16(zo);
        }
        dzi[e] = zi; dzf[e] = zf; dzg[e] = zg; dzo[e] = zo;
      }
      if (t > 0) {
        // K index of gate g, unit j is g*128 + j: k-group = g*16 + j/8, row in group = j%8
        const uint32_t off = (j >> 3) * (NB * 16) + (col0 >> 3) * 128 + (j & 7) * 16;
        const uint32_t gstride = 16 * (NB * 16);
        *reinterpret_cast<uint4*>(sZ + 0 * gstride + off) = make_uint4(pack_bf16x2(dzi[0], dzi[1]), pack_bf16x2(dzi[2], dzi[3]), pack_bf16x2(dzi[4], dzi[5]), pack_bf16x2(dzi[6], dzi[7]));
        *reinterpret_cast<uint4*>(sZ + 1 * gstride + off) = make_uint4(pack_bf16x2(dzf[0], dzf[1]), pack_bf16x2(dzf[2], dzf[3]), pack_bf16x2(dzf[4], dzf[5]), pack_bf16x2(dzf[6], dzf[7]));
        *reinterpret_cast<uint4*>(sZ + 2 * gstride + off) = make_uint4(pack_bf16x2(dzg[0], dzg[1]), pack_bf16x2(dzg[2], dzg[3]), pack_bf16x2(dzg[4], dzg[5]), pack_bf16x2(dzg[6], dzg[7]));
        *reinterpret_cast<uint4*>(sZ + 3 * gstride + off) = make_uint4(pack_bf16x2(dzo[0], dzo[1]), pack_bf16x2(dzo[2], dzo[3]), pack_bf16x2(dzo[4], dzo[5]), pack_bf16x2(dzo[6], dzo[7]));
      }
    }
    if (t > 0) {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
#pragma unroll 8
        for (int ks = 0; ks < LG / 16; ++ks) {
          umma_bf16(tmem, umma_desc(w_base + ks * 2 * (LH * 16), LH * 16, 128),
                    umma_desc(z_base + ks * 2 * (NB * 16), NB * 16, 128), idesc, ks > 0);
        }
        umma_commit(&mma_bar);
      }
    }
  }
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    atomicAdd(p.d_wih + g * LH + j, acc_w[g]);
    atomicAdd(p.d_bias + g * LH + j, acc_b[g]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, TM_COLS);
}

template <int NB>
static int launch_fwd(const LstmFwdParams& p, cudaStream_t st) {
  const size_t smem = (size_t)LH * LG * 2 + (size_t)LH * NB * 2 + (size_t)NB * p.T * 4;
  if (smem > 227 * 1024) { set_last_error("ddls_lstm128_forward: T=%d too long for a %d-row tile", p.T, NB); return 1; }
  auto kern = lstm_fwd_kernel<NB>;
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(p.B + NB - 1) / NB, L_THREADS, smem, st>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_fwd_kernel launch");
}

template <int NB>
static int launch_bwd(const LstmBwdParams& p, cudaStream_t st) {
  const size_t smem = (size_t)LH * LG * 2 + (size_t)LG * NB * 2;
  auto kern = lstm_bwd_kernel<NB>;
  DDLS_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(p.B + NB - 1) / NB, L_THREADS, smem, st>>>(p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_bwd_kernel launch");
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_lstm128_forward(const float* x, const float* w_ih, const float* w_hh, const float* bias, float* out,
                                    void* hs, void* gates, float* cs, int B, int T, int return_sequences,
                                    int save_for_backward, int tile, void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_lstm128_forward: empty problem"); return 1; }
  LstmFwdParams p{x, w_ih, w_hh, bias, out, static_cast<__nv_bfloat16*>(hs), static_cast<__nv_bfloat16*>(gates), cs,
                  B, T, return_sequences, save_for_backward};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tile <= 0) tile = (B <= 32 * 148) ? 32 : (B <= 64 * 148 ? 64 : 128);
  if (tile == 128 && (size_t)128 * T * 4 > 60 * 1024) tile = 64;
  switch (tile) {
    case 32: return launch_fwd<32>(p, st);
    case 64: return launch_fwd<64>(p, st);
    case 128: return launch_fwd<128>(p, st);
  }
  set_last_error("ddls_lstm128_forward: tile must be 32, 64 or 128");
  return 1;
}

extern "C" int ddls_lstm128_backward(const float* dout, const float* x, const float* w_hh, const void* gates,
                                     const float* cs, void* dz, float* d_wih, float* d_bias, int B, int T,
                                     int return_sequences, int tile, void* stream) {
  if (B <= 0 || T <= 0) { set_last_error("ddls_lstm128_backward: empty problem"); return 1; }
  LstmBwdParams p{dout, x, w_hh, static_cast<const __nv_bfloat16*>(gates), cs, static_cast<__nv_bfloat16*>(dz),
                  d_wih, d_bias, B, T, return_sequences};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tile <= 0) tile = (B <= 32 * 148) ? 32 : 64;
  switch (tile) {
    case 32: return launch_bwd<32>(p, st);
    case 64: return launch_bwd<64>(p, st);
  }
  set_last_error("ddls_lstm128_backward: tile must be 32 or 64");
  return 1;
}
