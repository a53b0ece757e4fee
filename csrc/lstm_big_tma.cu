// Second generation of the large-hidden LSTM step kernels (BASELINE.json config 4; K10 of SURVEY.md §2.5): the
// per-step GEMM runs on the TMA-fed warp-specialised mainloop of csrc/tma_pipe.cuh instead of the cp.async one.
//
// Forward step (grid = batch tiles x 4H/128 column tiles, one 128x128 tile per CTA):
//   warp 0   : TMA producer — the h_{t-1} K-blocks (K-major), the U K-blocks (N-major, permuted gate columns) and,
//              off the critical path, this tile's slice of the hoisted input projection zx_t into shared memory
//   warp 1   : MMA issuer, 16 K-blocks of Z = h_{t-1}·U into TMEM
//   warps 2-5: epilogue, one batch row per thread: reads the four gate pre-activations of 32 hidden units from
//              TMEM, adds zx (bias already folded in), applies the Keras gate math, and writes c_t (fp32),
//              h_t (bf16 = next step's A operand) and the post-activation gates for BPTT with 16-byte stores.
//              c_{t-1} is fetched into registers before the accumulator is ready.
// Column order of every 4H-wide tensor is the "tile-gate-unit" permutation of csrc/lstm_big.cu.
// (Keras LSTM semantics: /root/reference/ddl_nyiso_aztk.py:249-251.)
#include "tma_pipe.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

// zx[t][b][n] = x[b][t] * wih_perm[n] + bias_perm[n]   (first layer: input dimension 1)
__global__ void __launch_bounds__(256) lstm_big_zx1_kernel(const float* __restrict__ x, const float* __restrict__ wih,
                                                           const float* __restrict__ bias, __nv_bfloat16* __restrict__ zx,
                                                           int B, int T, int G) {
  const int row = blockIdx.x;                       // t * B + b
  const int t = row / B, b = row - t * B;
  const float xv = x[(size_t)b * T + t];
  uint4* dst = reinterpret_cast<uint4*>(zx + (size_t)row * G);
  for (int c = threadIdx.x; c < G / 8; c += blockDim.x) {
    const float4 w0 = reinterpret_cast<const float4*>(wih)[2 * c], w1 = reinterpret_cast<const float4*>(wih)[2 * c + 1];
    const float4 b0 = reinterpret_cast<const float4*>(bias)[2 * c], b1 = reinterpret_cast<const float4*>(bias)[2 * c + 1];
    uint4 o;
    o.x = pack_bf16x2(fmaf(xv, w0.x, b0.x), fmaf(xv, w0.y, b0.y));
    o.y = pack_bf16x2(fmaf(xv, w0.z, b0.z), fmaf(xv, w0.w, b0.w));
    o.z = pack_bf16x2(fmaf(xv, w1.x, b1.x), fmaf(xv, w1.y, b1.y));
    o.w = pack_bf16x2(fmaf(xv, w1.z, b1.z), fmaf(xv, w1.w, b1.w));
    dst[c] = o;
  }
}

struct BigFwd2 {
  const float* c_prev;       // [B][H]
  float* c_next;             // [B][H]
  __nv_bfloat16* h_next;     // [B][H]
  float* h_out_f32;          // [B][H] or nullptr
  __nv_bfloat16* gates;      // [B][4H] permuted, or nullptr
  int B, H, t;
};

constexpr int BF_THREADS = 192;
using FwdPipe = TmaPipe<128, false, true, 5>;
constexpr int BF_ZX_BYTES = 128 * 128 * 2;
constexpr int BF_SMEM = FwdPipe::STAGES * FwdPipe::STAGE_BYTES + BF_ZX_BYTES + 1024;

DDLS_DEVINL void unpack8(const uint4& q, float* f) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

__global__ void __launch_bounds__(BF_THREADS, 1)
lstm_big_fwd2_kernel(const __grid_constant__ CUtensorMap th, const __grid_constant__ CUtensorMap tu,
                     const __grid_constant__ CUtensorMap tz, BigFwd2 p) {
  using Pipe = FwdPipe;
  constexpr int S = Pipe::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* zx_s = smem + S * Pipe::STAGE_BYTES;
  __shared__ uint64_t full[S], empty[S], tfull, zfull;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * 128, n0 = blockIdx.y * 128;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&tfull, 1);
    mbar_init(&zfull, 1);
    mbar_fence_init();
    tma_prefetch_desc(&th); tma_prefetch_desc(&tu); tma_prefetch_desc(&tz);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const int kblocks = p.H / 64;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % S, ph = (kb / S) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], Pipe::STAGE_BYTES);
        uint8_t* st = smem + s * Pipe::STAGE_BYTES;
        Pipe::load_a(st, &full[s], &th, m0, kb * 64, p.t);
        Pipe::load_b(st, &full[s], &tu, n0, kb * 64, 0);
        if (kb == 1) {                               // after the first operands are on their way
          mbar_expect_tx(&zfull, BF_ZX_BYTES);
          tma_load_3d(smem_u32(zx_s), &tz, &zfull, n0, m0, p.t);
          tma_load_3d(smem_u32(zx_s) + 16384, &tz, &zfull, n0 + 64, m0, p.t);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % S, ph = (kb / S) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        Pipe::mma_stage(smem + s * Pipe::STAGE_BYTES, tmem, kb == 0);
        umma_commit(&empty[s]);
      }
      umma_commit(&tfull);
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane, b = m0 + r;
    const bool live = b < p.B;
    const int unit0 = (n0 >> 7) * 32;
    const int G = 4 * p.H;
    float cprev[32];
    if (live) {
      const float4* cp = reinterpret_cast<const float4*>(p.c_prev + (size_t)b * p.H + unit0);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 q = __ldg(cp + j);
        cprev[4 * j] = q.x; cprev[4 * j + 1] = q.y; cprev[4 * j + 2] = q.z; cprev[4 * j + 3] = q.w;
      }
    }
    mbar_wait(&zfull, 0);
    mbar_wait(&tfull, 0);
    tc_fence_after();
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      uint32_t acc[4][16];
#pragma unroll
      for (int g = 0; g < 4; ++g) tmem_ld16(tmem + (static_cast<uint32_t>(quad * 32) << 16) + g * 32 + hh * 16, acc[g]);
      tmem_ld_wait();
      float z[4][16];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        // columns g*32 + hh*16 .. +15 of this row: 16-byte chunks 4g+2hh and 4g+2hh+1 -> box g>>1, chunk (.. & 7), swizzled
        const uint8_t* box = zx_s + (g >> 1) * 16384 + r * 128;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int chunk = ((4 * g + 2 * hh + c) & 7) ^ (r & 7);
          const uint4 q = *reinterpret_cast<const uint4*>(box + chunk * 16);
          float f[8];
          unpack8(q, f);
#pragma unroll
          for (int j = 0; j < 8; ++j) z[g][c * 8 + j] = __uint_as_float(acc[g][c * 8 + j]) + f[j];
        }
      }
      if (live) {
        float cn[16], hv[16];
        uint32_t gp[4][8];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float gi = __saturatef(fmaf(0.2f, z[0][j], 0.5f)), gf = __saturatef(fmaf(0.2f, z[1][j], 0.5f));
          const float gg = fast_tanh(z[2][j]), go = __saturatef(fmaf(0.2f, z[3][j], 0.5f));
          cn[j] = fmaf(gf, cprev[hh * 16 + j], gi * gg);
          hv[j] = go * fast_tanh(cn[j]);
          z[0][j] = gi; z[1][j] = gf; z[2][j] = gg; z[3][j] = go;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
          for (int j = 0; j < 8; ++j) gp[g][j] = pack_bf16x2(z[g][2 * j], z[g][2 * j + 1]);
        const size_t o = (size_t)b * p.H + unit0 + hh * 16;
        float4* cd = reinterpret_cast<float4*>(p.c_next + o);
#pragma unroll
        for (int j = 0; j < 4; ++j) cd[j] = make_float4(cn[4 * j], cn[4 * j + 1], cn[4 * j + 2], cn[4 * j + 3]);
        uint4* hd = reinterpret_cast<uint4*>(p.h_next + o);
#pragma unroll
        for (int j = 0; j < 2; ++j)
          hd[j] = make_uint4(pack_bf16x2(hv[8 * j], hv[8 * j + 1]), pack_bf16x2(hv[8 * j + 2], hv[8 * j + 3]),
                             pack_bf16x2(hv[8 * j + 4], hv[8 * j + 5]), pack_bf16x2(hv[8 * j + 6], hv[8 * j + 7]));
        if (p.h_out_f32 != nullptr) {
          float4* fd = reinterpret_cast<float4*>(p.h_out_f32 + o);
#pragma unroll
          for (int j = 0; j < 4; ++j) fd[j] = make_float4(hv[4 * j], hv[4 * j + 1], hv[4 * j + 2], hv[4 * j + 3]);
        }
        if (p.gates != nullptr) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4* gd = reinterpret_cast<uint4*>(p.gates + (size_t)b * G + n0 + g * 32 + hh * 16);
            gd[0] = make_uint4(gp[g][0], gp[g][1], gp[g][2], gp[g][3]);
            gd[1] = make_uint4(gp[g][4], gp[g][5], gp[g][6], gp[g][7]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}


// ---------------------------------------------------------------------------------------------------------------
// Backward step: dh_{t-1} = dZ_t · Uᵀ is a [B x H] result with K = 4H.  32 output tiles of 128x128 would leave most
// SMs idle, and narrow tiles re-read the long-K operands once per tile (the first-generation kernel moved 160 MB
// per step).  Here four CTAs of a thread-block cluster split K (one gate-block of 4H each, 16 K-blocks at H = 1024)
// and exchange partial accumulators through distributed shared memory: CTA q of the cluster receives batch rows
// 32q..32q+31 of all four partials, sums them and runs the gate-derivative epilogue for those rows with
// lane = 4 consecutive hidden units (coalesced 8/16-byte accesses).  128 CTAs share both the GEMM and the
// elementwise work; nothing but dZ_{t-1} and the running dc is written.
struct BigBwd2 {
  const __nv_bfloat16* dout_bf16;  // [B][H] gradient from the layer above at this step, or nullptr
  const float* dout_f32;           // [B][H] (top layer, last step), or nullptr
  const __nv_bfloat16* gates;      // [B][4H] permuted gates of step t
  const float* c_t;                // [B][H]
  const float* c_prev;             // [B][H]
  float* dc;                       // [B][H] running cell-state gradient (read-modify-write)
  __nv_bfloat16* dz;               // [B][4H] permuted dZ_t (output)
  int B, H, t1;                    // t1 = t + 1: time index of dZ_{t+1} in the dz tensor map
  int use_acc;                     // 0 at t = T-1 (no dh from the future)
  int first;                       // 1 at t = T-1: dc starts from zero
};

using BwdPipe = TmaPipe<128, false, false, 4>;
constexpr int BB_RECV_LD = 132;                                   // floats per received row (padded: conflict-free v4)
constexpr int BB_RECV_BYTES = 4 * 32 * BB_RECV_LD * 4;
constexpr int BB_THREADS = 320;                                   // producer, MMA, 4 TMEM warps + 4 more for phase 2
constexpr int BB_ROWS_PER_WARP = 4;
constexpr int BB_SMEM = BwdPipe::STAGES * BwdPipe::STAGE_BYTES + BB_RECV_BYTES + 1024;

DDLS_DEVINL uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
DDLS_DEVINL void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
DDLS_DEVINL void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
DDLS_DEVINL uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
DDLS_DEVINL void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __cluster_dims__(1, 1, 4) __launch_bounds__(BB_THREADS, 1)
lstm_big_bwd2_kernel(const __grid_constant__ CUtensorMap tdz, const __grid_constant__ CUtensorMap tu, BigBwd2 p) {
  using Pipe = BwdPipe;
  constexpr int S = Pipe::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  float* recv = reinterpret_cast<float*>(smem + S * Pipe::STAGE_BYTES);     // [slot 4][row 32][BB_RECV_LD]
  __shared__ uint64_t full[S], empty[S], tfull;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * 128, n0 = blockIdx.y * 128;
  const uint32_t crank = cluster_ctarank();
  const int G = 4 * p.H;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&tfull, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tdz); tma_prefetch_desc(&tu);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_arrive();                                   // "this CTA is running": peers wait for it before their first remote store
  const uint32_t tmem = tmem_slot;
  const int kblocks = p.H / 64;                       // this CTA's quarter of K = 4H
  const int k_base = crank * p.H;
  // Phase-2 operands do not depend on the GEMM: warps 2..9 fetch theirs now, behind the mainloop.
  const int e = warp - 2;
  const int unit = n0 + 4 * lane;
  const int pc = (unit >> 5) * 128 + (unit & 31);
  uint2 pre_g[BB_ROWS_PER_WARP][4], pre_do[BB_ROWS_PER_WARP];
  float4 pre_ct[BB_ROWS_PER_WARP], pre_cp[BB_ROWS_PER_WARP], pre_dc[BB_ROWS_PER_WARP];
  if (warp >= 2) {
#pragma unroll
    for (int rr = 0; rr < BB_ROWS_PER_WARP; ++rr) {
      const int b = m0 + 32 * (int)crank + e * BB_ROWS_PER_WARP + rr;
      pre_dc[rr] = make_float4(0.f, 0.f, 0.f, 0.f);
      pre_do[rr] = make_uint2(0u, 0u);
      if (b < p.B) {
        const size_t o = (size_t)b * p.H + unit;
        const __nv_bfloat16* gs = p.gates + (size_t)b * G + pc;
#pragma unroll
        for (int g = 0; g < 4; ++g) pre_g[rr][g] = *reinterpret_cast<const uint2*>(gs + g * 32);
        pre_ct[rr] = *reinterpret_cast<const float4*>(p.c_t + o);
        pre_cp[rr] = *reinterpret_cast<const float4*>(p.c_prev + o);
        if (!p.first) pre_dc[rr] = *reinterpret_cast<const float4*>(p.dc + o);
        if (p.dout_bf16 != nullptr) pre_do[rr] = *reinterpret_cast<const uint2*>(p.dout_bf16 + o);
      }
    }
  }

  if (p.use_acc) {
    if (warp == 0) {
      if (lane == 0) {
        for (int kb = 0; kb < kblocks; ++kb) {
          const int s = kb % S, ph = (kb / S) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], Pipe::STAGE_BYTES);
          uint8_t* st = smem + s * Pipe::STAGE_BYTES;
          Pipe::load_a(st, &full[s], &tdz, m0, k_base + kb * 64, p.t1);
          Pipe::load_b(st, &full[s], &tu, n0, k_base + kb * 64, 0);
        }
      }
      __syncwarp();
    } else if (warp == 1) {
      if (lane == 0) {
        for (int kb = 0; kb < kblocks; ++kb) {
          const int s = kb % S, ph = (kb / S) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          Pipe::mma_stage(smem + s * Pipe::STAGE_BYTES, tmem, kb == 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull);
      }
      __syncwarp();
    }
  }
  cluster_wait();                                     // every CTA of the cluster has started
  if (p.use_acc && warp >= 2 && warp < 6) {
    // phase 1: rows 32q..32q+31 of this CTA's partial go to CTA q, slot = my rank
    const int quad = warp & 3;
    mbar_wait(&tfull, 0);
    tc_fence_after();
    const uint32_t dst = map_to_cta(smem_u32(recv + (crank * 32 + lane) * BB_RECV_LD), quad);
#pragma unroll
    for (int c = 0; c < 128; c += 16) {
      uint32_t v[16];
      tmem_ld16(tmem + (static_cast<uint32_t>(quad * 32) << 16) + c, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        st_cluster_v4(dst + (c + j) * 4, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                      __uint_as_float(v[j + 3]));
    }
  }
  __syncwarp();
  cluster_arrive();
  cluster_wait();                                     // all partials have landed; no remote access after this point
  if (warp >= 2) {
    // phase 2: this CTA owns batch rows m0 + 32*crank .. +31, units n0 .. n0+127; lane = 4 consecutive units
#pragma unroll
    for (int rr = 0; rr < BB_ROWS_PER_WARP; ++rr) {
      const int row = e * BB_ROWS_PER_WARP + rr;
      const int b = m0 + 32 * (int)crank + row;
      if (b >= p.B) continue;
      float dh[4] = {0.f, 0.f, 0.f, 0.f};
      if (p.use_acc) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const float4 q = *reinterpret_cast<const float4*>(recv + (s * 32 + row) * BB_RECV_LD + 4 * lane);
          dh[0] += q.x; dh[1] += q.y; dh[2] += q.z; dh[3] += q.w;
        }
      }
      const size_t o = (size_t)b * p.H + unit;
      if (p.dout_bf16 != nullptr) {
        const uint2 q = pre_do[rr];
        dh[0] += __uint_as_float(q.x << 16); dh[1] += __uint_as_float(q.x & 0xFFFF0000u);
        dh[2] += __uint_as_float(q.y << 16); dh[3] += __uint_as_float(q.y & 0xFFFF0000u);
      }
      if (p.dout_f32 != nullptr) {
        const float4 q = *reinterpret_cast<const float4*>(p.dout_f32 + o);
        dh[0] += q.x; dh[1] += q.y; dh[2] += q.z; dh[3] += q.w;
      }
      float gt[4][4];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const uint2 q = pre_g[rr][g];
        gt[g][0] = __uint_as_float(q.x << 16); gt[g][1] = __uint_as_float(q.x & 0xFFFF0000u);
        gt[g][2] = __uint_as_float(q.y << 16); gt[g][3] = __uint_as_float(q.y & 0xFFFF0000u);
      }
      const float4 ct4 = pre_ct[rr], cp4 = pre_cp[rr], dc4 = pre_dc[rr];
      const float ct[4] = {ct4.x, ct4.y, ct4.z, ct4.w}, cpv[4] = {cp4.x, cp4.y, cp4.z, cp4.w};
      float dcv[4] = {dc4.x, dc4.y, dc4.z, dc4.w};
      float zi[4], zf[4], zg[4], zo[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float gi = gt[0][j], gf = gt[1][j], gg = gt[2][j], go = gt[3][j];
        const float tc = fast_tanh(ct[j]);
        const float d = fmaf(dh[j] * go, 1.f - tc * tc, dcv[j]);
        zo[j] = dh[j] * tc * ((go > 0.f && go < 1.f) ? 0.2f : 0.f);
        zi[j] = d * gg * ((gi > 0.f && gi < 1.f) ? 0.2f : 0.f);
        zf[j] = d * cpv[j] * ((gf > 0.f && gf < 1.f) ? 0.2f : 0.f);
        zg[j] = d * gi * (1.f - gg * gg);
        dcv[j] = d * gf;
      }
      *reinterpret_cast<float4*>(p.dc + o) = make_float4(dcv[0], dcv[1], dcv[2], dcv[3]);
      __nv_bfloat16* zd = p.dz + (size_t)b * G + pc;
      *reinterpret_cast<uint2*>(zd) = make_uint2(pack_bf16x2(zi[0], zi[1]), pack_bf16x2(zi[2], zi[3]));
      *reinterpret_cast<uint2*>(zd + 32) = make_uint2(pack_bf16x2(zf[0], zf[1]), pack_bf16x2(zf[2], zf[3]));
      *reinterpret_cast<uint2*>(zd + 64) = make_uint2(pack_bf16x2(zg[0], zg[1]), pack_bf16x2(zg[2], zg[3]));
      *reinterpret_cast<uint2*>(zd + 96) = make_uint2(pack_bf16x2(zo[0], zo[1]), pack_bf16x2(zo[2], zo[3]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

}  // namespace ddls

using namespace ddls;
#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int ddls_lstm_big_zx1(const float* x, const float* wih_perm, const float* bias_perm, void* zx, int B, int T, int H,
                                 void* stream) {
  if (H % 128 != 0 || B <= 0 || T <= 0) { set_last_error("ddls_lstm_big_zx1: bad shape B=%d T=%d H=%d", B, T, H); return 1; }
  lstm_big_zx1_kernel<<<B * T, 256, 0, ST(stream)>>>(x, wih_perm, bias_perm, static_cast<__nv_bfloat16*>(zx), B, T, 4 * H);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_big_zx1_kernel launch");
}

// hs: [T+1][B][H] bf16 (hs[0] = initial state), zx: [T][B][4H] bf16 (bias folded in), cs: [T+1][B][H] fp32,
// gates: [T][B][4H] bf16 or nullptr.  Runs step t: reads hs[t], cs[t], zx[t]; writes hs[t+1], cs[t+1], gates[t].
extern "C" int ddls_lstm_big_step_fwd2(const void* hs, const void* u_perm, const void* zx, float* cs, void* gates,
                                       float* h_out_f32, int B, int T, int H, int t, void* stream) {
  if (H % 128 != 0 || B <= 0 || t < 0 || t >= T) { set_last_error("ddls_lstm_big_step_fwd2: bad shape"); return 1; }
  const int G = 4 * H;
  CUtensorMap th, tu, tz;
  {
    const uint64_t dims[3] = {(uint64_t)H, (uint64_t)B, (uint64_t)T + 1};
    const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * B};
    const uint32_t box[3] = {64, 128, 1};
    if (tmap_bf16_sw128(&th, hs, dims, strides, box)) return 1;
  }
  {
    const uint64_t dims[3] = {(uint64_t)G, (uint64_t)H, 1};
    const uint64_t strides[2] = {(uint64_t)G * 2, (uint64_t)G * 2 * H};
    const uint32_t box[3] = {64, 64, 1};
    if (tmap_bf16_sw128(&tu, u_perm, dims, strides, box)) return 1;
  }
  {
    const uint64_t dims[3] = {(uint64_t)G, (uint64_t)B, (uint64_t)T};
    const uint64_t strides[2] = {(uint64_t)G * 2, (uint64_t)G * 2 * B};
    const uint32_t box[3] = {64, 128, 1};
    if (tmap_bf16_sw128(&tz, zx, dims, strides, box)) return 1;
  }
  const size_t bh = (size_t)B * H;
  BigFwd2 p{cs + (size_t)t * bh, cs + (size_t)(t + 1) * bh, static_cast<__nv_bfloat16*>(const_cast<void*>(hs)) + (size_t)(t + 1) * bh,
            h_out_f32, gates ? static_cast<__nv_bfloat16*>(gates) + (size_t)t * B * G : nullptr, B, H, t};
  DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_big_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BF_SMEM));
  dim3 grid((B + 127) / 128, G / 128);
  lstm_big_fwd2_kernel<<<grid, BF_THREADS, BF_SMEM, ST(stream)>>>(th, tu, tz, p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_big_fwd2_kernel launch");
}

// dz: [T][B][4H] bf16 (permuted), gates: [T][B][4H], cs: [T+1][B][H] fp32, dc: [B][H] fp32 scratch.  Runs BPTT step t:
// reads dz[t+1] (unless t == T-1), gates[t], cs[t+1], cs[t], dout; writes dz[t] and dc.
extern "C" int ddls_lstm_big_step_bwd2(void* dz, const void* u_perm, const void* dout_bf16, const float* dout_f32,
                                       const void* gates, const float* cs, float* dc, int B, int T, int H, int t,
                                       void* stream) {
  if (H % 128 != 0 || B <= 0 || t < 0 || t >= T) { set_last_error("ddls_lstm_big_step_bwd2: bad shape"); return 1; }
  const int G = 4 * H;
  CUtensorMap tdz, tu;
  {
    const uint64_t dims[3] = {(uint64_t)G, (uint64_t)B, (uint64_t)T};
    const uint64_t strides[2] = {(uint64_t)G * 2, (uint64_t)G * 2 * B};
    const uint32_t box[3] = {64, 128, 1};
    if (tmap_bf16_sw128(&tdz, dz, dims, strides, box)) return 1;
  }
  {
    const uint64_t dims[3] = {(uint64_t)G, (uint64_t)H, 1};
    const uint64_t strides[2] = {(uint64_t)G * 2, (uint64_t)G * 2 * H};
    const uint32_t box[3] = {64, 128, 1};
    if (tmap_bf16_sw128(&tu, u_perm, dims, strides, box)) return 1;
  }
  const size_t bh = (size_t)B * H, bg = (size_t)B * G;
  const int last = t == T - 1;
  BigBwd2 p{static_cast<const __nv_bfloat16*>(dout_bf16), dout_f32, static_cast<const __nv_bfloat16*>(gates) + (size_t)t * bg,
            cs + (size_t)(t + 1) * bh, cs + (size_t)t * bh, dc, static_cast<__nv_bfloat16*>(dz) + (size_t)t * bg, B, H,
            last ? 0 : t + 1, last ? 0 : 1, last};
  DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_big_bwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BB_SMEM));
  dim3 grid((B + 127) / 128, H / 128, 4);
  lstm_big_bwd2_kernel<<<grid, BB_THREADS, BB_SMEM, ST(stream)>>>(tdz, tu, p);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "lstm_big_bwd2_kernel launch");
}
