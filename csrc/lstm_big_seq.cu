// Whole-sequence persistent kernels for the large-hidden LSTM (BASELINE.json config 4; K10 of SURVEY.md §2.5).
//
// The per-step kernels of csrc/lstm_big_tma.cu pay, 336 times per training step, for a launch, barrier/TMEM set-up,
// a cold TMA pipeline and a drain.  Here ONE cooperative launch per layer walks all T steps.  Each CTA keeps its
// output tile (128 batch rows x 32 hidden units, i.e. 128 permuted gate columns) for the whole sequence:
//   * the cell state c (forward) / the running dc (backward) stay in registers for all T steps,
//   * the weight K-blocks of step t+1 are prefetched by their own producer warp while step t's epilogue runs —
//     only the h_t (or dZ_t) operand has to wait for the step barrier,
//   * the step barrier is per batch tile (the 32 CTAs that produce the 128 rows a CTA consumes), a release-add on
//     one counter followed by acquire-polling in the operand producer; h_t goes generic-store -> fence.proxy.async
//     -> TMA load of the consumer.
// Forward roles: warp 0 weight producer, warp 1 MMA issuer, warps 2-9 epilogue, warp 10 h/zx producer.  The epilogue
// stores h_t first, fences and posts its arrival, and only then writes the BPTT stash (c_t, gates).
// Backward: K = 4H split over a 4-CTA cluster; partial accumulators are exchanged with bulk DSMEM copies that
// complete on the receiver's mbarrier (no per-step cluster barrier); warps 2-9 run the gate-derivative phase,
// warp 10 is the dZ producer.
// (Keras LSTM semantics: /root/reference/ddl_nyiso_aztk.py:249-251.)
#include <cstdio>
#include <cstdlib>

#include "tma_pipe.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

DDLS_DEVINL void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
DDLS_DEVINL void red_release_gpu_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DDLS_DEVINL void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
DDLS_DEVINL void red_relaxed_gpu_add(uint32_t* p, uint32_t v) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DDLS_DEVINL uint32_t ld_relaxed_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
DDLS_DEVINL uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Consumer-side fences after the step barrier (DDLS_SEQ_FENCE, default 0).  The producers make h_t / dZ_t visible at
// gpu scope (__threadfence) before their arrival is posted, and the consumer touches that data only through TMA, which
// reads L2 — so mode 0 issues the loads as soon as the counter is seen.  1 adds fence.acq_rel.gpu, 2 adds
// fence.proxy.async on top (the by-the-book sequence; ~1 us per step slower, measured with tools/lstm_seq_trace.py).
__device__ int g_fence_mode = 0;
// Spin until *p >= target.  Bounded: a lost arrival traps with a message instead of hanging the GPU.
DDLS_DEVINL void wait_counter(const uint32_t* p, uint32_t target, int step) {
  uint32_t spins = 0;
  while (ld_relaxed_gpu(p) < target) {
    if (++spins > (1u << 26)) {
      printf("ddls lstm_big_seq: step barrier timed out (block %d,%d,%d step %d: %u < %u)\n", blockIdx.x, blockIdx.y,
             blockIdx.z, step, ld_acquire_gpu(p), target);
      __trap();
    }
  }
  if (g_fence_mode >= 1) asm volatile("fence.acq_rel.gpu;" ::: "memory");
  if (g_fence_mode == 2) asm volatile("fence.proxy.async;" ::: "memory");
}

DDLS_DEVINL void stamp(long long* trace, int t, int slot) {
  if (trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) trace[t * 8 + slot] = clock64();
}

DDLS_DEVINL void stamp_all(long long* trace, int T, int t, int slot) {     // every CTA: global ns clock at slot (0/1)
  if (trace != nullptr) {
    unsigned long long ns;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
    const int cta = blockIdx.y * gridDim.x + blockIdx.x, ncta = gridDim.x * gridDim.y;
    trace[T * 8 + ((size_t)t * 2 + slot) * ncta + cta] = (long long)ns;
  }
}

DDLS_DEVINL void unpack8f(const uint4& q, float* f) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

struct BigSeqFwd {
  float* cs;               // [T+1][B][H]  cs[0] = initial state; cs[t+1] written when save != 0
  __nv_bfloat16* hs;       // [T+1][B][H]  hs[0] = initial state; hs[t+1] written every step
  __nv_bfloat16* gates;    // [T][B][4H] permuted post-activation gates, or nullptr
  float* h_out_f32;        // [B][H] final hidden state, or nullptr
  uint32_t* sync;          // one arrival counter per batch tile of this launch, zero on entry
  int B, H, T, m_tile0;
  long long* trace;        // optional [T][8] clock64 stamps of CTA (0,0) (tools/lstm_seq_trace.py), or nullptr
};

constexpr int SF_THREADS = 352;                                   // W-producer, MMA, 8 epilogue warps, h/zx-producer
using SeqFwdPipe = TmaPipe<128, false, true, 5>;
constexpr int SF_ZX_BYTES = 128 * 128 * 2;
constexpr int SF_SMEM = SeqFwdPipe::STAGES * SeqFwdPipe::STAGE_BYTES + SF_ZX_BYTES + 1024;

__global__ void __launch_bounds__(SF_THREADS, 1)
lstm_big_seq_fwd_kernel(const __grid_constant__ CUtensorMap th, const __grid_constant__ CUtensorMap tu,
                        const __grid_constant__ CUtensorMap tz, BigSeqFwd p) {
  using Pipe = SeqFwdPipe;
  constexpr int S = Pipe::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* zx_s = smem + S * Pipe::STAGE_BYTES;
  __shared__ uint64_t full[S], empty[S], tfull[2], tempty[2], zfull, zempty;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = (p.m_tile0 + blockIdx.x) * 128, n0 = blockIdx.y * 128;
  uint32_t* counter = p.sync + blockIdx.x * 32;              // one 128-byte line per batch tile
  const uint32_t group = gridDim.y;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 2); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 8); }
    mbar_init(&zfull, 1);
    mbar_init(&zempty, 8);
    mbar_fence_init();
    tma_prefetch_desc(&th); tma_prefetch_desc(&tu); tma_prefetch_desc(&tz);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const int kblocks = p.H / 64;
  const int T = p.T;

  if (warp == 0) {
    if (lane == 0) {                                  // weights: never wait for the recurrence
      int it = 0;
      for (int t = 0; t < T; ++t)
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], Pipe::B_BYTES);
          Pipe::load_b(smem + s * Pipe::STAGE_BYTES, &full[s], &tu, n0, kb * 64, 0);
        }
    }
  } else if (warp == 10) {
    if (lane == 0) {                                  // h_t and zx_t: gated by the step barrier
      int it = 0;
      for (int t = 0; t < T; ++t) {
        if (t > 0) {
          mbar_wait(&zempty, (t - 1) & 1);
          wait_counter(counter, group * (uint32_t)t, t);
        }
        stamp(p.trace, t, 0);                         // step barrier passed
        stamp_all(p.trace, T, t, 0);
        mbar_expect_tx(&zfull, SF_ZX_BYTES);
        tma_load_3d(smem_u32(zx_s), &tz, &zfull, n0, m0, t);
        tma_load_3d(smem_u32(zx_s) + 16384, &tz, &zfull, n0 + 64, m0, t);
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], Pipe::A_BYTES);
          Pipe::load_a(smem + s * Pipe::STAGE_BYTES, &full[s], &th, m0, kb * 64, t);
        }
        stamp(p.trace, t, 1);                         // last h K-block issued
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int it = 0;
      for (int t = 0; t < T; ++t) {
        const int acc = t & 1;
        if (t >= 2) mbar_wait(&tempty[acc], ((t - 2) >> 1) & 1);
        tc_fence_after();
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % S, ph = (it / S) & 1;
          mbar_wait(&full[s], ph);
          if (kb == 0) stamp(p.trace, t, 2);          // first K-block landed
          tc_fence_after();
          Pipe::mma_stage(smem + s * Pipe::STAGE_BYTES, tmem + acc * 128, kb == 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull[acc]);
        stamp(p.trace, t, 3);                         // last MMA issued
      }
    }
  } else {
    // epilogue: two warps per TMEM lane quadrant; a thread owns one batch row and 16 of the tile's 32 hidden units
    const int quad = warp & 3, hh = (warp - 2) >> 2;
    const int r = quad * 32 + lane, b = m0 + r;
    const bool live = b < p.B;
    const int unit0 = (n0 >> 7) * 32 + hh * 16;
    const int G = 4 * p.H;
    const size_t bh = (size_t)p.B * p.H;
    const size_t row_o = (size_t)b * p.H + unit0;
    float c[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) c[j] = 0.f;
    if (live) {
      const float4* cp = reinterpret_cast<const float4*>(p.cs + row_o);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 q = __ldg(cp + j);
        c[4 * j] = q.x; c[4 * j + 1] = q.y; c[4 * j + 2] = q.z; c[4 * j + 3] = q.w;
      }
    }
    for (int t = 0; t < T; ++t) {
      const int acc = t & 1;
      mbar_wait(&zfull, t & 1);
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 64) stamp(p.trace, t, 4);    // accumulator complete
      uint32_t av[4][16];
#pragma unroll
      for (int g = 0; g < 4; ++g)
        tmem_ld16(tmem + (static_cast<uint32_t>(quad * 32) << 16) + acc * 128 + g * 32 + hh * 16, av[g]);
      tmem_ld_wait();
      float z[4][16];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const uint8_t* box = zx_s + (g >> 1) * 16384 + r * 128;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int chunk = ((4 * g + 2 * hh + cc) & 7) ^ (r & 7);
          const uint4 q = *reinterpret_cast<const uint4*>(box + chunk * 16);
          float f[8];
          unpack8f(q, f);
#pragma unroll
          for (int j = 0; j < 8; ++j) z[g][cc * 8 + j] = __uint_as_float(av[g][cc * 8 + j]) + f[j];
        }
      }
      tc_fence_before();                               // accumulator and zx tile are free again
      __syncwarp();
      if (lane == 0) { mbar_arrive(&tempty[acc]); mbar_arrive(&zempty); }
      float hv[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float gi = __saturatef(fmaf(0.2f, z[0][j], 0.5f)), gf = __saturatef(fmaf(0.2f, z[1][j], 0.5f));
        const float gg = fast_tanh(z[2][j]), go = __saturatef(fmaf(0.2f, z[3][j], 0.5f));
        const float cn = fmaf(gf, c[j], gi * gg);
        c[j] = cn;
        hv[j] = go * fast_tanh(cn);
        z[0][j] = gi; z[1][j] = gf; z[2][j] = gg; z[3][j] = go;
      }
      const size_t o = (size_t)(t + 1) * bh + row_o;
      // 1. h_t, the only thing the next step waits for
      if (live) {
        uint4* hd = reinterpret_cast<uint4*>(p.hs + o);
#pragma unroll
        for (int j = 0; j < 2; ++j)
          hd[j] = make_uint4(pack_bf16x2(hv[8 * j], hv[8 * j + 1]), pack_bf16x2(hv[8 * j + 2], hv[8 * j + 3]),
                             pack_bf16x2(hv[8 * j + 4], hv[8 * j + 5]), pack_bf16x2(hv[8 * j + 6], hv[8 * j + 7]));
      }
      if (threadIdx.x == 64) stamp(p.trace, t, 5);    // h stores issued
      fence_acq_rel_gpu();                            // this thread's stores are visible at gpu scope
      named_bar_sync(1, 256);
      if (warp == 2 && lane == 0) {
        stamp(p.trace, t, 6);                         // fences done
        red_relaxed_gpu_add(counter, 1u);
        stamp(p.trace, t, 7);
        stamp_all(p.trace, T, t, 1);
      }
      // 2. the BPTT stash, off the critical path
      if (live) {
        if (p.gates != nullptr) {
          float4* cd = reinterpret_cast<float4*>(p.cs + o);
#pragma unroll
          for (int j = 0; j < 4; ++j) cd[j] = make_float4(c[4 * j], c[4 * j + 1], c[4 * j + 2], c[4 * j + 3]);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4* gd = reinterpret_cast<uint4*>(p.gates + ((size_t)t * p.B + b) * G + n0 + g * 32 + hh * 16);
            gd[0] = make_uint4(pack_bf16x2(z[g][0], z[g][1]), pack_bf16x2(z[g][2], z[g][3]),
                               pack_bf16x2(z[g][4], z[g][5]), pack_bf16x2(z[g][6], z[g][7]));
            gd[1] = make_uint4(pack_bf16x2(z[g][8], z[g][9]), pack_bf16x2(z[g][10], z[g][11]),
                               pack_bf16x2(z[g][12], z[g][13]), pack_bf16x2(z[g][14], z[g][15]));
          }
        }
        if (t == T - 1 && p.h_out_f32 != nullptr) {
          float4* fd = reinterpret_cast<float4*>(p.h_out_f32 + row_o);
#pragma unroll
          for (int j = 0; j < 4; ++j) fd[j] = make_float4(hv[4 * j], hv[4 * j + 1], hv[4 * j + 2], hv[4 * j + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}


// ---------------------------------------------------------------------------------------------------------------
struct BigSeqBwd {
  __nv_bfloat16* dz;             // [T][B][4H] permuted (output; also next step's A operand through the tensor map)
  const __nv_bfloat16* dout;     // [T][B][H] gradient from the layer above per step, or nullptr
  const float* dout_last;        // [B][H] fp32 gradient at t = T-1 (top layer), or nullptr
  const __nv_bfloat16* gates;    // [T][B][4H]
  const float* cs;               // [T+1][B][H]
  uint32_t* sync;
  int B, H, T, m_tile0;
  long long* trace;              // optional [T][8] clock64 stamps of CTA (0,0,0), or nullptr
};

constexpr int SB_THREADS = 352;                                   // W-producer, MMA, 8 phase warps (4 read TMEM), dZ-producer
constexpr int SB_ROWS = 4;                                        // batch rows per phase-2 warp
using SeqBwdPipe = TmaPipe<128, false, false, 4>;
constexpr int SB_RECV_LD = 132;
constexpr int SB_RECV_BYTES = 4 * 32 * SB_RECV_LD * 4;
constexpr int SB_SMEM = SeqBwdPipe::STAGES * SeqBwdPipe::STAGE_BYTES + SB_RECV_BYTES + 1024;

DDLS_DEVINL uint32_t seq_cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
DDLS_DEVINL void seq_cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
DDLS_DEVINL void seq_cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
DDLS_DEVINL uint32_t seq_map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
DDLS_DEVINL void seq_st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(SB_THREADS, 1)
lstm_big_seq_bwd_kernel(const __grid_constant__ CUtensorMap tdz, const __grid_constant__ CUtensorMap tu, BigSeqBwd p) {
  using Pipe = SeqBwdPipe;
  constexpr int S = Pipe::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  float* recv = reinterpret_cast<float*>(smem + S * Pipe::STAGE_BYTES);     // [slot 4][row 32][SB_RECV_LD]
  __shared__ uint64_t full[S], empty[S], tfull, tempty, recv_full, stage_free;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = (p.m_tile0 + blockIdx.x) * 128, n0 = blockIdx.y * 128;
  const uint32_t crank = seq_cluster_ctarank();
  uint32_t* counter = p.sync + blockIdx.x * 32;
  const uint32_t group = gridDim.y * gridDim.z;
  const int G = 4 * p.H, T = p.T;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 2); mbar_init(&empty[s], 1); }
    mbar_init(&tfull, 1);
    mbar_init(&tempty, 8);
    mbar_init(&recv_full, 1);
    mbar_init(&stage_free, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tdz); tma_prefetch_desc(&tu);
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  seq_cluster_arrive();                               // all four CTAs are running before any remote store
  seq_cluster_wait();
  const uint32_t tmem = tmem_slot;
  const int kblocks = p.H / 64;
  const int k_base = crank * p.H;
  const size_t bh = (size_t)p.B * p.H, bg = (size_t)p.B * G;

  // phase-2 ownership: batch rows m0 + 32*crank + e*4 .. +3, hidden units n0 + 4*lane .. +3
  const int e = warp - 2;
  const int unit = n0 + 4 * lane;
  const int pc = (unit >> 5) * 128 + (unit & 31);
  float4 dc[SB_ROWS];
#pragma unroll
  for (int rr = 0; rr < SB_ROWS; ++rr) dc[rr] = make_float4(0.f, 0.f, 0.f, 0.f);

  // Steps run t = T-1 .. 0; step i = T-1-t.  Step 0 has no dh from the future (no GEMM).
  for (int i = 0; i < T; ++i) {
    const int t = T - 1 - i;
    if (warp == 0) {
      if (lane == 0 && i > 0) {
        if (i >= 2) mbar_wait(&stage_free, (i - 2) & 1);     // last step's staging copies have left the ring
        for (int kb = 0; kb < kblocks; ++kb) {
          const int it = (i - 1) * kblocks + kb, s = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], Pipe::B_BYTES);
          Pipe::load_b(smem + s * Pipe::STAGE_BYTES, &full[s], &tu, n0, k_base + kb * 64, 0);
        }
      }
    } else if (warp == 10) {
      if (lane == 0 && i > 0) {
        wait_counter(counter, group * (uint32_t)i, i);
        if (i >= 2) mbar_wait(&stage_free, (i - 2) & 1);
        stamp(p.trace, i, 0);
        for (int kb = 0; kb < kblocks; ++kb) {
          const int it = (i - 1) * kblocks + kb, s = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], Pipe::A_BYTES);
          Pipe::load_a(smem + s * Pipe::STAGE_BYTES, &full[s], &tdz, m0, k_base + kb * 64, t + 1);
        }
      }
    } else if (warp == 1) {
      if (lane == 0 && i > 0) {
        if (i >= 2) mbar_wait(&tempty, (i - 2) & 1);
        tc_fence_after();
        for (int kb = 0; kb < kblocks; ++kb) {
          const int it = (i - 1) * kblocks + kb, s = it % S, ph = (it / S) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          if (kb == 0) stamp(p.trace, i, 1);
          Pipe::mma_stage(smem + s * Pipe::STAGE_BYTES, tmem, kb == 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull);
        stamp(p.trace, i, 2);
      }
    }
    // Everything of the gate-derivative phase that does not depend on dh is fetched AND evaluated behind the mainloop:
    //   d = dh*cA + dc;  dZ_o = dh*cO;  dZ_i = d*cI;  dZ_f = d*cF;  dZ_g = d*cG;  dc' = d*cD.
    float cA[SB_ROWS][4], cO[SB_ROWS][4], cI[SB_ROWS][4], cF[SB_ROWS][4], cG[SB_ROWS][4], cD[SB_ROWS][4];
    uint2 pre_do[SB_ROWS];
    if (warp >= 2 && warp < 10) {
      uint2 pre_g[SB_ROWS][4];
      float4 pre_ct[SB_ROWS], pre_cp[SB_ROWS];
#pragma unroll
      for (int rr = 0; rr < SB_ROWS; ++rr) {
        const int b = m0 + 32 * (int)crank + e * SB_ROWS + rr;
        pre_do[rr] = make_uint2(0u, 0u);
        pre_ct[rr] = make_float4(0.f, 0.f, 0.f, 0.f);
        pre_cp[rr] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int g = 0; g < 4; ++g) pre_g[rr][g] = make_uint2(0u, 0u);
        if (b < p.B) {
          const size_t o = (size_t)b * p.H + unit;
          const __nv_bfloat16* gs = p.gates + (size_t)t * bg + (size_t)b * G + pc;
#pragma unroll
          for (int g = 0; g < 4; ++g) pre_g[rr][g] = *reinterpret_cast<const uint2*>(gs + g * 32);
          pre_ct[rr] = *reinterpret_cast<const float4*>(p.cs + (size_t)(t + 1) * bh + o);
          pre_cp[rr] = *reinterpret_cast<const float4*>(p.cs + (size_t)t * bh + o);
          if (p.dout != nullptr) pre_do[rr] = *reinterpret_cast<const uint2*>(p.dout + (size_t)t * bh + o);
        }
      }
#pragma unroll
      for (int rr = 0; rr < SB_ROWS; ++rr) {
        float gt[4][4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint2 q = pre_g[rr][g];
          gt[g][0] = __uint_as_float(q.x << 16); gt[g][1] = __uint_as_float(q.x & 0xFFFF0000u);
          gt[g][2] = __uint_as_float(q.y << 16); gt[g][3] = __uint_as_float(q.y & 0xFFFF0000u);
        }
        const float ct[4] = {pre_ct[rr].x, pre_ct[rr].y, pre_ct[rr].z, pre_ct[rr].w};
        const float cpv[4] = {pre_cp[rr].x, pre_cp[rr].y, pre_cp[rr].z, pre_cp[rr].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gi = gt[0][j], gf = gt[1][j], gg = gt[2][j], go = gt[3][j];
          const float tc = fast_tanh(ct[j]);
          cA[rr][j] = go * (1.f - tc * tc);
          cO[rr][j] = tc * ((go > 0.f && go < 1.f) ? 0.2f : 0.f);
          cI[rr][j] = gg * ((gi > 0.f && gi < 1.f) ? 0.2f : 0.f);
          cF[rr][j] = cpv[j] * ((gf > 0.f && gf < 1.f) ? 0.2f : 0.f);
          cG[rr][j] = gi * (1.f - gg * gg);
          cD[rr][j] = gf;
        }
      }
    }
    if (i > 0 && warp >= 2 && warp < 10) {
      // phase 1: the finished partial accumulator goes TMEM -> local staging (aliasing the operand ring, idle now)
      // -> four bulk DSMEM copies: rows 32q..32q+31 to CTA q's receive slot [my rank], completing on CTA q's
      // recv_full barrier.  No cluster barrier: the receiver waits on its own mbarrier.
      const int quad = warp & 3, chalf = (warp - 2) >> 2;
      float* stg = reinterpret_cast<float*>(smem);           // [row 128][SB_RECV_LD]
      mbar_wait(&tfull, (i - 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 64) stamp(p.trace, i, 3);
      float* srow = stg + (quad * 32 + lane) * SB_RECV_LD;
#pragma unroll
      for (int c = chalf * 64; c < chalf * 64 + 64; c += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + (static_cast<uint32_t>(quad * 32) << 16) + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; j += 4)
          *reinterpret_cast<float4*>(srow + c + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                 __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty);
      fence_proxy_async_smem();                              // staging was written by the generic proxy
      named_bar_sync(2, 256);
      if (warp == 2 && lane == 0) {
        constexpr uint32_t SLOT_BYTES = 32 * SB_RECV_LD * 4;
        mbar_expect_tx(&recv_full, 4 * SLOT_BYTES);          // four senders (one of them is this CTA)
#pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
          const uint32_t dst = seq_map_to_cta(smem_u32(recv + crank * 32 * SB_RECV_LD), q);
          const uint32_t bar = seq_map_to_cta(smem_u32(&recv_full), q);
          asm volatile(
              "cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
              "r"(smem_u32(stg + q * 32 * SB_RECV_LD)), "r"(SLOT_BYTES), "r"(bar)
              : "memory");
        }
        tma_bulk_commit();
        tma_bulk_wait_read();                                // staging has been read: the operand ring may be refilled
        mbar_arrive(&stage_free);
      }
      mbar_wait(&recv_full, (i - 1) & 1);                    // all four partials of my rows have landed
      if (threadIdx.x == 64) stamp(p.trace, i, 4);
    }
    if (warp >= 2 && warp < 10) {
#pragma unroll
      for (int rr = 0; rr < SB_ROWS; ++rr) {
        const int row = e * SB_ROWS + rr;
        const int b = m0 + 32 * (int)crank + row;
        if (b >= p.B) continue;
        float dh[4] = {0.f, 0.f, 0.f, 0.f};
        if (i > 0) {
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const float4 q = *reinterpret_cast<const float4*>(recv + (s * 32 + row) * SB_RECV_LD + 4 * lane);
            dh[0] += q.x; dh[1] += q.y; dh[2] += q.z; dh[3] += q.w;
          }
        }
        const size_t o = (size_t)b * p.H + unit;
        if (p.dout != nullptr) {
          const uint2 q = pre_do[rr];
          dh[0] += __uint_as_float(q.x << 16); dh[1] += __uint_as_float(q.x & 0xFFFF0000u);
          dh[2] += __uint_as_float(q.y << 16); dh[3] += __uint_as_float(q.y & 0xFFFF0000u);
        }
        if (i == 0 && p.dout_last != nullptr) {
          const float4 q = *reinterpret_cast<const float4*>(p.dout_last + o);
          dh[0] += q.x; dh[1] += q.y; dh[2] += q.z; dh[3] += q.w;
        }
        float dcv[4] = {dc[rr].x, dc[rr].y, dc[rr].z, dc[rr].w};
        float zi[4], zf[4], zg[4], zo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float d = fmaf(dh[j], cA[rr][j], dcv[j]);
          zo[j] = dh[j] * cO[rr][j];
          zi[j] = d * cI[rr][j];
          zf[j] = d * cF[rr][j];
          zg[j] = d * cG[rr][j];
          dcv[j] = d * cD[rr][j];
        }
        dc[rr] = make_float4(dcv[0], dcv[1], dcv[2], dcv[3]);
        __nv_bfloat16* zd = p.dz + (size_t)t * bg + (size_t)b * G + pc;
        *reinterpret_cast<uint2*>(zd) = make_uint2(pack_bf16x2(zi[0], zi[1]), pack_bf16x2(zi[2], zi[3]));
        *reinterpret_cast<uint2*>(zd + 32) = make_uint2(pack_bf16x2(zf[0], zf[1]), pack_bf16x2(zf[2], zf[3]));
        *reinterpret_cast<uint2*>(zd + 64) = make_uint2(pack_bf16x2(zg[0], zg[1]), pack_bf16x2(zg[2], zg[3]));
        *reinterpret_cast<uint2*>(zd + 96) = make_uint2(pack_bf16x2(zo[0], zo[1]), pack_bf16x2(zo[2], zo[3]));
      }
      // publish dZ_t
      if (threadIdx.x == 64) stamp(p.trace, i, 5);
      fence_acq_rel_gpu();                            // this thread's stores are visible at gpu scope
      named_bar_sync(1, 256);
      if (warp == 2 && lane == 0) {
        stamp(p.trace, i, 6);
        red_relaxed_gpu_add(counter, 1u);
        stamp(p.trace, i, 7);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  seq_cluster_arrive();                               // no CTA leaves while a peer could still address its memory
  seq_cluster_wait();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

}  // namespace ddls

using namespace ddls;
#define ST(s) static_cast<cudaStream_t>(s)

namespace {
int sm_count_seq() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

template <typename... Args>
int launch_coop(void (*kern)(Args...), dim3 grid, int threads, int smem, cudaStream_t st, int cluster_z, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeCooperative;
  attrs[0].val.cooperative = 1;
  int n = 1;
  if (cluster_z > 1) {
    attrs[1].id = cudaLaunchAttributeClusterDimension;
    attrs[1].val.clusterDim.x = 1; attrs[1].val.clusterDim.y = 1; attrs[1].val.clusterDim.z = cluster_z;
    n = 2;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = n;
  return check_cuda(cudaLaunchKernelEx(&cfg, kern, args...), "cooperative launch");
}
}  // namespace

// Largest number of 128-row batch tiles one cooperative launch can hold (0: the layer is too wide for this path).
extern "C" int ddls_lstm_big_seq_tiles_per_launch(int H, int backward) {
  const int per_tile = backward ? (H / 128) * 4 : (4 * H / 128);
  return sm_count_seq() / per_tile;
}

static long long* g_seq_trace = nullptr;
// Debug hook: device buffer of T*8 int64 that receives clock64 stamps of CTA (0,0) of the next forward launches.
extern "C" int ddls_lstm_big_seq_set_fence_mode(int m) {
  return check_cuda(cudaMemcpyToSymbol(ddls::g_fence_mode, &m, sizeof(int)), "set fence mode");
}
static int apply_env_fence_mode() {
  static bool done = false;
  if (done) return 0;
  done = true;
  const char* e = getenv("DDLS_SEQ_FENCE");
  return (e != nullptr) ? ddls_lstm_big_seq_set_fence_mode(atoi(e)) : 0;
}
extern "C" void ddls_lstm_big_seq_set_trace(void* buf) { g_seq_trace = static_cast<long long*>(buf); }

extern "C" int ddls_lstm_big_seq_fwd(void* hs, const void* u_perm, const void* zx, float* cs, void* gates, float* h_out_f32,
                                     void* sync, int B, int T, int H, void* stream) {
  if (H % 128 != 0 || B <= 0 || T <= 0) { set_last_error("ddls_lstm_big_seq_fwd: bad shape B=%d T=%d H=%d", B, T, H); return 1; }
  if (apply_env_fence_mode()) return 1;
  const int G = 4 * H;
  const int per_launch = ddls_lstm_big_seq_tiles_per_launch(H, 0);
  if (per_launch < 1) { set_last_error("ddls_lstm_big_seq_fwd: 4H/128 = %d tiles exceed the SM count", G / 128); return 1; }
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
  cudaStream_t st = ST(stream);
  const int m_tiles = (B + 127) / 128;
  DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_big_seq_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SF_SMEM));
  DDLS_CUDA_OK(cudaMemsetAsync(sync, 0, sizeof(uint32_t) * 32 * m_tiles, st));
  for (int m = 0; m < m_tiles; m += per_launch) {
    const int n = (m_tiles - m < per_launch) ? m_tiles - m : per_launch;
    BigSeqFwd p{cs, static_cast<__nv_bfloat16*>(hs), static_cast<__nv_bfloat16*>(gates), h_out_f32,
                static_cast<uint32_t*>(sync) + m * 32, B, H, T, m, g_seq_trace};
    if (launch_coop(lstm_big_seq_fwd_kernel, dim3(n, G / 128), SF_THREADS, SF_SMEM, st, 1, th, tu, tz, p)) return 1;
    count_launch(1);
  }
  return 0;
}

static long long* g_seq_trace_bwd = nullptr;
extern "C" void ddls_lstm_big_seq_set_trace_bwd(void* buf) { g_seq_trace_bwd = static_cast<long long*>(buf); }

extern "C" int ddls_lstm_big_seq_bwd(void* dz, const void* u_perm, const void* dout, const float* dout_last, const void* gates,
                                     const float* cs, void* sync, int B, int T, int H, void* stream) {
  if (H % 128 != 0 || B <= 0 || T <= 0) { set_last_error("ddls_lstm_big_seq_bwd: bad shape B=%d T=%d H=%d", B, T, H); return 1; }
  if (apply_env_fence_mode()) return 1;
  const int G = 4 * H;
  const int per_launch = ddls_lstm_big_seq_tiles_per_launch(H, 1);
  if (per_launch < 1) { set_last_error("ddls_lstm_big_seq_bwd: 4*H/128 = %d CTAs exceed the SM count", 4 * H / 128); return 1; }
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
  cudaStream_t st = ST(stream);
  const int m_tiles = (B + 127) / 128;
  DDLS_CUDA_OK(cudaFuncSetAttribute(lstm_big_seq_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SB_SMEM));
  DDLS_CUDA_OK(cudaMemsetAsync(sync, 0, sizeof(uint32_t) * 32 * m_tiles, st));
  for (int m = 0; m < m_tiles; m += per_launch) {
    const int n = (m_tiles - m < per_launch) ? m_tiles - m : per_launch;
    BigSeqBwd p{static_cast<__nv_bfloat16*>(dz), static_cast<const __nv_bfloat16*>(dout), dout_last,
                static_cast<const __nv_bfloat16*>(gates), cs, static_cast<uint32_t*>(sync) + m * 32, B, H, T, m, g_seq_trace_bwd};
    if (launch_coop(lstm_big_seq_bwd_kernel, dim3(n, H / 128, 4), SB_THREADS, SB_SMEM, st, 4, tdz, tu, p)) return 1;
    count_launch(1);
  }
  return 0;
}
