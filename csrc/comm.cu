// Peer-memory collectives over NVLink 5 / NVSwitch (K15-K19 of SURVEY.md §2.5).
//
// Every rank maps every peer's symmetric heap (csrc/symm.cpp), so a collective is ONE kernel that
// moves data with plain stores/loads on peer pointers (or `multimem.*` through the switch), fused with
// the 1/N (or 1/window) scale and the bf16 <-> fp32 wire cast, and synchronises with epoch-numbered
// flag words (`st.release.sys` / `ld.acquire.sys`) that live in the heaps:
//
//   one-shot  : push my whole message into a staging slot on every peer, one flag round, reduce locally
//               in rank order (bit-identical result on every rank).  Latency-optimal; default for the
//               200 KB - 4 MB gradients of the reference's models.
//   two-shot  : reduce-scatter by push, reduce my slice, all-gather by push.  2(N-1)/N bytes per rank.
//   nvls      : `multimem.ld_reduce` my slice through the switch + `multimem.st` broadcast.
//   broadcast : root stores through the multicast mapping (or pushes to each peer).
//   allgather : every rank pushes its chunk into every peer's output.
//   adag      : center += scale*(w - w_pulled) via `multimem.red`, then pull (asynchronous, no barrier).
//
// All addresses are (heap base of rank r) + (symmetric offset), so the same kernel also runs in an
// emulation mode on one GPU: blockIdx.y enumerates the emulated ranks inside a single cooperative
// launch (never as separate launches that wait on each other).
// The reference's equivalent is dist-keras' pickle-over-TCP commit/pull.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "shadow.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int MAXR = DDLS_MAX_RANKS;
constexpr int COMM_THREADS = 512;
constexpr int MAX_CHANNELS = DDLS_COMM_MAX_BLOCKS;

struct CommCtx {
  uint8_t* base[MAXR];   // local mappings of every rank's heap
  uint8_t* mc_base;      // multicast mapping of the heaps (nullptr if unavailable)
  int world;
  int rank_base;         // this process' rank (0 in emulation mode, where blockIdx.y adds the rank)
  uint64_t flags_off;    // uint32 [MAX_CHANNELS][MAXR] written by peers
  uint64_t seq_off;      // uint32 [2]: {completed epochs, arrival ticket}
  uint64_t timeout_ns;
  int fence_mode;        // 0: fences chosen per call site (below); 1: every flag round has release + acquire fences
};

// Which system-scope fences a flag round needs.  A fence.acq_rel.sys costs ~3 us on this part, so a round only pays for
// the ones its data flow requires:
//   release  before signalling, when THIS kernel wrote data a peer will read after the round (stores pushed to peers,
//            multimem.st).  Not needed when the data was written by earlier kernels of the stream: kernel completion has
//            already made it visible system-wide.
//   acquire  after the wait, when the data is then read through a path that could hold a stale copy.  The kernels here
//            read peer-written data only through L2 (ld.cg / multimem / first-touch lines of an L1 that was invalidated
//            at kernel start), after the control dependency on the polled flag, so no call site needs it; the stand-alone
//            barrier keeps it.  DDLS_COMM_FENCES=strict restores both fences everywhere.
constexpr int FENCE_NONE = 0, FENCE_REL = 1, FENCE_ACQ = 2, FENCE_BOTH = 3;

DDLS_DEVINL uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Flag barrier among same-index blocks of all ranks.  Called by the whole block; data written before the
// call (by any thread of the block) is visible to peers' readers after their matching wait.
DDLS_DEVINL void peer_barrier(const CommCtx& c, int rank, int channel, uint32_t epoch, int fences) {
  if (c.fence_mode == 1) fences = FENCE_BOTH;
  // One system-scope fence per signalling thread, not one per thread of the block: the bar.sync orders every thread's
  // earlier writes before the fence of the signalling threads, and the fence is cumulative (the pattern of a cooperative
  // grid barrier).  The poll is a relaxed load — an acquire load would pay a fence on every spin — followed by one fence.
  // Measured here: the former form (__threadfence_system by all threads, st.release, ld.acquire polling) cost 10-12 us
  // per flag round with 64-127 blocks in flight.
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int r = threadIdx.x;
    uint32_t* remote = reinterpret_cast<uint32_t*>(c.base[r] + c.flags_off) + channel * MAXR + rank;
    if (fences & FENCE_REL) fence_acq_rel_sys();
    st_relaxed_sys(remote, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(c.base[rank] + c.flags_off) + channel * MAXR + r;
    if (static_cast<int32_t>(ld_relaxed_sys(mine) - epoch) < 0) {
      const uint64_t t0 = globaltimer_ns();
      uint32_t spins = 0;
      while (static_cast<int32_t>(ld_relaxed_sys(mine) - epoch) < 0) {
        if ((++spins & 255u) == 0 && globaltimer_ns() - t0 > c.timeout_ns) {
          printf("ddls comm: rank %d channel %d timed out waiting for rank %d at epoch %u (saw %u)\n", rank, channel, r,
                 epoch, ld_relaxed_sys(mine));
          __trap();
        }
      }
    }
    if (fences & FENCE_ACQ) fence_acq_rel_sys();
  }
  __syncthreads();
}

// Epoch protocol: the per-rank counter is even between collectives; every collective consumes two values
// (e0+1 for its first flag round, e0+2 for its second, if any) and the staging parity (e0 >> 1) & 1 flips on
// every call, so a fast rank pushing for call k+1 never lands in a slot a slow rank still reads for call k.
// All blocks read the epoch counter on entry; the last block to leave advances it (cf. publish_step).
DDLS_DEVINL uint32_t read_epoch(const CommCtx& c, int rank) {
  return *reinterpret_cast<volatile uint32_t*>(c.base[rank] + c.seq_off);
}
DDLS_DEVINL void advance_epoch(const CommCtx& c, int rank, uint32_t new_value) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* seq = reinterpret_cast<uint32_t*>(c.base[rank] + c.seq_off);
    __threadfence();
    const uint32_t ticket = atomicAdd(seq + 1, 1u);
    if (ticket == gridDim.x - 1) { seq[0] = new_value; seq[1] = 0; __threadfence(); }
  }
}

template <typename Wire> struct WireOps;
template <> struct WireOps<float> {
  using vec = float4;                       // 4 elements per 16 B
  static constexpr int EPV = 4;
  DDLS_DEVINL static vec pack(const float* v) { return make_float4(v[0], v[1], v[2], v[3]); }
  DDLS_DEVINL static void unpack(const vec& p, float* v) { v[0] = p.x; v[1] = p.y; v[2] = p.z; v[3] = p.w; }
};
template <> struct WireOps<__nv_bfloat16> {
  using vec = uint2;                        // 4 elements per 8 B
  static constexpr int EPV = 4;
  DDLS_DEVINL static vec pack(const float* v) { return make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3])); }
  DDLS_DEVINL static void unpack(const vec& p, float* v) {
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&p.x), b = *reinterpret_cast<const __nv_bfloat162*>(&p.y);
    v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
  }
};

struct ArParams {
  uint64_t in_off, out_off;   // fp32 [n] in each heap (may alias)
  uint64_t stage_off;         // staging: two parity halves, each [world][n] wire elements
  uint64_t parity_bytes;      // size of one half — FIXED per staging region, not derived from n: a fast rank may already
                              // push call k+1 (other parity, possibly larger n) while a slow one still reads call k
  uint64_t n;                 // elements, multiple of 4
  float scale;
  int mc_push;                // flag-in-data: push each packet with ONE store through the multicast mapping (all peers at once)
};

// Optional optimizer epilogue (K12/K13 "fused into the allreduce epilogue"): the thread that has just summed a
// gradient vector applies the Keras update to the matching parameter / slot vectors, so a data-parallel step of a
// single-bucket model is ONE kernel after backward instead of allreduce + optimizer.
struct OptFuse {
  int kind;                   // 0 none, 1 adam, 2 adagrad, 3 sgd
  float* p; float* s1; float* s2;   // params, m|accumulator, v
  int* step;                  // [completed steps, ticket] as in csrc/elementwise.cu
  float lr, b1, b2, eps;
  int zero_grads;
  uint64_t nvec;              // vectors the optimizer owns; the message may be longer (metric tail: reduced only)
  int advance_step;           // the LAST bucket kernel of a training step publishes step+1 (multi-bucket models)
};

DDLS_DEVINL void opt_apply(const OptFuse& o, uint64_t i, float (&g)[4], float lr_t) {
  float4* P = reinterpret_cast<float4*>(o.p) + i;
  float4 pi = *P;
  float pv[4] = {pi.x, pi.y, pi.z, pi.w};
  if (o.kind == 1) {
    float4* M = reinterpret_cast<float4*>(o.s1) + i;
    float4* V = reinterpret_cast<float4*>(o.s2) + i;
    const float4 m4 = *M, v4 = *V;
    float mv[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mv[j] = o.b1 * mv[j] + (1.0f - o.b1) * g[j];
      vv[j] = o.b2 * vv[j] + (1.0f - o.b2) * g[j] * g[j];
      pv[j] -= lr_t * mv[j] / (sqrtf(vv[j]) + o.eps);
    }
    *M = make_float4(mv[0], mv[1], mv[2], mv[3]);
    *V = make_float4(vv[0], vv[1], vv[2], vv[3]);
  } else if (o.kind == 2) {
    float4* A = reinterpret_cast<float4*>(o.s1) + i;
    const float4 a4 = *A;
    float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      av[j] += g[j] * g[j];
      pv[j] -= o.lr * g[j] / (sqrtf(av[j]) + o.eps);
    }
    *A = make_float4(av[0], av[1], av[2], av[3]);
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) pv[j] -= o.lr * g[j];
  }
  *P = make_float4(pv[0], pv[1], pv[2], pv[3]);
  if (o.zero_grads) { g[0] = g[1] = g[2] = g[3] = 0.f; }
}

// ---------------------------------------------------------------------------------------------------------
// one-shot push allreduce
// ---------------------------------------------------------------------------------------------------------
template <typename Wire>
__global__ void __launch_bounds__(COMM_THREADS) allreduce_oneshot_kernel(CommCtx c, ArParams p) {
  using W = WireOps<Wire>;
  using V = typename W::vec;
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t e0 = read_epoch(c, rank);
  const uint32_t epoch = e0 + 1;
  const uint64_t nvec = p.n / 4;
  const uint64_t per = (nvec + gridDim.x - 1) / gridDim.x;
  const uint64_t lo = blockIdx.x * per, hi = min(nvec, lo + per);
  const float4* in = reinterpret_cast<const float4*>(c.base[rank] + p.in_off);
  const uint64_t slot_vecs = nvec;                       // one slot = n wire elements
  const uint64_t parity_vecs = p.parity_bytes / sizeof(V);
  const uint64_t par = (e0 >> 1) & 1u;
  // push my slice to every peer's staging slot [parity][rank]
#pragma unroll 4
  for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const float4 x = in[i];
    const float xv[4] = {x.x, x.y, x.z, x.w};
    const V w = W::pack(xv);
    for (int d = 1; d < c.world; ++d) {
      const int r = (rank + d) % c.world;                // stagger destinations across ranks
      V* dst = reinterpret_cast<V*>(c.base[r] + p.stage_off) + par * parity_vecs + (uint64_t)rank * slot_vecs + i;
      *dst = w;
    }
  }
  peer_barrier(c, rank, blockIdx.x, epoch, FENCE_REL);
  const V* stage = reinterpret_cast<const V*>(c.base[rank] + p.stage_off) + par * parity_vecs;
  float4* out = reinterpret_cast<float4*>(c.base[rank] + p.out_off);
#pragma unroll 2
  for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float4 mine = in[i];
#pragma unroll 8
    for (int r = 0; r < c.world; ++r) {                   // fixed rank order => identical bits on every rank
      float v[4];
      if (r == rank) {
        const float mv[4] = {mine.x, mine.y, mine.z, mine.w};
        const V w = W::pack(mv);                         // round my own contribution like the wire does
        W::unpack(w, v);
      } else {
        W::unpack(stage[(uint64_t)r * slot_vecs + i], v);
      }
      acc[0] += v[0]; acc[1] += v[1]; acc[2] += v[2]; acc[3] += v[3];
    }
    out[i] = make_float4(acc[0] * p.scale, acc[1] * p.scale, acc[2] * p.scale, acc[3] * p.scale);
  }
  advance_epoch(c, rank, e0 + 2);
}

// ---------------------------------------------------------------------------------------------------------
// one-shot, "flag-in-data" variant for small messages (<= 256 KB by default): every 8-byte packet carries 4 bytes of
// payload and the epoch number, so there is NO separate flag round and NO system-scope fence — a receiver simply
// polls each packet until its flag equals the current epoch (8-byte stores are single-copy atomic over NVLink).
// Costs 2x the wire bytes, removes one NVLink round trip (the release fence) from the critical path.
// Staging: its own zero-initialised region, [2 parities][world][n] packets.
// ---------------------------------------------------------------------------------------------------------
DDLS_DEVINL void st_volatile_v4(uint4* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.volatile.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
DDLS_DEVINL uint4 ld_volatile_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
DDLS_DEVINL uint4 ll_wait(const CommCtx& c, const uint4* p, uint32_t epoch, int rank, int from) {
  uint4 v = ld_volatile_v4(p);
  if (v.y == epoch && v.w == epoch) return v;
  const uint64_t t0 = globaltimer_ns();
  uint32_t spins = 0;
  for (;;) {
    v = ld_volatile_v4(p);
    if (v.y == epoch && v.w == epoch) return v;
    if ((++spins & 1023u) == 0 && globaltimer_ns() - t0 > c.timeout_ns) {
      printf("ddls comm (ll): rank %d timed out waiting for rank %d at epoch %u (saw %u/%u)\n", rank, from, epoch, v.y, v.w);
      __trap();
    }
  }
}

template <typename Wire>
__global__ void __launch_bounds__(COMM_THREADS) allreduce_oneshot_ll_kernel(CommCtx c, ArParams p, OptFuse opt) {
  constexpr bool F32 = sizeof(Wire) == 4;
  pdl_wait();
  int t_step = 0;
  float lr_t = 0.f;
  if (opt.kind != 0) {
    t_step = *reinterpret_cast<volatile int*>(opt.step) + 1;
    lr_t = (opt.kind == 1) ? opt.lr * sqrtf(1.0f - powf(opt.b2, (float)t_step)) / (1.0f - powf(opt.b1, (float)t_step)) : opt.lr;
  }
  constexpr int Q = F32 ? 2 : 1;                          // 16-byte stores per 4 elements
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t e0 = read_epoch(c, rank);
  const uint32_t epoch = e0 + 1;
  const uint64_t nvec = p.n / 4;
  const uint64_t slot_q = nvec * Q, parity_q = p.parity_bytes / 16;
  const uint64_t par = (e0 >> 1) & 1u;
  const float4* in = reinterpret_cast<const float4*>(c.base[rank] + p.in_off);
  float4* out = reinterpret_cast<float4*>(c.base[rank] + p.out_off);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t first = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (uint64_t i = first; i < nvec; i += stride) {
    const float4 x = in[i];
    uint32_t w[4];
    if (F32) { w[0] = __float_as_uint(x.x); w[1] = __float_as_uint(x.y); w[2] = __float_as_uint(x.z); w[3] = __float_as_uint(x.w); }
    else { w[0] = pack_bf16x2(x.x, x.y); w[1] = pack_bf16x2(x.z, x.w); w[2] = 0; w[3] = 0; }
    if (p.mc_push) {
      // a store to the multicast mapping is replicated by the switch into the same offset of every heap: one 16-byte
      // store per packet pair instead of N-1 (my own replica of my slot is written too and never read)
      uint4* dst = reinterpret_cast<uint4*>(c.mc_base + p.stage_off) + par * parity_q + (uint64_t)rank * slot_q + i * Q;
      st_volatile_v4(dst, w[0], epoch, w[1], epoch);
      if (F32) st_volatile_v4(dst + 1, w[2], epoch, w[3], epoch);
    } else {
      for (int d = 1; d < c.world; ++d) {
        const int r = (rank + d) % c.world;
        uint4* dst = reinterpret_cast<uint4*>(c.base[r] + p.stage_off) + par * parity_q + (uint64_t)rank * slot_q + i * Q;
        st_volatile_v4(dst, w[0], epoch, w[1], epoch);
        if (F32) st_volatile_v4(dst + 1, w[2], epoch, w[3], epoch);
      }
    }
  }
  const uint4* stage = reinterpret_cast<const uint4*>(c.base[rank] + p.stage_off) + par * parity_q;
  for (uint64_t i = first; i < nvec; i += stride) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float4 mine = in[i];
    // Issue the loads of EVERY peer's packet(s) before testing any flag: up to 14 independent 16-byte loads are in flight
    // instead of seven dependent poll loops (8 GPUs: 256 KB took 21 us with the serial form).  Packets that have not
    // landed yet are then re-polled one by one.
    uint4 pa[MAXR], pb[MAXR];
#pragma unroll
    for (int r = 0; r < MAXR; ++r) {
      if (r < c.world && r != rank) {
        const uint4* src = stage + (uint64_t)r * slot_q + i * Q;
        pa[r] = ld_volatile_v4(src);
        if (F32) pb[r] = ld_volatile_v4(src + 1);
      }
    }
#pragma unroll
    for (int r = 0; r < MAXR; ++r) {                      // fixed rank order => identical bits on every rank
      if (r >= c.world) break;
      float v[4];
      if (r == rank) {
        if (F32) { v[0] = mine.x; v[1] = mine.y; v[2] = mine.z; v[3] = mine.w; }
        else {
          const uint32_t a = pack_bf16x2(mine.x, mine.y), b = pack_bf16x2(mine.z, mine.w);
          v[0] = __uint_as_float(a << 16); v[1] = __uint_as_float(a & 0xFFFF0000u);
          v[2] = __uint_as_float(b << 16); v[3] = __uint_as_float(b & 0xFFFF0000u);
        }
      } else {
        const uint4* src = stage + (uint64_t)r * slot_q + i * Q;
        uint4 a = pa[r];
        if (a.y != epoch || a.w != epoch) a = ll_wait(c, src, epoch, rank, r);
        if (F32) {
          uint4 b = pb[r];
          if (b.y != epoch || b.w != epoch) b = ll_wait(c, src + 1, epoch, rank, r);
          v[0] = __uint_as_float(a.x); v[1] = __uint_as_float(a.z); v[2] = __uint_as_float(b.x); v[3] = __uint_as_float(b.z);
        } else {
          v[0] = __uint_as_float(a.x << 16); v[1] = __uint_as_float(a.x & 0xFFFF0000u);
          v[2] = __uint_as_float(a.z << 16); v[3] = __uint_as_float(a.z & 0xFFFF0000u);
        }
      }
      acc[0] += v[0]; acc[1] += v[1]; acc[2] += v[2]; acc[3] += v[3];
    }
    float g[4] = {acc[0] * p.scale, acc[1] * p.scale, acc[2] * p.scale, acc[3] * p.scale};
    if (opt.kind != 0 && i < opt.nvec) opt_apply(opt, i, g, lr_t);
    out[i] = make_float4(g[0], g[1], g[2], g[3]);
  }
  // last block out: advance the collective epoch and, with a fused optimizer, its step counter
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* seq = reinterpret_cast<uint32_t*>(c.base[rank] + c.seq_off);
    __threadfence();
    const uint32_t ticket = atomicAdd(seq + 1, 1u);
    if (ticket == gridDim.x - 1) {
      seq[0] = e0 + 2; seq[1] = 0;
      if (opt.kind != 0 && opt.advance_step) opt.step[0] = t_step;
      __threadfence();
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// two-shot: reduce-scatter by push -> reduce own slice -> all-gather by push.  out must be symmetric.
// ---------------------------------------------------------------------------------------------------------
template <typename Wire>
__global__ void __launch_bounds__(COMM_THREADS) allreduce_twoshot_kernel(CommCtx c, ArParams p) {
  using W = WireOps<Wire>;
  using V = typename W::vec;
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t e0 = read_epoch(c, rank);
  const uint32_t epoch = e0 + 1;
  const uint64_t nvec = p.n / 4;
  // block slice, then rank sub-slice inside it
  const uint64_t per_blk = (nvec + gridDim.x - 1) / gridDim.x;
  const uint64_t blo = blockIdx.x * per_blk, bhi = min(nvec, blo + per_blk);
  const uint64_t per_rank = (bhi > blo) ? ((bhi - blo) + c.world - 1) / c.world : 0;
  const float4* in = reinterpret_cast<const float4*>(c.base[rank] + p.in_off);
  const uint64_t par = (e0 >> 1) & 1u;
  const uint64_t slot_vecs = nvec;
  const uint64_t parity_vecs = p.parity_bytes / sizeof(V);
  for (int d = 1; d < c.world; ++d) {
    const int r = (rank + d) % c.world;
    const uint64_t lo = blo + r * per_rank, hi = min(bhi, lo + per_rank);
    V* dst = reinterpret_cast<V*>(c.base[r] + p.stage_off) + par * parity_vecs + (uint64_t)rank * slot_vecs;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      const float4 x = in[i];
      const float xv[4] = {x.x, x.y, x.z, x.w};
      dst[i] = W::pack(xv);
    }
  }
  peer_barrier(c, rank, blockIdx.x, epoch, FENCE_REL);
  {
    const uint64_t lo = blo + rank * per_rank, hi = min(bhi, lo + per_rank);
    const V* stage = reinterpret_cast<const V*>(c.base[rank] + p.stage_off) + par * parity_vecs;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const float4 mine = in[i];
      for (int r = 0; r < c.world; ++r) {
        float v[4];
        if (r == rank) {
          const float mv[4] = {mine.x, mine.y, mine.z, mine.w};
          W::unpack(W::pack(mv), v);
        } else {
          W::unpack(stage[(uint64_t)r * slot_vecs + i], v);
        }
        acc[0] += v[0]; acc[1] += v[1]; acc[2] += v[2]; acc[3] += v[3];
      }
      const float4 res = make_float4(acc[0] * p.scale, acc[1] * p.scale, acc[2] * p.scale, acc[3] * p.scale);
      for (int d = 0; d < c.world; ++d) {
        const int r = (rank + d) % c.world;
        reinterpret_cast<float4*>(c.base[r] + p.out_off)[i] = res;
      }
    }
  }
  peer_barrier(c, rank, blockIdx.x, epoch + 1, FENCE_REL);
  advance_epoch(c, rank, epoch + 1);
}

// ---------------------------------------------------------------------------------------------------------
// NVLS: in-switch reduction + multicast store.  in/out symmetric, bound to the multicast object.
// ---------------------------------------------------------------------------------------------------------
DDLS_DEVINL float4 mc_ld_reduce_nc(const float* mc_ptr);    // defined next to the sharded step kernel below

__global__ void __launch_bounds__(COMM_THREADS) allreduce_nvls_kernel(CommCtx c, ArParams p) {
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t epoch = read_epoch(c, rank) + 1;
  const uint64_t nvec = p.n / 4;
  const uint64_t per_blk = (nvec + gridDim.x - 1) / gridDim.x;
  const uint64_t blo = blockIdx.x * per_blk, bhi = min(nvec, blo + per_blk);
  const uint64_t per_rank = (bhi > blo) ? ((bhi - blo) + c.world - 1) / c.world : 0;
  const uint64_t lo = blo + rank * per_rank, hi = min(bhi, lo + per_rank);
  peer_barrier(c, rank, blockIdx.x, epoch, FENCE_NONE);               // every rank's input is final
  const float* mc_in = reinterpret_cast<const float*>(c.mc_base + p.in_off);
  float* mc_out = reinterpret_cast<float*>(c.mc_base + p.out_off);
  // NVLS_U switch round trips in flight per thread: the loads of a round are all issued before the first store (a
  // load/store pair per iteration leaves ONE load in flight per warp)
  constexpr int NVLS_U = 4;
  for (uint64_t i0 = lo + threadIdx.x; i0 < hi; i0 += (uint64_t)NVLS_U * blockDim.x) {
    float4 v[NVLS_U];
#pragma unroll
    for (int k = 0; k < NVLS_U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i < hi) v[k] = mc_ld_reduce_nc(mc_in + 4 * i);
    }
#pragma unroll
    for (int k = 0; k < NVLS_U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i < hi) {
        v[k].x *= p.scale; v[k].y *= p.scale; v[k].z *= p.scale; v[k].w *= p.scale;
        multimem_st_f32x4(mc_out + 4 * i, v[k]);
      }
    }
  }
  peer_barrier(c, rank, blockIdx.x, epoch + 1, FENCE_REL);
  advance_epoch(c, rank, epoch + 1);
}

// NVLS with bf16 on the wire (Compression.bf16 on buckets above the one-shot range): the fp32 input is cast into a bf16
// staging buffer of the heap, the switch reduces the bf16 replicas with fp32 accumulation
// (`multimem.ld_reduce.add.acc::f32.v4.bf16x2`, 8 elements per 16 bytes), the scaled sum goes back to every replica as
// bf16 by multicast store, and each rank widens it to fp32 locally: half the NVLink bytes of the fp32 form in both
// directions.  Staging: one parity half of the push-collective region, n bf16 elements; n % 8 == 0.
DDLS_DEVINL uint4 mc_ld_reduce_bf16x8(const void* mc_ptr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc_ptr));
  return v;
}
DDLS_DEVINL void mc_st_b128(void* mc_ptr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_ptr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
}
DDLS_DEVINL uint32_t scale_bf16x2(uint32_t w, float s) {
  return pack_bf16x2(__uint_as_float(w << 16) * s, __uint_as_float(w & 0xFFFF0000u) * s);
}

__global__ void __launch_bounds__(COMM_THREADS) allreduce_nvls_bf16_kernel(CommCtx c, ArParams p) {
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t e0 = read_epoch(c, rank);
  const uint64_t nv8 = p.n / 8;                            // 16-byte bf16 vectors
  const uint64_t per_blk = (nv8 + gridDim.x - 1) / gridDim.x;
  const uint64_t blo = blockIdx.x * per_blk, bhi = min(nv8, blo + per_blk);
  const uint64_t per_rank = (bhi > blo) ? ((bhi - blo) + c.world - 1) / c.world : 0;
  const uint64_t lo = blo + rank * per_rank, hi = min(bhi, lo + per_rank);
  const uint64_t par_off = ((e0 >> 1) & 1u) * p.parity_bytes;
  const float4* in = reinterpret_cast<const float4*>(c.base[rank] + p.in_off);
  uint4* st_local = reinterpret_cast<uint4*>(c.base[rank] + p.stage_off + par_off);
  uint8_t* st_mc = c.mc_base + p.stage_off + par_off;
  // cast my whole block range (every rank's slice of it: the owners reduce MY replica too)
  for (uint64_t i = blo + threadIdx.x; i < bhi; i += blockDim.x) {
    const float4 a = in[2 * i], b = in[2 * i + 1];
    st_local[i] = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
  peer_barrier(c, rank, blockIdx.x, e0 + 1, FENCE_REL);
  constexpr int U = 4;
  for (uint64_t i0 = lo + threadIdx.x; i0 < hi; i0 += (uint64_t)U * blockDim.x) {
    uint4 v[U];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i < hi) v[k] = mc_ld_reduce_bf16x8(st_mc + 16 * i);
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i < hi) {
        if (p.scale != 1.0f)
          v[k] = make_uint4(scale_bf16x2(v[k].x, p.scale), scale_bf16x2(v[k].y, p.scale), scale_bf16x2(v[k].z, p.scale),
                            scale_bf16x2(v[k].w, p.scale));
        mc_st_b128(st_mc + 16 * i, v[k]);
      }
    }
  }
  peer_barrier(c, rank, blockIdx.x, e0 + 2, FENCE_REL);
  float4* out = reinterpret_cast<float4*>(c.base[rank] + p.out_off);
  for (uint64_t i = blo + threadIdx.x; i < bhi; i += blockDim.x) {
    const uint4 w = __ldcg(st_local + i);                  // written through NVLink: read from L2
    out[2 * i] = make_float4(__uint_as_float(w.x << 16), __uint_as_float(w.x & 0xFFFF0000u), __uint_as_float(w.y << 16),
                             __uint_as_float(w.y & 0xFFFF0000u));
    out[2 * i + 1] = make_float4(__uint_as_float(w.z << 16), __uint_as_float(w.z & 0xFFFF0000u), __uint_as_float(w.w << 16),
                                 __uint_as_float(w.w & 0xFFFF0000u));
  }
  advance_epoch(c, rank, e0 + 2);
}

// ---------------------------------------------------------------------------------------------------------
// Sharded data-parallel step: reduce-scatter -> Keras optimizer on my 1/N slice -> parameter all-gather, ONE kernel per
// gradient bucket (the training loop of /root/reference/ddl_mnist_aztk.py:216-219 does commit + pull + optimizer as three
// host-side phases).
//
//   barrier A   every rank's gradients of this bucket are final (and nobody reads this bucket's parameters any more)
//   my slice    g = scale * sum_ranks(grad)   `multimem.ld_reduce` through the switch (or N peer loads in rank order)
//               m, v, p update on the slice only: the optimizer pass costs 1/N and its slots are sharded by rank
//               new parameters -> every replica with ONE `multimem.st` (or N peer stores); the reduced gradient itself
//               never crosses NVLink again
//   barrier B   every slice of this block index has landed everywhere
//   local pass  clear my gradient buffer and refresh my bf16 shadow weights for the slices other ranks own
//
// Rank r owns vectors [r*per_rank, (r+1)*per_rank) of the bucket (per_rank = ceil(nvec/world)), independent of the grid,
// so the host knows which part of the optimizer slots is valid on which rank (checkpoints gather them).  Block b of every
// rank works on sub-slice b of its rank's slice; barriers pair same-index blocks.
// ---------------------------------------------------------------------------------------------------------
struct DpStep {
  uint64_t grad_off, param_off;   // symmetric offsets of the bucket's first element (fp32) in every heap
  uint64_t s1_off, s2_off;        // optimizer slots (m | accumulator, v) of the bucket, offsets in MY heap
  uint64_t n, n_opt;              // elements in the bucket / elements the optimizer owns (rest: metric tail, reduced only)
  uint64_t elem0;                 // flat parameter index of the bucket's first element (shadow specs use flat indices)
  int* step;                      // optimizer step counter {steps, ticket}
  float scale, lr, b1, b2, eps;
  int kind;                       // 1 adam, 2 adagrad, 3 sgd
  int zero_grads, advance_step;
  unsigned long long* trace;      // optional: 5 globaltimer stamps per block (entry, barrier A, slice done, barrier B, end)
};

DDLS_DEVINL float4 ldcg4(const float4* p) { return __ldcg(p); }
// multimem.ld_reduce WITHOUT the "memory" clobber of the common.cuh wrapper: several of these must be in flight per
// thread together with the local slot loads (a clobber serialises them; the flag barrier before the first one is the
// ordering point).
DDLS_DEVINL float4 mc_ld_reduce_nc(const float* mc_ptr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc_ptr));
  return v;
}

constexpr int DP_U = 4;          // vectors per thread per round: every remote load of a round is issued before the first use
constexpr int DP_THREADS = 256;  // ~120 registers per thread: a 256-thread CTA leaves room for a compute CTA on the same SM

template <bool MC, int KIND>
__global__ void __launch_bounds__(DP_THREADS, 1) dp_step_kernel(CommCtx c, DpStep p, ShadowSet sh) {
  const int rank = c.rank_base + blockIdx.y;
  pdl_wait();
  const uint32_t e0 = read_epoch(c, rank);
  const int t_step = *reinterpret_cast<volatile int*>(p.step) + 1;
  __shared__ float s_lr_t;
  if (threadIdx.x == 0)
    s_lr_t = (KIND == 1) ? p.lr * sqrtf(1.0f - powf(p.b2, (float)t_step)) / (1.0f - powf(p.b1, (float)t_step)) : p.lr;
  const uint64_t nvec = p.n / 4, nopt = p.n_opt / 4;
  const uint64_t per_rank = (nvec + c.world - 1) / c.world;
  const uint64_t per_blk = (per_rank + gridDim.x - 1) / gridDim.x;
  const uint64_t rlo = min(nvec, (uint64_t)rank * per_rank), rhi = min(nvec, rlo + per_rank);
  const uint64_t lo = min(rhi, rlo + blockIdx.x * per_blk), hi = min(rhi, lo + per_blk);
  float4* P = reinterpret_cast<float4*>(c.base[rank] + p.param_off);
  float4* G = reinterpret_cast<float4*>(c.base[rank] + p.grad_off);
  float4* S1 = reinterpret_cast<float4*>(c.base[rank] + p.s1_off);
  float4* S2 = reinterpret_cast<float4*>(c.base[rank] + p.s2_off);
  unsigned long long* tr = (p.trace != nullptr && threadIdx.x == 0) ? p.trace + 5 * (blockIdx.y * gridDim.x + blockIdx.x) : nullptr;
  if (tr) tr[0] = globaltimer_ns();
  peer_barrier(c, rank, blockIdx.x, e0 + 1, FENCE_NONE);
  if (tr) tr[1] = globaltimer_ns();
  const float lr_t = s_lr_t;
  const float ob1 = 1.0f - p.b1, ob2 = 1.0f - p.b2;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  for (uint64_t i0 = lo + threadIdx.x; i0 < hi; i0 += (uint64_t)DP_U * blockDim.x) {
    float4 g[DP_U], pi[DP_U], s1[DP_U], s2[DP_U];
    // round 1: every remote load of this thread
#pragma unroll
    for (int k = 0; k < DP_U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i < hi) {
        if (MC) {
          g[k] = mc_ld_reduce_nc(reinterpret_cast<const float*>(c.mc_base + p.grad_off) + 4 * i);
        } else {
          g[k] = zero4;
#pragma unroll
          for (int r = 0; r < MAXR; ++r) {                  // fixed rank order
            if (r >= c.world) break;
            const float4 x = ldcg4(reinterpret_cast<const float4*>(c.base[r] + p.grad_off) + i);
            g[k].x += x.x; g[k].y += x.y; g[k].z += x.z; g[k].w += x.w;
          }
        }
      }
    }
    // round 2: the local parameter / slot vectors (independent of round 1: they overlap the switch round trip)
#pragma unroll
    for (int k = 0; k < DP_U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i < hi && i < nopt) {
        pi[k] = P[i];
        if (KIND <= 2) s1[k] = S1[i];
        if (KIND == 1) s2[k] = S2[i];
      }
    }
#pragma unroll
    for (int k = 0; k < DP_U; ++k) {
      const uint64_t i = i0 + (uint64_t)k * blockDim.x;
      if (i >= hi) continue;
      float4 gg = g[k];
      gg.x *= p.scale; gg.y *= p.scale; gg.z *= p.scale; gg.w *= p.scale;
      if (i < nopt) {
        float4 q = pi[k];
        if (KIND == 1) {
          float4 mi = s1[k], vi = s2[k];
          mi.x = p.b1 * mi.x + ob1 * gg.x; mi.y = p.b1 * mi.y + ob1 * gg.y; mi.z = p.b1 * mi.z + ob1 * gg.z; mi.w = p.b1 * mi.w + ob1 * gg.w;
          vi.x = p.b2 * vi.x + ob2 * gg.x * gg.x; vi.y = p.b2 * vi.y + ob2 * gg.y * gg.y;
          vi.z = p.b2 * vi.z + ob2 * gg.z * gg.z; vi.w = p.b2 * vi.w + ob2 * gg.w * gg.w;
          q.x -= lr_t * mi.x / (sqrtf(vi.x) + p.eps); q.y -= lr_t * mi.y / (sqrtf(vi.y) + p.eps);
          q.z -= lr_t * mi.z / (sqrtf(vi.z) + p.eps); q.w -= lr_t * mi.w / (sqrtf(vi.w) + p.eps);
          S1[i] = mi; S2[i] = vi;
        } else if (KIND == 2) {
          float4 ai = s1[k];
          ai.x += gg.x * gg.x; ai.y += gg.y * gg.y; ai.z += gg.z * gg.z; ai.w += gg.w * gg.w;
          q.x -= p.lr * gg.x / (sqrtf(ai.x) + p.eps); q.y -= p.lr * gg.y / (sqrtf(ai.y) + p.eps);
          q.z -= p.lr * gg.z / (sqrtf(ai.z) + p.eps); q.w -= p.lr * gg.w / (sqrtf(ai.w) + p.eps);
          S1[i] = ai;
        } else {
          q.x -= p.lr * gg.x; q.y -= p.lr * gg.y; q.z -= p.lr * gg.z; q.w -= p.lr * gg.w;
        }
        if (MC) {
          multimem_st_f32x4(reinterpret_cast<float*>(c.mc_base + p.param_off) + 4 * i, q);
          if (!p.zero_grads) multimem_st_f32x4(reinterpret_cast<float*>(c.mc_base + p.grad_off) + 4 * i, gg);
        } else {
#pragma unroll
          for (int d = 0; d < MAXR; ++d) {
            if (d >= c.world) break;
            const int r = (rank + d) % c.world;
            reinterpret_cast<float4*>(c.base[r] + p.param_off)[i] = q;
            if (!p.zero_grads) reinterpret_cast<float4*>(c.base[r] + p.grad_off)[i] = gg;
          }
        }
        if (p.zero_grads) G[i] = zero4;           // my replica of MY slice: nobody else reads it
        const float pv[4] = {q.x, q.y, q.z, q.w};
        write_shadows(sh, p.elem0 + 4 * i, pv);
      } else {                                    // metric tail: the reduced value goes to every replica, nothing is cleared
        if (MC) {
          multimem_st_f32x4(reinterpret_cast<float*>(c.mc_base + p.grad_off) + 4 * i, gg);
        } else {
          for (int r = 0; r < c.world; ++r) reinterpret_cast<float4*>(c.base[r] + p.grad_off)[i] = gg;
        }
      }
    }
  }
  if (p.trace != nullptr) __syncthreads();
  if (tr) tr[2] = globaltimer_ns();
  peer_barrier(c, rank, blockIdx.x, e0 + 2, FENCE_REL);
  if (tr) tr[3] = globaltimer_ns();
  // local pass over the sub-slices the OTHER ranks' same-index blocks have just finished
  if (p.zero_grads || sh.n != 0) {
    for (int d = 1; d < c.world; ++d) {
      const int r = (rank + d) % c.world;
      const uint64_t qlo = min(nvec, (uint64_t)r * per_rank), qhi = min(nvec, qlo + per_rank);
      const uint64_t l2 = min(qhi, qlo + blockIdx.x * per_blk), h2 = min(min(qhi, l2 + per_blk), nopt);
      const bool shadowed = sh.n != 0 && p.elem0 + 4 * h2 > sh.lo && p.elem0 + 4 * l2 < sh.hi;
      for (uint64_t i0 = l2 + threadIdx.x; i0 < h2; i0 += (uint64_t)DP_U * blockDim.x) {
        float4 q[DP_U];
#pragma unroll
        for (int k = 0; k < DP_U; ++k) {
          const uint64_t i = i0 + (uint64_t)k * blockDim.x;
          if (i < h2) {
            if (p.zero_grads) G[i] = zero4;
            if (shadowed) q[k] = ldcg4(P + i);    // written by a peer through NVLink: read it from L2, not a stale L1 line
          }
        }
        if (shadowed) {
#pragma unroll
          for (int k = 0; k < DP_U; ++k) {
            const uint64_t i = i0 + (uint64_t)k * blockDim.x;
            if (i < h2) {
              const float pv[4] = {q[k].x, q[k].y, q[k].z, q[k].w};
              write_shadows(sh, p.elem0 + 4 * i, pv);
            }
          }
        }
      }
    }
  }
  __syncthreads();
  if (tr) tr[4] = globaltimer_ns();
  if (threadIdx.x == 0) {
    uint32_t* seq = reinterpret_cast<uint32_t*>(c.base[rank] + c.seq_off);
    __threadfence();
    const uint32_t ticket = atomicAdd(seq + 1, 1u);
    if (ticket == gridDim.x - 1) {
      seq[0] = e0 + 2; seq[1] = 0;
      if (p.advance_step) p.step[0] = t_step;
      __threadfence();
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// broadcast: root's buffer -> every rank's buffer at the same offset (16-byte vectors)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(COMM_THREADS) broadcast_kernel(CommCtx c, uint64_t off, uint64_t nvec16, int root,
                                                                 int use_mc) {
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t epoch = read_epoch(c, rank) + 1;
  const uint64_t per = (nvec16 + gridDim.x - 1) / gridDim.x;
  const uint64_t lo = blockIdx.x * per, hi = min(nvec16, lo + per);
  peer_barrier(c, rank, blockIdx.x, epoch, FENCE_NONE);               // receivers are done with the old contents
  if (rank == root) {
    const float4* src = reinterpret_cast<const float4*>(c.base[rank] + off);
    if (use_mc && c.mc_base != nullptr) {
      float* dst = reinterpret_cast<float*>(c.mc_base + off);
      for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) multimem_st_f32x4(dst + 4 * i, src[i]);
    } else {
      for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const float4 v = src[i];
        for (int d = 1; d < c.world; ++d)
          reinterpret_cast<float4*>(c.base[(rank + d) % c.world] + off)[i] = v;
      }
    }
  }
  peer_barrier(c, rank, blockIdx.x, epoch + 1, FENCE_REL);
  advance_epoch(c, rank, epoch + 1);
}

// allgather: out[r*chunk : (r+1)*chunk] on every rank = rank r's in (16-byte vectors; out symmetric)
__global__ void __launch_bounds__(COMM_THREADS) allgather_kernel(CommCtx c, uint64_t in_off, uint64_t out_off,
                                                                 uint64_t chunk_vec16) {
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t epoch = read_epoch(c, rank) + 1;
  const uint64_t per = (chunk_vec16 + gridDim.x - 1) / gridDim.x;
  const uint64_t lo = blockIdx.x * per, hi = min(chunk_vec16, lo + per);
  const float4* src = reinterpret_cast<const float4*>(c.base[rank] + in_off);
  peer_barrier(c, rank, blockIdx.x, epoch, FENCE_NONE);
  for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const float4 v = src[i];
    for (int d = 0; d < c.world; ++d) {
      const int r = (rank + d) % c.world;
      reinterpret_cast<float4*>(c.base[r] + out_off)[(uint64_t)rank * chunk_vec16 + i] = v;
    }
  }
  peer_barrier(c, rank, blockIdx.x, epoch + 1, FENCE_REL);
  advance_epoch(c, rank, epoch + 1);
}

__global__ void __launch_bounds__(32) barrier_kernel(CommCtx c) {
  const int rank = c.rank_base + blockIdx.y;
  const uint32_t epoch = read_epoch(c, rank) + 1;
  peer_barrier(c, rank, 0, epoch, FENCE_BOTH);
  advance_epoch(c, rank, epoch + 1);
}

// ---------------------------------------------------------------------------------------------------------
// ADAG / DynSGD commit + pull against a center variable replicated in every heap (K19).
//   delta = (w - pulled) * scale ;  center (all replicas) += delta  via multimem.red (or peer atomics)
//   then w = pulled = local center replica.  No barrier: ranks run asynchronously, like the original
//   parameter server.  `updates_off` is a replicated uint32 commit counter.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(COMM_THREADS) adag_commit_kernel(CommCtx c, float* __restrict__ w,
                                                                   float* __restrict__ pulled, uint64_t center_off,
                                                                   uint64_t updates_off, uint64_t n, float scale,
                                                                   int staleness_aware, uint32_t* last_update) {
  const int rank = c.rank_base + blockIdx.y;
  const uint64_t nvec = n / 4;
  float s = scale;
  if (staleness_aware) {
    const uint32_t now = ld_relaxed_sys(reinterpret_cast<const uint32_t*>(c.base[rank] + updates_off));
    s = scale / (float)(now - *last_update + 1u);
  }
  float4* w4 = reinterpret_cast<float4*>(w);
  float4* p4 = reinterpret_cast<float4*>(pulled);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nvec; i += (uint64_t)gridDim.x * blockDim.x) {
    const float4 a = w4[i], b = p4[i];
    const float4 d = make_float4((a.x - b.x) * s, (a.y - b.y) * s, (a.z - b.z) * s, (a.w - b.w) * s);
    if (c.mc_base != nullptr) {
      multimem_red_add_f32x4(reinterpret_cast<float*>(c.mc_base + center_off) + 4 * i, d);
    } else {
      for (int r = 0; r < c.world; ++r) {
        float* dst = reinterpret_cast<float*>(c.base[r] + center_off) + 4 * i;
        atomicAdd(dst + 0, d.x); atomicAdd(dst + 1, d.y); atomicAdd(dst + 2, d.z); atomicAdd(dst + 3, d.w);
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (c.mc_base != nullptr) multimem_red_add_u32(reinterpret_cast<uint32_t*>(c.mc_base + updates_off), 1u);
    else for (int r = 0; r < c.world; ++r) atomicAdd(reinterpret_cast<uint32_t*>(c.base[r] + updates_off), 1u);
  }
}

__global__ void __launch_bounds__(COMM_THREADS) adag_pull_kernel(CommCtx c, float* __restrict__ w,
                                                                 float* __restrict__ pulled, uint64_t center_off,
                                                                 uint64_t updates_off, uint64_t n, uint32_t* last_update) {
  const int rank = c.rank_base + blockIdx.y;
  const float4* cen = reinterpret_cast<const float4*>(c.base[rank] + center_off);
  float4* w4 = reinterpret_cast<float4*>(w);
  float4* p4 = reinterpret_cast<float4*>(pulled);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n / 4; i += (uint64_t)gridDim.x * blockDim.x) {
    const float4 v = cen[i];
    w4[i] = v; p4[i] = v;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    *last_update = ld_relaxed_sys(reinterpret_cast<const uint32_t*>(c.base[rank] + updates_off));
}

// A kernel that uses (almost) no shared memory leaves its SM configured with the smallest shared-memory carve-out, and a
// CTA that needs ~200 KB (the TMA convolution / GEMM kernels these collectives overlap with) cannot join that SM until
// every resident CTA has left.  Asking for the largest carve-out keeps the SM in the configuration the compute kernels need.
template <typename K>
static void prefer_max_smem(K kern) {
  static thread_local const void* done[48];
  static thread_local int ndone = 0;
  for (int i = 0; i < ndone; ++i) if (done[i] == (const void*)kern) return;
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (ndone < 48) done[ndone++] = (const void*)kern;
}

static int make_ctx(const DdlsCommDesc* d, CommCtx* c) {
  if (d->world < 1 || d->world > MAXR) { set_last_error("comm: world must be in [1,%d]", MAXR); return 1; }
  for (int r = 0; r < MAXR; ++r) c->base[r] = r < d->world ? static_cast<uint8_t*>(d->base[r]) : nullptr;
  c->mc_base = static_cast<uint8_t*>(d->mc_base);
  c->world = d->world;
  c->rank_base = d->emulate ? 0 : d->rank;
  c->flags_off = d->flags_off;
  c->seq_off = d->seq_off;
  c->timeout_ns = d->timeout_ms ? (uint64_t)d->timeout_ms * 1000000ull : 20000000000ull;
  static const int fence_mode = [] { const char* e = getenv("DDLS_COMM_FENCES"); return (e && e[0] == 's') ? 1 : 0; }();
  c->fence_mode = fence_mode;
  return 0;
}

static bool is_dp_step_kernel(const void* k) {
  return k == (const void*)dp_step_kernel<true, 1> || k == (const void*)dp_step_kernel<true, 2> || k == (const void*)dp_step_kernel<true, 3> ||
         k == (const void*)dp_step_kernel<false, 1> || k == (const void*)dp_step_kernel<false, 2> || k == (const void*)dp_step_kernel<false, 3>;
}

template <typename... Args>
static int launch(const DdlsCommDesc* d, void (*kern)(Args...), int blocks, int threads, cudaStream_t st, Args... args) {
  if (d->emulate && blocks * d->world > 128) blocks = 128 / d->world;   // all emulated blocks must be co-resident
  dim3 grid(blocks, d->emulate ? d->world : 1);
  prefer_max_smem(kern);
  if (d->emulate) {
    // emulated ranks spin on each other: all blocks must be co-resident => cooperative launch
    void* ptrs[] = {(void*)&args...};
    DDLS_CUDA_OK(cudaLaunchCooperativeKernel((const void*)kern, grid, dim3(threads), ptrs, 0, st));
  } else {
    // only the kernels that call pdl_wait() may be launched with the programmatic-serialization attribute
    if ((const void*)kern == (const void*)allreduce_oneshot_ll_kernel<float> ||
        (const void*)kern == (const void*)allreduce_oneshot_ll_kernel<__nv_bfloat16> || is_dp_step_kernel((const void*)kern)) {
      DDLS_CUDA_OK(launch_k(kern, grid, dim3(threads), 0, st, args...));
    } else {
      kern<<<grid, threads, 0, st>>>(args...);
    }
  }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "comm kernel launch");
}

static int pick_blocks(uint64_t n, int requested) {
  if (requested > 0) return requested > MAX_CHANNELS ? MAX_CHANNELS : requested;
  const uint64_t bytes = n * 4;
  int b = (int)((bytes + (32 << 10) - 1) / (32 << 10));     // ~32 KB of payload per block (4 x 16 B per thread)
  if (b < 1) b = 1;
  if (b > 64) b = 64;
  return b;
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_comm_allreduce(const DdlsCommDesc* d, int algo, uint64_t in_off, uint64_t out_off,
                                   uint64_t stage_off, uint64_t stage_bytes, uint64_t n, float scale, int wire_flags,
                                   int blocks, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  if (n == 0) return 0;
  const int wire_bf16 = wire_flags & DDLS_WIRE_BF16;
  const int mc_push = (wire_flags & DDLS_LL_MC_PUSH) && c.mc_base != nullptr && !d->emulate;
  if (n % 4 != 0 || (in_off | out_off | stage_off) % 16 != 0) {
    set_last_error("ddls_comm_allreduce: n must be a multiple of 4 and offsets 16-byte aligned");
    return 1;
  }
  const uint64_t parity_bytes = (stage_bytes / 2) & ~static_cast<uint64_t>(15);
  if (algo != DDLS_AR_NVLS) {
    const uint64_t per_elem = (algo == DDLS_AR_ONESHOT_LL) ? (wire_bf16 ? 4 : 8) : (wire_bf16 ? 2 : 4);
    if ((uint64_t)d->world * n * per_elem > parity_bytes) {
      set_last_error("ddls_comm_allreduce: staging region of %llu bytes is too small for %llu elements x %d ranks",
                     (unsigned long long)stage_bytes, (unsigned long long)n, d->world);
      return 1;
    }
  }
  ArParams p{in_off, out_off, stage_off, parity_bytes, n, scale, mc_push};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = pick_blocks(n, blocks);
  switch (algo) {
    case DDLS_AR_ONESHOT:
      return wire_bf16 ? launch(d, allreduce_oneshot_kernel<__nv_bfloat16>, nb, COMM_THREADS, st, c, p)
                       : launch(d, allreduce_oneshot_kernel<float>, nb, COMM_THREADS, st, c, p);
    case DDLS_AR_ONESHOT_LL: {
      int lb = blocks > 0 ? blocks : (int)((n / 4 + COMM_THREADS - 1) / COMM_THREADS);     // one 16-byte vector per thread
      if (lb > 64) lb = 64;
      if (lb < 1) lb = 1;
      OptFuse none{};
      return wire_bf16 ? launch(d, allreduce_oneshot_ll_kernel<__nv_bfloat16>, lb, COMM_THREADS, st, c, p, none)
                       : launch(d, allreduce_oneshot_ll_kernel<float>, lb, COMM_THREADS, st, c, p, none);
    }
    case DDLS_AR_TWOSHOT:
      return wire_bf16 ? launch(d, allreduce_twoshot_kernel<__nv_bfloat16>, nb, COMM_THREADS, st, c, p)
                       : launch(d, allreduce_twoshot_kernel<float>, nb, COMM_THREADS, st, c, p);
    case DDLS_AR_NVLS:
      if (c.mc_base == nullptr || d->emulate) { set_last_error("ddls_comm_allreduce: NVLS needs a multicast mapping"); return 1; }
      if (wire_bf16) {
        if (n % 8 != 0) { set_last_error("ddls_comm_allreduce: NVLS with a bf16 wire needs n %% 8 == 0 (got %llu)", (unsigned long long)n); return 1; }
        if (n * 2 > parity_bytes) {
          set_last_error("ddls_comm_allreduce: NVLS with a bf16 wire needs %llu bytes of staging per parity half (got %llu)",
                         (unsigned long long)(n * 2), (unsigned long long)parity_bytes);
          return 1;
        }
        return launch(d, allreduce_nvls_bf16_kernel, nb, COMM_THREADS, st, c, p);
      }
      return launch(d, allreduce_nvls_kernel, nb, COMM_THREADS, st, c, p);
  }
  set_last_error("ddls_comm_allreduce: unknown algorithm %d", algo);
  return 1;
}

extern "C" int ddls_comm_allreduce_opt(const DdlsCommDesc* d, uint64_t grad_off, uint64_t stage_off, uint64_t stage_bytes,
                                       uint64_t n, uint64_t n_opt, float scale, int wire_flags, int kind, float* params, float* slot1,
                                       float* slot2, int* step, float lr, float b1, float b2, float eps, int zero_grads,
                                       int advance_step, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  if (n == 0) return 0;
  const int wire_bf16 = wire_flags & DDLS_WIRE_BF16;
  const int mc_push = (wire_flags & DDLS_LL_MC_PUSH) && c.mc_base != nullptr && !d->emulate;
  if (d->emulate) { set_last_error("ddls_comm_allreduce_opt: not available in emulation mode (one parameter set per rank)"); return 1; }
  if (kind < 1 || kind > 3 || params == nullptr || step == nullptr || (kind <= 2 && slot1 == nullptr) || (kind == 1 && slot2 == nullptr)) {
    set_last_error("ddls_comm_allreduce_opt: bad optimizer description (kind %d)", kind);
    return 1;
  }
  if (n % 4 != 0 || n_opt % 4 != 0 || n_opt > n || (grad_off | stage_off) % 16 != 0 || (reinterpret_cast<uintptr_t>(params) & 15)) {
    set_last_error("ddls_comm_allreduce_opt: n and n_opt (<= n) must be multiples of 4 and buffers 16-byte aligned");
    return 1;
  }
  const uint64_t parity_bytes = (stage_bytes / 2) & ~static_cast<uint64_t>(15);
  if ((uint64_t)d->world * n * (wire_bf16 ? 4 : 8) > parity_bytes) {
    set_last_error("ddls_comm_allreduce_opt: %llu elements do not fit the flag-in-data staging region", (unsigned long long)n);
    return 1;
  }
  ArParams p{grad_off, grad_off, stage_off, parity_bytes, n, scale, mc_push};
  OptFuse o{kind, params, slot1, slot2, step, lr, b1, b2, eps, zero_grads, n_opt / 4, advance_step};
  int lb = (int)((n / 4 + COMM_THREADS - 1) / COMM_THREADS);
  if (lb > 64) lb = 64;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return wire_bf16 ? launch(d, allreduce_oneshot_ll_kernel<__nv_bfloat16>, lb, COMM_THREADS, st, c, p, o)
                   : launch(d, allreduce_oneshot_ll_kernel<float>, lb, COMM_THREADS, st, c, p, o);
}

// Debug aid (tools/dp_step_trace.py): a device buffer of 5 uint64 per block that later ddls_comm_dp_step launches of this
// process fill with globaltimer stamps; NULL switches it off.
static unsigned long long* g_dp_trace = nullptr;
extern "C" int ddls_comm_dp_step_trace(unsigned long long* buf) { g_dp_trace = buf; return 0; }

extern "C" int ddls_comm_dp_step(const DdlsCommDesc* d, uint64_t grad_off, uint64_t param_off, uint64_t slot1_off,
                                 uint64_t slot2_off, uint64_t n, uint64_t n_opt, uint64_t elem0, float scale, int kind,
                                 int* step, float lr, float b1, float b2, float eps, int zero_grads, int advance_step,
                                 const DdlsShadowSpec* shadows, int n_shadows, int blocks, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  if (n == 0) return 0;
  if (kind < 1 || kind > 3 || step == nullptr) { set_last_error("ddls_comm_dp_step: bad optimizer description (kind %d)", kind); return 1; }
  if (n % 4 != 0 || n_opt % 4 != 0 || n_opt > n || elem0 % 4 != 0 ||
      (grad_off | param_off | slot1_off | slot2_off) % 16 != 0) {
    set_last_error("ddls_comm_dp_step: n, n_opt (<= n), elem0 must be multiples of 4 and offsets 16-byte aligned");
    return 1;
  }
  DpStep p{grad_off, param_off, slot1_off, slot2_off, n, n_opt, elem0, step, scale, lr, b1, b2, eps, kind, zero_grads, advance_step,
           g_dp_trace};
  const ShadowSet sh = make_shadows(shadows, n_shadows);
  // the kernel is latency-bound (two flag rounds + one switch round trip): enough blocks that a thread's share of the slice
  // is ONE round of DP_U vectors, all in flight together - and enough that the local pass after barrier B (clearing the
  // gradients and refreshing the bf16 shadows of the (world-1)/world of the bucket the OTHER ranks own: 15.6 us of a
  // 39.6 us call at 8 GPUs with blocks sized by the slice alone) is spread as widely; `blocks` > 0 caps the SMs taken
  // from the backward kernels this overlaps with
  const uint64_t per_block = (uint64_t)DP_THREADS * DP_U * 16 * 2;
  int nb = (int)((n * 4 + per_block - 1) / per_block);
  if (nb < 1) nb = 1;
  if (nb > MAX_CHANNELS) nb = MAX_CHANNELS;
  if (blocks > 0 && blocks < nb) nb = blocks;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool mc = c.mc_base != nullptr && !d->emulate;
#define DDLS_DP_LAUNCH(M, K) launch(d, dp_step_kernel<M, K>, nb, DP_THREADS, st, c, p, sh)
  if (mc) return kind == 1 ? DDLS_DP_LAUNCH(true, 1) : kind == 2 ? DDLS_DP_LAUNCH(true, 2) : DDLS_DP_LAUNCH(true, 3);
  return kind == 1 ? DDLS_DP_LAUNCH(false, 1) : kind == 2 ? DDLS_DP_LAUNCH(false, 2) : DDLS_DP_LAUNCH(false, 3);
#undef DDLS_DP_LAUNCH
}

extern "C" int ddls_comm_broadcast(const DdlsCommDesc* d, uint64_t off, uint64_t bytes, int root, int use_multicast,
                                   int blocks, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  if (bytes == 0) return 0;
  if (bytes % 16 != 0 || off % 16 != 0) { set_last_error("ddls_comm_broadcast: 16-byte granularity required"); return 1; }
  return launch(d, broadcast_kernel, pick_blocks(bytes / 4, blocks), COMM_THREADS, static_cast<cudaStream_t>(stream), c, off,
                bytes / 16, root, use_multicast);
}

extern "C" int ddls_comm_allgather(const DdlsCommDesc* d, uint64_t in_off, uint64_t out_off, uint64_t chunk_bytes,
                                   int blocks, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  if (chunk_bytes == 0) return 0;
  if (chunk_bytes % 16 != 0 || (in_off | out_off) % 16 != 0) { set_last_error("ddls_comm_allgather: 16-byte granularity required"); return 1; }
  return launch(d, allgather_kernel, pick_blocks(chunk_bytes / 4, blocks), COMM_THREADS, static_cast<cudaStream_t>(stream),
                c, in_off, out_off, chunk_bytes / 16);
}

extern "C" int ddls_comm_barrier(const DdlsCommDesc* d, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  return launch(d, barrier_kernel, 1, 32, static_cast<cudaStream_t>(stream), c);
}

extern "C" int ddls_comm_adag_commit(const DdlsCommDesc* d, float* w, float* pulled, uint64_t center_off,
                                     uint64_t updates_off, uint64_t n, float scale, int staleness_aware,
                                     unsigned* last_update, void* stream) {
  CommCtx c;
  if (make_ctx(d, &c)) return 1;
  if (n % 4 != 0) { set_last_error("ddls_comm_adag_commit: n must be a multiple of 4"); return 1; }
  int nb = (int)((n / 4 + COMM_THREADS - 1) / COMM_THREADS);
  if (nb > 148) nb = 148;
  if (nb < 1) nb = 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid(nb, 1);
  adag_commit_kernel<<<grid, COMM_THREADS, 0, st>>>(c, w, pulled, center_off, updates_off, n, scale, staleness_aware,
                                                     last_update);
  adag_pull_kernel<<<grid, COMM_THREADS, 0, st>>>(c, w, pulled, center_off, updates_off, n, last_update);
  count_launch(2);
  return check_cuda(cudaGetLastError(), "adag kernels launch");
}
