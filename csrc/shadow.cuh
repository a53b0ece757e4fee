// bf16 shadow copies of parameter matrices, shared by the optimizer kernels (csrc/elementwise.cu) and the sharded
// data-parallel step kernel (csrc/comm.cu).
#pragma once
#include "common.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

// bf16 "shadow" copies of selected parameter matrices, refreshed by the same kernel that updates the fp32
// masters (the fused cast of K12): tensor-core kernels then read 16-byte-aligned bf16 rows (ld padded to a
// multiple of 8) instead of converting fp32 on every use.
struct ShadowSet {
  DdlsShadowSpec s[DDLS_MAX_SHADOWS];
  int n;
  unsigned long long lo, hi;   // union of the shadowed element ranges (cheap early-out)
};

DDLS_DEVINL void write_shadows(const ShadowSet& sh, size_t e0, const float (&v)[4]) {
  if (sh.n == 0 || e0 + 4 <= sh.lo || e0 >= sh.hi) return;
#pragma unroll 1
  for (int k = 0; k < sh.n; ++k) {
    const DdlsShadowSpec& sp = sh.s[k];
    const size_t end = sp.begin + (size_t)sp.rows * sp.cols;
    if (e0 + 4 <= sp.begin || e0 >= end) continue;
    __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(sp.dst);
    if (e0 >= sp.begin && e0 + 4 <= end && end - sp.begin < 0xFFFFFFFFull) {
      // common case: all four elements inside the matrix — one 32-bit division, then walk (row, col)
      const uint32_t rel = (uint32_t)(e0 - sp.begin), cols = (uint32_t)sp.cols;
      uint32_t r = rel / cols, c = rel - r * cols;
      if (c + 4 <= cols && ((r * (uint32_t)sp.ld + c) & 1u) == 0) {          // same row, 4-byte aligned: two bf16x2 stores
        uint32_t* d2 = reinterpret_cast<uint32_t*>(dst + (size_t)r * sp.ld + c);
        d2[0] = pack_bf16x2(v[0], v[1]);
        d2[1] = pack_bf16x2(v[2], v[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          dst[(size_t)r * sp.ld + c] = __float2bfloat16(v[j]);
          if (++c == cols) { c = 0; ++r; }
        }
      }
      continue;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const size_t e = e0 + j;
      if (e >= sp.begin && e < end) {
        const size_t rel = e - sp.begin;
        const uint32_t r = (uint32_t)(rel / sp.cols), c = (uint32_t)(rel - (size_t)r * sp.cols);
        dst[(size_t)r * sp.ld + c] = __float2bfloat16(v[j]);
      }
    }
  }
}

static inline ShadowSet make_shadows(const DdlsShadowSpec* specs, int n) {
  ShadowSet sh;
  sh.n = 0; sh.lo = ~0ull; sh.hi = 0;
  for (int k = 0; specs != nullptr && k < n && k < DDLS_MAX_SHADOWS; ++k) {
    sh.s[sh.n++] = specs[k];
    const unsigned long long b = specs[k].begin, e = b + (unsigned long long)specs[k].rows * specs[k].cols;
    if (b < sh.lo) sh.lo = b;
    if (e > sh.hi) sh.hi = e;
  }
  return sh;
}

}  // namespace ddls
