// Warp-specialised tcgen05 mainloop fed by tensor-map TMA with 128-byte swizzle.
//
// The cp.async mainloop of gemm_pipe.cuh spends issue slots of every thread on address math and is limited by how
// many 16-byte requests the CTA keeps in flight; for the large GEMMs of the scaled stacked LSTM (BASELINE.json
// config 4) that left the tensor pipe at ~13% of peak.  Here ONE thread issues box loads (`cp.async.bulk.tensor`),
// ONE thread issues the MMAs, and four warps drain finished accumulators from TMEM while the next tile's MMAs run
// (two accumulator buffers), so nothing on the critical path executes per-element instructions.
//
// Shared-memory operand images (one 64-wide K block per stage, bf16):
//   K-major  [rows][64 k] : one box {64 k, rows}; row r at r*128 B, 16-byte chunks XOR-swizzled with r%8.
//                           UMMA descriptor: SWIZZLE_128B, SBO = 1024 B (8 rows), K16 step = +32 B.
//   MN-major [64 k][rows] : rows/64 boxes {64 mn, 64 k}, each 8 KB (k-row at k*128 B, swizzled with k%8).
//                           UMMA descriptor: SWIZZLE_128B, LBO = 8192 B (next 64 mn), SBO = 1024 B (8 k-rows),
//                           K16 step = +2048 B.
// Tensor maps are always rank 3 (the third coordinate selects a time step / batch slab; extent 1 for plain 2-D).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace ddls {

// host side (csrc/tma.cpp): bf16 tensor map, SWIZZLE_128B, zero fill out of bounds.  dims[0] is the contiguous one;
// strides_bytes[i] is the byte stride of dims[i+1].
int tmap_bf16_sw128(CUtensorMap* out, const void* base, const uint64_t dims[3], const uint64_t strides_bytes[2],
                    const uint32_t box[3]);

int tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, int swizzle);

DDLS_DEVINL void tma_load_4d(uint32_t smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

DDLS_DEVINL void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
DDLS_DEVINL void tma_load_3d(uint32_t smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

constexpr uint32_t SW64_DESC_HI = ((512u >> 4) & 0x3FFFu) | (1u << 14) | (4u << 29);     // SBO 512, version 1, SWIZZLE_64B
constexpr uint32_t SW128_DESC_HI = ((1024u >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);   // SBO 1024, version 1, SWIZZLE_128B

template <int BN_, bool A_MN_, bool B_MN_, int STAGES_>
struct TmaPipe {
  static constexpr int BM = 128, BN = BN_, BK = 64, STAGES = STAGES_;
  static constexpr bool A_MN = A_MN_, B_MN = B_MN_;
  static constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;     // + alignment slack
  static constexpr uint32_t IDESC = umma_idesc_bf16(BM, BN, A_MN, B_MN);
  static_assert(BN % 64 == 0 && BN <= 256, "BN must be 64, 128, 192 or 256");

  // --- producer (one thread) ---
  // A operand: K-major tensor map dims {K, M, Z}; MN-major dims {M, K, Z}.  Likewise B with N.
  static DDLS_DEVINL void load_a(uint8_t* stage, uint64_t* full, const CUtensorMap* ta, int m0, int k0, int z) {
    const uint32_t a = smem_u32(stage);
    if constexpr (A_MN) {
#pragma unroll
      for (int j = 0; j < BM / 64; ++j) tma_load_3d(a + j * 8192, ta, full, m0 + j * 64, k0, z);
    } else {
      tma_load_3d(a, ta, full, k0, m0, z);
    }
  }
  static DDLS_DEVINL void load_b(uint8_t* stage, uint64_t* full, const CUtensorMap* tb, int n0, int k0, int z) {
    const uint32_t b = smem_u32(stage) + A_BYTES;
    if constexpr (B_MN) {
#pragma unroll
      for (int j = 0; j < BN / 64; ++j) tma_load_3d(b + j * 8192, tb, full, n0 + j * 64, k0, z);
    } else {
      tma_load_3d(b, tb, full, k0, n0, z);
    }
  }

  // Separate A / B rings (the operands of one K-block need not sit next to each other).
  static DDLS_DEVINL void load_a_at(uint32_t a, uint64_t* full, const CUtensorMap* ta, int m0, int k0, int z) {
    if constexpr (A_MN) {
#pragma unroll
      for (int j = 0; j < BM / 64; ++j) tma_load_3d(a + j * 8192, ta, full, m0 + j * 64, k0, z);
    } else {
      tma_load_3d(a, ta, full, k0, m0, z);
    }
  }
  static DDLS_DEVINL void load_b_at(uint32_t b, uint64_t* full, const CUtensorMap* tb, int n0, int k0, int z) {
    if constexpr (B_MN) {
#pragma unroll
      for (int j = 0; j < BN / 64; ++j) tma_load_3d(b + j * 8192, tb, full, n0 + j * 64, k0, z);
    } else {
      tma_load_3d(b, tb, full, k0, n0, z);
    }
  }

  // --- MMA issuer (one thread): the four K=16 MMAs of one stage ---
  static DDLS_DEVINL void mma_stage(const uint8_t* stage, uint32_t tmem_d, bool first_block) {
    mma_at(smem_u32(stage), smem_u32(stage) + A_BYTES, tmem_d, first_block);
  }
  static DDLS_DEVINL void mma_at(uint32_t a, uint32_t b, uint32_t tmem_d, bool first_block) {
#pragma unroll
    for (int k16 = 0; k16 < BK / 16; ++k16) {
      const uint32_t a_lo = A_MN ? umma_desc_lo(a + k16 * 2048, 8192) : umma_desc_lo(a + k16 * 32, 16);
      const uint32_t b_lo = B_MN ? umma_desc_lo(b + k16 * 2048, 8192) : umma_desc_lo(b + k16 * 32, 16);
      umma_bf16_lohi(tmem_d, a_lo, SW128_DESC_HI, b_lo, SW128_DESC_HI, IDESC, !(first_block && k16 == 0));
    }
  }
};

DDLS_DEVINL uint8_t* align1024(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~static_cast<uintptr_t>(1023));
}

}  // namespace ddls
