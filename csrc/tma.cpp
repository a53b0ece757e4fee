// Host side of the tensor-map TMA paths: cuTensorMapEncodeTiled resolved at run time (no link against libcuda,
// so the library also loads on the GPU-less build box).
#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>

namespace ddls {
void set_last_error(const char* fmt, ...);

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeFn g_encode = nullptr;
static std::once_flag g_once;

int tmap_bf16_sw128(CUtensorMap* out, const void* base, const uint64_t dims[3], const uint64_t strides_bytes[2],
                    const uint32_t box[3]) {
  std::call_once(g_once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &st) == cudaSuccess &&
        st == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<EncodeFn>(p);
  });
  if (g_encode == nullptr) { set_last_error("cuTensorMapEncodeTiled is not available from this driver"); return 1; }
  cuuint64_t d[3] = {dims[0], dims[1], dims[2]};
  cuuint64_t s[2] = {strides_bytes[0], strides_bytes[1]};
  cuuint32_t b[3] = {box[0], box[1], box[2]};
  cuuint32_t e[3] = {1, 1, 1};
  const CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), d, s, b, e,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed (%d): base %p dims {%llu,%llu,%llu} strides {%llu,%llu} box {%u,%u,%u}",
                   (int)r, base, (unsigned long long)d[0], (unsigned long long)d[1], (unsigned long long)d[2],
                   (unsigned long long)s[0], (unsigned long long)s[1], b[0], b[1], b[2]);
    return 1;
  }
  return 0;
}


// General form: rank 2..5, swizzle 0 = none, 1 = 32 B, 2 = 64 B, 3 = 128 B.
int tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, int swizzle) {
  uint64_t d3[3] = {1, 1, 1};
  uint64_t s2[2] = {16, 16};
  uint32_t b3[3] = {1, 1, 1};
  if (tmap_bf16_sw128(out, base, d3, s2, b3) != 0 && g_encode == nullptr) return 1;   // resolves g_encode (error already set)
  if (rank < 2 || rank > 5) { set_last_error("tmap_bf16: rank must be 2..5"); return 1; }
  cuuint64_t d[5]; cuuint64_t s[4]; cuuint32_t b[5]; cuuint32_t e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
  const CUtensorMapSwizzle sw = swizzle == 3 ? CU_TENSOR_MAP_SWIZZLE_128B : swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_64B
                              : swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  const CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, e,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("cuTensorMapEncodeTiled failed (%d) for a rank-%d map", (int)r, rank); return 1; }
  return 0;
}

}  // namespace ddls
