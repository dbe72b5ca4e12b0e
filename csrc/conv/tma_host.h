// Host-side helpers to build TMA tensor maps (cuTensorMapEncodeTiled resolved at run time through
// the runtime's driver entry point, so the library has no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>

namespace b200 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  return fn;
}

inline CUtensorMapSwizzle swizzle_for_bytes(int inner_bytes) {
  if (inner_bytes <= 0) return CU_TENSOR_MAP_SWIZZLE_NONE;       // plain rows (warp-private store slabs)
  if (inner_bytes >= 128) return CU_TENSOR_MAP_SWIZZLE_128B;
  if (inner_bytes >= 64) return CU_TENSOR_MAP_SWIZZLE_64B;
  return CU_TENSOR_MAP_SWIZZLE_32B;
}

// rank-R bf16 tensor map; dims[0] is the contiguous one; strides in elements for dims 1..R-1.
inline bool make_map(CUtensorMap* m, const void* base, int rank, const long long* dims, const long long* strides_el,
                     const int* box, int inner_bytes) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = (cuuint64_t)dims[i]; b[i] = (cuuint32_t)box[i]; e[i] = 1; }
  for (int i = 1; i < rank; ++i) s[i - 1] = (cuuint64_t)strides_el[i] * 2;
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, e,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for_bytes(inner_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}


// Same for fp32 tensors with no swizzle (reduce-add epilogues).
inline bool make_map_f32(CUtensorMap* m, const void* base, int rank, const long long* dims,
                         const long long* strides_el, const int* box) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = (cuuint64_t)dims[i]; b[i] = (cuuint32_t)box[i]; e[i] = 1; }
  for (int i = 1; i < rank; ++i) s[i - 1] = (cuuint64_t)strides_el[i] * 4;
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, e,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

}  // namespace b200
