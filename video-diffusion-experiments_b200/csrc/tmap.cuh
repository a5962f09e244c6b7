// Host-side TMA descriptor construction.  cuTensorMapEncodeTiled is fetched through the runtime's
// driver entry point so the library links against cudart only (and still loads on a GPU-less box).
#pragma once
#include "common.cuh"

namespace b200ad {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_tiled();   // api.cu

// Tiled tensor map, innermost dim contiguous.  dims/box innermost-first; strides_bytes for dims 1..rank-1.
inline int make_tmap(CUtensorMap* out, CUtensorMapDataType dtype, const void* gptr, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_tiled();
  B200AD_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  B200AD_REQUIRE((reinterpret_cast<uintptr_t>(gptr) & 15) == 0, "TMA base pointer must be 16-byte aligned");
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    B200AD_REQUIRE(box[i] >= 1 && box[i] <= 256, "TMA box dim %d = %u out of range", i, box[i]);
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    B200AD_REQUIRE((strides_bytes[i] & 15) == 0, "TMA stride %d = %llu not a multiple of 16 bytes", i,
                   (unsigned long long)strides_bytes[i]);
  }
  CUresult r = fn(out, dtype, (cuuint32_t)rank, const_cast<void*>(gptr),
                  gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200AD_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

inline int make_tmap_bf16(CUtensorMap* out, const void* gptr, int rank, const uint64_t* dims,
                          const uint64_t* strides_bytes, const uint32_t* box,
                          CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  return make_tmap(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, gptr, rank, dims, strides_bytes, box, swz);
}
inline int make_tmap_f32(CUtensorMap* out, const void* gptr, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box,
                         CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  return make_tmap(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, gptr, rank, dims, strides_bytes, box, swz);
}

}  // namespace b200ad
