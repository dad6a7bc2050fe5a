// TMA descriptor helpers shared by the tcgen05 kernels (gram_tc.cu, kf3_tc.cu).
// cuTensorMapEncodeTiled is resolved at run time through cudaGetDriverEntryPoint, so the library
// loads on a box without a GPU driver.
#include <cuda.h>

#include "gemm_tc.cuh"

namespace sigsvgd {
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

}  // namespace

// row-major matrix [outer x inner] of 2-byte (fp16) or 4-byte (fp32) elements with row pitch ld
// (elements); box = [box_outer x box_inner], 128-byte swizzle (box_inner * elem_bytes == 128)
static int make_tmap_any(CUtensorMap* tm, const void* base, int elem_bytes, int64_t inner,
                         int64_t outer, int64_t ld, uint32_t box_inner, uint32_t box_outer) {
  // swizzle span = bytes of one box row (128 or 64)
  const CUtensorMapSwizzle swz =
      box_inner * elem_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return SIGSVGD_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || ((ld * elem_bytes) & 15)) {
    set_error("TMA operands must be 16-byte aligned with a row pitch that is a multiple of 16 bytes");
    return SIGSVGD_ERR_INVALID;
  }
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * (cuuint64_t)elem_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = fn(tm, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                   2, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld", (int)rc,
              (long long)inner, (long long)outer, (long long)ld);
    return SIGSVGD_ERR_CUDA;
  }
  return 0;
}

int make_tmap_f16(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_any(tm, base, 2, inner, outer, ld, box_inner, box_outer);
}
int make_tmap_f32(CUtensorMap* tm, const float* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_any(tm, base, 4, inner, outer, ld, box_inner, box_outer);
}

bool tc_available() { return encode_fn() != nullptr; }

}  // namespace sigsvgd
