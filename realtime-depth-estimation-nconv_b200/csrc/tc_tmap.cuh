// Host-side encoder of the TMA tensor maps shared by the tcgen05 kernels (tc_conv.cu, tc_conv_x.cuh, nconv_tc.cu).
#pragma once
#include "common.cuh"
#include <cuda.h>
#include <cuda_runtime.h>

namespace nconv {

typedef CUresult (*TcEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// P8 / P8Q source seen as [planes][rows][pitch px * 4 words] (32-bit words: a staged row of P pixels is ONE
// contiguous inner extent of 16*P bytes) with box [4P][box_rows][1].
static inline int tc_make_tmap(CUtensorMap* tm, const void* base, int planes, int rows, int pitch, int P, int box_rows) {
  static TcEncodeFn enc = nullptr;
  if (!enc) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) return NCONV_ERR_UNSUPPORTED;
    enc = (TcEncodeFn)fn;
  }
  if (P * 4 > 256 || box_rows > 256) return NCONV_ERR_UNSUPPORTED;
  const cuuint64_t gdim[3] = {(cuuint64_t)pitch * 4, (cuuint64_t)rows, (cuuint64_t)planes};
  const cuuint64_t gstr[2] = {(cuuint64_t)pitch * 16, (cuuint64_t)rows * pitch * 16};
  const cuuint32_t box[3] = {(cuuint32_t)P * 4, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void*>(base), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : NCONV_ERR_BAD_ARG;
}


}  // namespace nconv
