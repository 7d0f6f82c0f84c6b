// TMA tensor maps (cuTensorMapEncodeTiled through the runtime's driver entry point: no libcuda link) and the
// tensor-tile bulk load.  Used for operands that are consumed K-major with 128-byte swizzle (acx_umma.cuh).
#pragma once
#include <cuda.h>
#include "acx_common.cuh"

// 2-D fp32 tensor, `inner` contiguous elements per row, `rows` rows, row pitch in bytes (multiple of 16);
// box = box_inner x box_rows elements, 128-byte swizzle (box_inner * 4 must be 128), out-of-range elements read 0.
inline int acx_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows, uint64_t pitch_bytes,
                       uint32_t box_inner, uint32_t box_rows) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr) return -1;
    fn = reinterpret_cast<EncodeFn>(p);
  }
  const cuuint64_t gdim[2] = {inner, rows};
  const cuuint64_t gstride[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box_inner, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)r;
}

#ifdef __CUDACC__
// box at (c_inner, c_row) -> shared memory (1024-byte aligned), completion counted in bytes on `bar`
__device__ inline void tma_load_2d(void* dst_smem, const CUtensorMap* map, int c_inner, int c_row, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(acx_smem_u32(dst_smem)), "l"(reinterpret_cast<uint64_t>(map)), "r"(c_inner), "r"(c_row),
        "r"(acx_smem_u32(bar))
      : "memory");
}
#endif
