// Tensor maps for TMA row gathers (cp.async.bulk.tensor.2d ... tile::gather4): a [rows, cols] fp32
// matrix with a byte pitch, box = one 128-byte K-block of ONE row, SWIZZLE_128B.  A gather4 names
// four row indices and lands their K-blocks as four consecutive 128-byte rows of a K-major
// SWIZZLE_128B operand tile -- the layout tcgen05.mma reads -- without any thread touching the data.
// The encoder is a driver entry point; it is fetched through the runtime so that the library does
// not link against libcuda.
#pragma once
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>

#include "common.cuh"

namespace sgnn {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn tensormap_encoder() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// rows x cols fp32 at `base` (16-byte aligned), `pitch_bytes` between rows (multiple of 16)
static inline int make_row_gather_map(CUtensorMap* map, const void* base, long rows, int cols, long pitch_bytes) {
  EncodeTiledFn enc = tensormap_encoder();
  if (!enc) {
    sgnn_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return SGNN_ERR_CUDA;
  }
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)kKBlock, 1u};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    sgnn_set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return SGNN_ERR_CUDA;
  }
  return SGNN_OK;
}

// same matrix, box = `box_rows` rows x one 128-byte K-block: one TMA tile load lands a whole K-block of
// a row-major [rows, cols] matrix as a K-major SWIZZLE_128B operand tile (rows past the end read as 0)
static inline int make_row_tile_map(CUtensorMap* map, const void* base, long rows, int cols, long pitch_bytes,
                                    int box_rows) {
  EncodeTiledFn enc = tensormap_encoder();
  if (!enc) {
    sgnn_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return SGNN_ERR_CUDA;
  }
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)kKBlock, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    sgnn_set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return SGNN_ERR_CUDA;
  }
  return SGNN_OK;
}

#ifdef __CUDACC__
// one box of a 2D tensor map (inner coordinate `col`, outer coordinate `row`) -> shared memory
__device__ __forceinline__ void tma_tile_2d(void* smem_dst, const CUtensorMap* map, int col, int row, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(col), "r"(row)
      : "memory");
}
// four rows (row indices r0..r3) x one 128-byte K-block starting at column `col` -> 512 bytes at dst
__device__ __forceinline__ void tma_gather4(void* smem_dst, const CUtensorMap* map, int col, int r0, int r1, int r2,
                                            int r3, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}
__device__ __forceinline__ void tensormap_prefetch(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
#endif

}  // namespace sgnn
