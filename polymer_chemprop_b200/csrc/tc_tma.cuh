// Tensor-map (TMA) plumbing: host-side encoding of 2-D fp32 row-major matrices and the device-side
// cp.async.bulk.tensor load.  Maps are encoded with data type TFLOAT32, so the TMA engine rounds fp32 to
// tf32 while it copies: operands loaded this way need no transform pass before tcgen05.mma.
// cuTensorMapEncodeTiled is fetched through cudaGetDriverEntryPoint (no link-time libcuda dependency).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace wd {

typedef CUresult (*PFN_wdEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_wdEncodeTiled tma_encode_fn() {
  static PFN_wdEncodeTiled fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<PFN_wdEncodeTiled>(p);
  }();
  return fn;
}

enum TmaSwizzle { TMA_SW128 = 0, TMA_SW128_ATOM32 = 1, TMA_NO_SWIZZLE = 2 };   // NO_SWIZZLE: rows land back to back as they are

// Map over X[rows, cols] (row stride ld floats; base and rows 16-byte aligned), box = box_rows x 32 columns.
// Elements outside [rows, cols] read as zero.
inline int tma_make_2d(CUtensorMap* map, const float* X, int64_t rows, int cols, int ld, int box_rows, TmaSwizzle sw,
                       bool round_tf32 = true) {
  PFN_wdEncodeTiled enc = tma_encode_fn();
  if (!enc) return fail(WD_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  const cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(map, round_tf32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(X), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE,
                         sw == TMA_SW128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                         : (sw == TMA_NO_SWIZZLE ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B),
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(WD_ERR_CUDA, "cuTensorMapEncodeTiled failed with code %d", (int)r);
  return WD_OK;
}

namespace ptx {
// box of `map` at (column c0, row c1) -> shared memory; completes `bar` with the full box byte count
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// the same load issued by a CTA of a pair (cta_group::2): the box lands in THIS CTA's shared memory, its bytes are
// counted on the barrier at `bar`'s offset in the LEADER's shared memory
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
}  // namespace ptx

}  // namespace wd
