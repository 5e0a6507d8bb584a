// tensormap.cu — host-side construction of TMA descriptors (CUtensorMap) for the conv/GEMM kernels.
// The driver entry points are fetched at run time (cudaGetDriverEntryPoint), so the library has no link-time
// dependency on libcuda and still loads on a machine without a driver.
#include "conv/tensormap.h"

#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

#include "common.cuh"

namespace b200 {
namespace {

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
using EncodeIm2colFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn g_tiled = nullptr;
EncodeIm2colFn g_im2col = nullptr;
std::once_flag g_once;
int g_init_rc = 0;

void init_entry_points() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    g_init_rc = -1;
    return;
  }
  g_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    g_init_rc = -1;
    return;
  }
  g_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
}

int ensure() {
  std::call_once(g_once, init_entry_points);
  if (g_init_rc != 0) set_error("cuTensorMapEncode* entry points are unavailable (no CUDA driver?)");
  return g_init_rc;
}

}  // namespace

int tmap_tiled_2d(CUtensorMap* map, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                  uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, bool swizzle128) {
  if (ensure() != 0) return -1;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = g_tiled(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu stride=%llu box=%ux%u", (int)r,
              (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)row_stride_bytes, box_inner,
              box_outer);
    return -1;
  }
  return 0;
}

int tmap_chunked_rows(CUtensorMap* map, const void* base, uint64_t rows, uint64_t row_stride_bytes, uint32_t chunks,
                      uint32_t box_rows, uint32_t box_chunks) {
  // [rows][chunks*64] bf16 viewed as {64, rows, chunks}: one box = box_chunks planes of [box_rows][64], i.e. the
  // planes land back to back in shared memory ([chunk][row][128 B]) with a single TMA instruction
  if (ensure() != 0) return -1;
  cuuint64_t dims[3] = {64, rows, chunks};
  cuuint64_t strides[2] = {row_stride_bytes, 128};
  cuuint32_t box[3] = {64, box_rows, box_chunks};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_tiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (chunked rows) failed (%d)", (int)r);
    return -1;
  }
  return 0;
}

int tmap_nhwc_tile(CUtensorMap* map, const void* base, int n, int h, int w, int c, uint32_t box_c, uint32_t box_w,
                   uint32_t box_h) {
  // plain tiled view of an NHWC bf16 tensor {C, W, H, N}; box = box_h x box_w pixels x box_c channels of one image.
  // Coordinates may be negative / past the edge: those elements read as zero and are not written by stores.
  if (ensure() != 0) return -1;
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)w * c * 2, (cuuint64_t)h * w * c * 2};
  cuuint32_t box[4] = {box_c, box_w, box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = g_tiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (nhwc tile) failed (%d): nhwc=%d,%d,%d,%d box=%u,%u,%u", (int)r, n, h, w, c, box_c,
              box_w, box_h);
    return -1;
  }
  return 0;
}

int tmap_im2col_nhwc(CUtensorMap* map, const void* base, int n, int h, int w, int c, int lower_w, int lower_h,
                     int upper_w, int upper_h, int stride_w, int stride_h, uint32_t channels_per_pixel,
                     uint32_t pixels_per_column) {
  if (ensure() != 0) return -1;
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)w * c * 2, (cuuint64_t)h * w * c * 2};
  int lower[2] = {lower_w, lower_h};
  int upper[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride_w, (cuuint32_t)stride_h, 1};
  CUresult r = g_im2col(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper,
                        channels_per_pixel, pixels_per_column, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeIm2col failed (%d): nhwc=%d,%d,%d,%d lower=%d,%d upper=%d,%d stride=%d,%d cpp=%u ppc=%u",
              (int)r, n, h, w, c, lower_w, lower_h, upper_w, upper_h, stride_w, stride_h, channels_per_pixel,
              pixels_per_column);
    return -1;
  }
  return 0;
}

}  // namespace b200
