// tensormap.h — TMA descriptor helpers (see tensormap.cu).
#pragma once

#include <cuda.h>
#include <stdint.h>

namespace b200 {

// 2-D row-major tensor [outer][inner] of elem_bytes elements; box = box_outer rows x box_inner elements.
int tmap_tiled_2d(CUtensorMap* map, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                  uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, bool swizzle128);

// [rows][chunks*64] bf16 read as `box_chunks` consecutive 64-column planes of `box_rows` rows with ONE TMA instruction;
// the planes land back to back in shared memory. Returns -1 (with b200_last_error) if the driver rejects the layout.
int tmap_chunked_rows(CUtensorMap* map, const void* base, uint64_t rows, uint64_t row_stride_bytes, uint32_t chunks,
                      uint32_t box_rows, uint32_t box_chunks);

// Tiled view {C, W, H, N} of an NHWC bf16 tensor with a box of box_h x box_w pixels x box_c channels (128-byte swizzle).
int tmap_nhwc_tile(CUtensorMap* map, const void* base, int n, int h, int w, int c, uint32_t box_c, uint32_t box_w,
                   uint32_t box_h);

// im2col view of an NHWC bf16 tensor. lower/upper are the bounding-box corners of the filter's base pixel
// (fprop: lower = -pad_lo, upper = pad_hi - (filter-1)); stride_* are the traversal strides.
int tmap_im2col_nhwc(CUtensorMap* map, const void* base, int n, int h, int w, int c, int lower_w, int lower_h,
                     int upper_w, int upper_h, int stride_w, int stride_h, uint32_t channels_per_pixel,
                     uint32_t pixels_per_column);

}  // namespace b200
