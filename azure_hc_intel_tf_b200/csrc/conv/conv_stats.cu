// conv_stats.cu — folds the per-row-group column statistics emitted by the conv epilogue into per-channel sums.
// The fold is over a fixed order, so BatchNorm statistics stay bitwise reproducible.
#include "b200conv.h"
#include "common.cuh"

namespace b200 {
namespace {

// grid.x covers channels in chunks of 32, block = 32 (channels) x 8 (row slices)
__global__ void __launch_bounds__(256) stats_reduce_kernel(const float* __restrict__ part, int64_t rows, int k,
                                                           float* __restrict__ sums) {
  __shared__ float sm[2][8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int ch = blockIdx.x * 32 + tx;
  float s1 = 0.0f, s2 = 0.0f;
  if (ch < k) {
    for (int64_t r = ty; r < rows; r += 8) {
      const float* p = part + r * 2 * k;
      s1 += p[ch];
      s2 += p[k + ch];
    }
  }
  sm[0][ty][tx] = s1;
  sm[1][ty][tx] = s2;
  __syncthreads();
  if (ty == 0 && ch < k) {
    float a = 0.0f, b = 0.0f;
#pragma unroll
    for (int y = 0; y < 8; ++y) {
      a += sm[0][y][tx];
      b += sm[1][y][tx];
    }
    sums[ch] = a;
    sums[k + ch] = b;
  }
}

}  // namespace
}  // namespace b200

extern "C" int b200_conv_stats_reduce(const float* col_stats, int64_t rows, int32_t k, float* sums, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(col_stats && sums && rows > 0 && k > 0, "conv_stats_reduce: bad args");
  stats_reduce_kernel<<<(k + 31) / 32, 256, 0, static_cast<cudaStream_t>(stream)>>>(col_stats, rows, k, sums);
  B200_LAUNCH_CHECK();
  return 0;
}
