// loss.cu — softmax cross-entropy forward + backward in one pass (K8 in SURVEY.md §2.6; TF's
// softmax_cross_entropy_with_logits under tf_cnn_benchmarks' loss_function). One warp-group per row:
// 1001 classes fit in registers/shared memory, so logits are read once and dlogits written once.
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kT = 256;

__device__ __forceinline__ float block_reduce(float v, float* sm, bool is_max) {
  v = is_max ? warp_max(v) : warp_sum(v);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  float r = sm[0];
#pragma unroll
  for (int k = 1; k < kT / 32; ++k) r = is_max ? fmaxf(r, sm[k]) : r + sm[k];
  return r;
}

__global__ void __launch_bounds__(kT) xent_rows_kernel(const float* __restrict__ logits, int ld,
                                                       const int64_t* __restrict__ labels, int classes,
                                                       float* __restrict__ loss_rows, b200_bf16* __restrict__ dlogits,
                                                       int ld_d, float gscale) {
  __shared__ float sm[kT / 32];
  const int row = blockIdx.x;
  const float* x = logits + (int64_t)row * ld;
  float mx = -INFINITY;
  for (int j = threadIdx.x; j < classes; j += kT) mx = fmaxf(mx, x[j]);
  mx = block_reduce(mx, sm, true);
  float se = 0.0f;
  for (int j = threadIdx.x; j < classes; j += kT) se += __expf(x[j] - mx);
  se = block_reduce(se, sm, false);
  const int lab = (int)labels[row];
  const float lse = mx + __logf(se);
  if (threadIdx.x == 0) loss_rows[row] = lse - x[lab];
  if (dlogits) {
    __nv_bfloat16* d = reinterpret_cast<__nv_bfloat16*>(dlogits) + (int64_t)row * ld_d;
    const float inv = 1.0f / se;
    for (int j = threadIdx.x; j < ld_d; j += kT) {
      float v = 0.0f;
      if (j < classes) v = (__expf(x[j] - mx) * inv - (j == lab ? 1.0f : 0.0f)) * gscale;
      d[j] = __float2bfloat16(v);
    }
  }
}

__global__ void __launch_bounds__(kT) mean_kernel(const float* __restrict__ v, int n, float* __restrict__ out) {
  __shared__ float sm[kT / 32];
  float s = 0.0f;
  for (int j = threadIdx.x; j < n; j += kT) s += v[j];
  s = block_reduce(s, sm, false);
  if (threadIdx.x == 0) *out = s / (float)n;
}

__global__ void __launch_bounds__(256) normalize_u8_kernel(const uint8_t* __restrict__ x, b200_bf16* __restrict__ y,
                                                           int64_t pixels, int cin, int cout) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += (int64_t)gridDim.x * blockDim.x) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(y) + p * cout;
    for (int c = 0; c < cout; ++c) {
      float v = c < cin ? (float)x[p * cin + c] / 127.5f - 1.0f : 0.0f;
      o[c] = __float2bfloat16(v);
    }
  }
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_softmax_xent(const float* logits, int32_t ld, const int64_t* labels, int32_t b, int32_t classes,
                      float* loss_rows, float* loss, b200_bf16* dlogits, int32_t ld_d, float grad_scale, void* stream) {
  B200_CHECK_ARG(logits && labels && loss_rows && loss, "softmax_xent: null pointer");
  B200_CHECK_ARG(b > 0 && classes > 0 && ld >= classes && (!dlogits || ld_d >= classes), "softmax_xent: bad shape");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  xent_rows_kernel<<<b, kT, 0, st>>>(logits, ld, labels, classes, loss_rows, dlogits, ld_d, grad_scale / (float)b);
  B200_LAUNCH_CHECK();
  mean_kernel<<<1, kT, 0, st>>>(loss_rows, b, loss);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_normalize_u8(const uint8_t* x, b200_bf16* y, int64_t pixels, int32_t c_in, int32_t c_out, void* stream) {
  B200_CHECK_ARG(x && y && pixels > 0 && c_in > 0 && c_out >= c_in, "normalize_u8: bad args");
  normalize_u8_kernel<<<grid_for(pixels, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, pixels, c_in, c_out);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
