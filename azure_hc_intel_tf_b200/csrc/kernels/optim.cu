// optim.cu — flat-buffer momentum SGD and LARS (single-GPU / post-allreduce variants).
//
// Role in the reference stack: TensorFlow's ApplyMomentum kernel run once per variable (161 launches for
// ResNet-50) under --optimizer=momentum (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:73); K10 / K-O1 in
// SURVEY.md. Here all parameters live in one plan-ordered flat buffer, so the whole model (or one fusion
// bucket) is one launch; the fused allreduce+update variant lives in csrc/comm/.
#include "common.cuh"

namespace b200 {
namespace {

__global__ void __launch_bounds__(256) sgd_flat_kernel(float* __restrict__ w, float* __restrict__ m,
                                                       const float* __restrict__ g, b200_bf16* __restrict__ shadow,
                                                       int64_t nblk, const int32_t* __restrict__ block_tensor,
                                                       const float* __restrict__ tdecay, const float* __restrict__ tlars,
                                                       const float* __restrict__ trust, float lr, float mu, float wd,
                                                       float gscale, const float* __restrict__ lr_dev) {
  if (lr_dev) lr = __ldg(lr_dev);    // a schedule can change the rate under a captured CUDA graph
  // one CTA handles one 1024-element block per iteration: 256 threads x float4
  for (int64_t b = blockIdx.x; b < nblk; b += gridDim.x) {
    float dec = wd, tr = 1.0f;
    if (block_tensor) {
      int t = block_tensor[b];
      if (t < 0) continue;     // padding between tensors: leave untouched (stays zero)
      dec = wd * tdecay[t];
      if (trust && tlars[t] != 0.0f) tr = trust[t];
    }
    int64_t i = b * 1024 + threadIdx.x * 4;
    float4 wv = *reinterpret_cast<const float4*>(w + i);
    float4 mv = *reinterpret_cast<const float4*>(m + i);
    float4 gv = __ldcs(reinterpret_cast<const float4*>(g + i));
    float ww[4] = {wv.x, wv.y, wv.z, wv.w}, mm[4] = {mv.x, mv.y, mv.z, mv.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float ge = (gg[k] * gscale + dec * ww[k]) * tr;
      mm[k] = mu * mm[k] + ge;
      ww[k] -= lr * mm[k];
    }
    *reinterpret_cast<float4*>(w + i) = make_float4(ww[0], ww[1], ww[2], ww[3]);
    *reinterpret_cast<float4*>(m + i) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    if (shadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(ww[0], ww[1]), hi = __floats2bfloat162_rn(ww[2], ww[3]);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(shadow + i) = pk;
    }
  }
}

// RMSProp and Adam on the same flat buffers (TensorFlow semantics, as tf_cnn_benchmarks --optimizer=rmsprop|adam builds
// them): kRule 1: ms = rho*ms + (1-rho)*g^2; mom = mu*mom + lr*g/sqrt(ms + eps); w -= mom   (ApplyRMSProp)
//       kRule 2: m = b1*m + (1-b1)*g; v = b2*v + (1-b2)*g^2; w -= lr_t*m/(sqrt(v) + eps)     (ApplyAdam; lr_t carries
//                the bias correction lr*sqrt(1-b2^t)/(1-b1^t), computed on the host and read from lr_dev)
// s1 is the first-moment-like slot (mom / m), s2 the second-moment slot (ms / v). g includes 1/N and weight decay.
template <int kRule>
__global__ void __launch_bounds__(256) adaptive_flat_kernel(float* __restrict__ w, float* __restrict__ s1, float* __restrict__ s2,
                                                            const float* __restrict__ g, b200_bf16* __restrict__ shadow,
                                                            int64_t nblk, const int32_t* __restrict__ block_tensor,
                                                            const float* __restrict__ tdecay, float lr, float a, float b,
                                                            float eps, float wd, float gscale,
                                                            const float* __restrict__ lr_dev) {
  if (lr_dev) lr = __ldg(lr_dev);
  for (int64_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
    float dec = wd;
    if (block_tensor) {
      int t = block_tensor[blk];
      if (t < 0) continue;
      dec = wd * tdecay[t];
    }
    int64_t i = blk * 1024 + threadIdx.x * 4;
    float4 wv = *reinterpret_cast<const float4*>(w + i);
    float4 av = *reinterpret_cast<const float4*>(s1 + i);
    float4 bv = *reinterpret_cast<const float4*>(s2 + i);
    float4 gv = __ldcs(reinterpret_cast<const float4*>(g + i));
    float ww[4] = {wv.x, wv.y, wv.z, wv.w}, aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
    float gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float ge = gg[k] * gscale + dec * ww[k];
      if (kRule == 1) {
        bb[k] = a * bb[k] + (1.0f - a) * ge * ge;
        aa[k] = b * aa[k] + lr * ge * rsqrtf(bb[k] + eps);
        ww[k] -= aa[k];
      } else {
        aa[k] = a * aa[k] + (1.0f - a) * ge;
        bb[k] = b * bb[k] + (1.0f - b) * ge * ge;
        ww[k] -= lr * aa[k] / (sqrtf(bb[k]) + eps);
      }
    }
    *reinterpret_cast<float4*>(w + i) = make_float4(ww[0], ww[1], ww[2], ww[3]);
    *reinterpret_cast<float4*>(s1 + i) = make_float4(aa[0], aa[1], aa[2], aa[3]);
    *reinterpret_cast<float4*>(s2 + i) = make_float4(bb[0], bb[1], bb[2], bb[3]);
    if (shadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(ww[0], ww[1]), hi = __floats2bfloat162_rn(ww[2], ww[3]);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(shadow + i) = pk;
    }
  }
}

__global__ void __launch_bounds__(256) lars_norms_kernel(const float* __restrict__ w, const float* __restrict__ g,
                                                         int64_t nblk, const int32_t* __restrict__ block_tensor,
                                                         float* __restrict__ norms, float gscale) {
  __shared__ float red[2][8];
  for (int64_t b = blockIdx.x; b < nblk; b += gridDim.x) {
    int t = block_tensor[b];
    if (t < 0) continue;
    int64_t i = b * 1024 + threadIdx.x * 4;
    float4 wv = *reinterpret_cast<const float4*>(w + i);
    float4 gv = *reinterpret_cast<const float4*>(g + i);
    float sw = wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w;
    float sg = (gv.x * gv.x + gv.y * gv.y + gv.z * gv.z + gv.w * gv.w) * gscale * gscale;
    sw = warp_sum(sw);
    sg = warp_sum(sg);
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
      red[0][warp] = sw;
      red[1][warp] = sg;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
      float s = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += red[threadIdx.x][k];
      atomicAdd(norms + 2 * t + threadIdx.x, s);
    }
    __syncthreads();
  }
}

__global__ void lars_trust_kernel(const float* __restrict__ norms, const float* __restrict__ tdecay,
                                  const float* __restrict__ tlars, float* __restrict__ trust, int n, float eta,
                                  float wd, float eps) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  float r = 1.0f;
  if (tlars[t] != 0.0f) {
    float wn = sqrtf(norms[2 * t]), gn = sqrtf(norms[2 * t + 1]);
    if (wn > 0.0f && gn > 0.0f) r = eta * wn / (gn + wd * tdecay[t] * wn + eps);
  }
  trust[t] = r;
}

// Deterministic LARS norms: (1) one (sum w^2, sum g^2) pair per 1024-element block, reduced by a fixed tree;
// (2) one CTA per tensor folds that tensor's blocks in a fixed pattern and writes the trust ratio. Identical inputs
// give bitwise identical trust ratios on every rank, which keeps data-parallel replicas bit-identical.
__global__ void __launch_bounds__(256) lars_block_norms_kernel(const float* __restrict__ w, const float* __restrict__ g,
                                                               int64_t nblk, float* __restrict__ partials, float gscale) {
  __shared__ float red[2][8];
  for (int64_t b = blockIdx.x; b < nblk; b += gridDim.x) {
    int64_t i = b * 1024 + threadIdx.x * 4;
    float4 wv = *reinterpret_cast<const float4*>(w + i);
    float4 gv = *reinterpret_cast<const float4*>(g + i);
    float sw = wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w;
    float sg = (gv.x * gv.x + gv.y * gv.y + gv.z * gv.z + gv.w * gv.w) * gscale * gscale;
    sw = warp_sum(sw);
    sg = warp_sum(sg);
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
      red[0][warp] = sw;
      red[1][warp] = sg;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
      float s = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += red[threadIdx.x][k];
      partials[2 * b + threadIdx.x] = s;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) lars_fold_trust_kernel(const float* __restrict__ partials,
                                                              const int32_t* __restrict__ blk_lo,
                                                              const int32_t* __restrict__ blk_cnt,
                                                              const float* __restrict__ tdecay,
                                                              const float* __restrict__ tlars, float* __restrict__ trust,
                                                              float eta, float wd, float eps) {
  __shared__ float red[2][8];
  const int t = blockIdx.x;
  const int lo = blk_lo[t], cnt = blk_cnt[t];
  float sw = 0.0f, sg = 0.0f;
  for (int b = threadIdx.x; b < cnt; b += 256) {
    sw += partials[2 * (lo + b)];
    sg += partials[2 * (lo + b) + 1];
  }
  sw = warp_sum(sw);
  sg = warp_sum(sg);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    red[0][warp] = sw;
    red[1][warp] = sg;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.0f, b2 = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a += red[0][k];
      b2 += red[1][k];
    }
    float r = 1.0f;
    if (tlars[t] != 0.0f) {
      float wn = sqrtf(a), gn = sqrtf(b2);
      if (wn > 0.0f && gn > 0.0f) r = eta * wn / (gn + wd * tdecay[t] * wn + eps);
    }
    trust[t] = r;
  }
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_sgd_momentum_flat(float* master, float* mom, const float* grad, b200_bf16* shadow, int64_t n,
                           const int32_t* block_tensor, const float* tensor_decay, const float* tensor_lars,
                           const float* trust, float lr, float mu, float wd, float grad_scale, const float* lr_dev,
                           void* stream) {
  B200_CHECK_ARG(master && mom && grad, "sgd_momentum_flat: null pointer");
  B200_CHECK_ARG(n > 0 && n % 1024 == 0, "sgd_momentum_flat: n must be a positive multiple of 1024");
  B200_CHECK_ARG(!block_tensor || tensor_decay, "sgd_momentum_flat: block table needs tensor_decay");
  B200_CHECK_ARG(!trust || tensor_lars, "sgd_momentum_flat: trust needs tensor_lars");
  int64_t nblk = n / 1024;
  int grid = (int)(nblk < (int64_t)kNumSMs * 8 ? nblk : (int64_t)kNumSMs * 8);
  sgd_flat_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(master, mom, grad, shadow, nblk, block_tensor,
                                                                      tensor_decay, tensor_lars, trust, lr, mu, wd,
                                                                      grad_scale, lr_dev);
  B200_LAUNCH_CHECK();
  return 0;
}

static int adaptive_flat(int rule, float* master, float* s1, float* s2, const float* grad, b200_bf16* shadow, int64_t n,
                         const int32_t* block_tensor, const float* tensor_decay, float lr, float a, float b, float eps,
                         float wd, float grad_scale, const float* lr_dev, void* stream) {
  B200_CHECK_ARG(master && s1 && s2 && grad, "adaptive_flat: null pointer");
  B200_CHECK_ARG(n > 0 && n % 1024 == 0, "adaptive_flat: n must be a positive multiple of 1024");
  B200_CHECK_ARG(!block_tensor || tensor_decay, "adaptive_flat: block table needs tensor_decay");
  int64_t nblk = n / 1024;
  int grid = (int)(nblk < (int64_t)kNumSMs * 8 ? nblk : (int64_t)kNumSMs * 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (rule == 1)
    adaptive_flat_kernel<1><<<grid, 256, 0, st>>>(master, s1, s2, grad, shadow, nblk, block_tensor, tensor_decay, lr, a, b,
                                                  eps, wd, grad_scale, lr_dev);
  else
    adaptive_flat_kernel<2><<<grid, 256, 0, st>>>(master, s1, s2, grad, shadow, nblk, block_tensor, tensor_decay, lr, a, b,
                                                  eps, wd, grad_scale, lr_dev);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_rmsprop_flat(float* master, float* mom, float* ms, const float* grad, b200_bf16* shadow, int64_t n,
                      const int32_t* block_tensor, const float* tensor_decay, float lr, float rho, float mu, float eps,
                      float wd, float grad_scale, const float* lr_dev, void* stream) {
  return adaptive_flat(1, master, mom, ms, grad, shadow, n, block_tensor, tensor_decay, lr, rho, mu, eps, wd, grad_scale,
                       lr_dev, stream);
}

int b200_adam_flat(float* master, float* m, float* v, const float* grad, b200_bf16* shadow, int64_t n,
                   const int32_t* block_tensor, const float* tensor_decay, float lr_t, float beta1, float beta2, float eps,
                   float wd, float grad_scale, const float* lr_dev, void* stream) {
  return adaptive_flat(2, master, m, v, grad, shadow, n, block_tensor, tensor_decay, lr_t, beta1, beta2, eps, wd, grad_scale,
                       lr_dev, stream);
}

int b200_lars_norms(const float* master, const float* grad, int64_t n, const int32_t* block_tensor, float* norms,
                    float grad_scale, void* stream) {
  B200_CHECK_ARG(master && grad && block_tensor && norms, "lars_norms: null pointer");
  B200_CHECK_ARG(n > 0 && n % 1024 == 0, "lars_norms: n must be a positive multiple of 1024");
  int64_t nblk = n / 1024;
  int grid = (int)(nblk < (int64_t)kNumSMs * 8 ? nblk : (int64_t)kNumSMs * 8);
  lars_norms_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(master, grad, nblk, block_tensor, norms,
                                                                        grad_scale);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_lars_trust_det(const float* master, const float* grad, int64_t n, float* block_partials,
                        const int32_t* tensor_blk_lo, const int32_t* tensor_blk_cnt, const float* tensor_decay,
                        const float* tensor_lars, float* trust, int32_t num_tensors, float eta, float wd, float eps,
                        float grad_scale, void* stream) {
  B200_CHECK_ARG(master && grad && block_partials && tensor_blk_lo && tensor_blk_cnt && tensor_decay && tensor_lars && trust,
                 "lars_trust_det: null pointer");
  B200_CHECK_ARG(n > 0 && n % 1024 == 0 && num_tensors > 0, "lars_trust_det: n must be a positive multiple of 1024");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int64_t nblk = n / 1024;
  int grid = (int)(nblk < (int64_t)kNumSMs * 8 ? nblk : (int64_t)kNumSMs * 8);
  lars_block_norms_kernel<<<grid, 256, 0, st>>>(master, grad, nblk, block_partials, grad_scale);
  B200_LAUNCH_CHECK();
  lars_fold_trust_kernel<<<num_tensors, 256, 0, st>>>(block_partials, tensor_blk_lo, tensor_blk_cnt, tensor_decay,
                                                      tensor_lars, trust, eta, wd, eps);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_lars_trust(const float* norms, const float* tensor_decay, const float* tensor_lars, float* trust,
                    int32_t num_tensors, float eta, float wd, float eps, void* stream) {
  B200_CHECK_ARG(norms && tensor_decay && tensor_lars && trust && num_tensors > 0, "lars_trust: bad args");
  lars_trust_kernel<<<(num_tensors + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      norms, tensor_decay, tensor_lars, trust, num_tensors, eta, wd, eps);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
