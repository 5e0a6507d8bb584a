// norm.cu — training-mode BatchNorm (+ReLU, +residual) forward/backward and bias(+ReLU) for NHWC bf16.
//
// Role in the reference stack: MKL-DNN's fused batch-normalisation and eltwise primitives that TensorFlow
// calls under --mkl=TRUE (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76); K4/K5/K6 in SURVEY.md §2.6.
//
// All reductions over rows share one column-reduce skeleton: a block owns TY rows x all channels per
// iteration (each thread 8 channels = one 16-byte load), accumulates in fp32 registers, folds TY in shared
// memory, writes one partial per block into caller-provided scratch, and the last block to finish (ticket
// counter in scratch[0]) folds the partials in a fixed order and runs the policy's finalisation. Results are
// therefore bitwise reproducible run to run.
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kThreads = 256;
constexpr int kMaxBlocks = 2 * kNumSMs;
constexpr int kScratchHeader = 64;  // floats reserved in front of the partials (ticket lives at [0])

struct ColGeom {
  int tx;       // threads across channels (each 8 channels)
  int ty;       // rows per block iteration
  int cvecs;    // c / 8
  int ychunks;  // grid.y: channel chunks of tx vectors
  int blocks;   // grid.x
};

inline ColGeom col_geom(int64_t rows, int c) {
  ColGeom g;
  g.cvecs = c / 8;
  g.tx = g.cvecs < kThreads ? g.cvecs : kThreads;
  g.ty = kThreads / g.tx;
  g.ychunks = (g.cvecs + g.tx - 1) / g.tx;
  int64_t b = (rows + (int64_t)g.ty * 8 - 1) / ((int64_t)g.ty * 8);
  int cap = kMaxBlocks / g.ychunks;
  if (cap < 1) cap = 1;
  g.blocks = (int)(b < 1 ? 1 : (b > cap ? cap : b));
  return g;
}

// ---- policies -------------------------------------------------------------------------------------

struct StatsPolicy {
  const b200_bf16* x;
  const float *gamma, *beta;
  float *running_mean, *running_var;
  float decay, eps;
  float *mean, *rstd, *scale, *shift;
  int64_t rows;
  int c;

  __device__ __forceinline__ void load(int64_t row, int cv, float a[8], float b[8]) const {
    float f[8];
    unpack8(ld8_stream(x + row * c + cv * 8), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] += f[i];
      b[i] += f[i] * f[i];
    }
  }
  __device__ __forceinline__ void finalize(int ch, float s1, float s2) const {
    bn_finish(ch, s1, s2);
  }
  __device__ __forceinline__ void bn_finish(int ch, float s1, float s2) const {
    float inv = 1.0f / (float)rows;
    float m = s1 * inv;
    float var = fmaxf(s2 * inv - m * m, 0.0f);
    float r = rsqrtf(var + eps);
    float g = gamma ? gamma[ch] : 1.0f;
    float sc = g * r;
    mean[ch] = m;
    rstd[ch] = r;
    scale[ch] = sc;
    shift[ch] = (beta ? beta[ch] : 0.0f) - m * sc;
    if (running_mean) {
      float unbiased = rows > 1 ? var * ((float)rows / (float)(rows - 1)) : var;
      running_mean[ch] = decay * running_mean[ch] + (1.0f - decay) * m;
      running_var[ch] = decay * running_var[ch] + (1.0f - decay) * unbiased;
    }
  }
};

struct BwdReducePolicy {
  const b200_bf16 *dz, *x, *z;
  const float *mean, *rstd, *scale, *shift;
  int relu;
  float *dgamma, *dbeta;
  int c;

  __device__ __forceinline__ void load(int64_t row, int cv, float a[8], float b[8]) const {
    float g[8], xv[8];
    int ch = cv * 8;
    unpack8(ld8_stream(dz + row * c + ch), g);
    unpack8(ld8_stream(x + row * c + ch), xv);
    if (relu) {
      if (z) {
        float zv[8];
        unpack8(ld8_stream(z + row * c + ch), zv);
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = zv[i] > 0.0f ? g[i] : 0.0f;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = (scale[ch + i] * xv[i] + shift[ch + i]) > 0.0f ? g[i] : 0.0f;
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] += g[i];                                              // dbeta
      b[i] += g[i] * (xv[i] - mean[ch + i]) * rstd[ch + i];      // dgamma
    }
  }
  __device__ __forceinline__ void finalize(int ch, float s1, float s2) const {
    dbeta[ch] = s1;
    dgamma[ch] = s2;
  }
};

struct BiasBwdPolicy {
  const b200_bf16 *dz, *z;
  b200_bf16* dx;
  float* dbias;
  int relu;
  int c;

  __device__ __forceinline__ void load(int64_t row, int cv, float a[8], float b[8]) const {
    float g[8];
    int ch = cv * 8;
    unpack8(ld8_stream(dz + row * c + ch), g);
    if (relu) {
      float zv[8];
      unpack8(ld8_stream(z + row * c + ch), zv);
#pragma unroll
      for (int i = 0; i < 8; ++i) g[i] = zv[i] > 0.0f ? g[i] : 0.0f;
      if (dx) st8(dx + row * c + ch, pack8(g));
    } else if (dx && dx != dz) {
      st8(dx + row * c + ch, pack8(g));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] += g[i];
  }
  __device__ __forceinline__ void finalize(int ch, float s1, float) const { dbias[ch] = s1; }
};

// ---- skeleton -------------------------------------------------------------------------------------

template <class Policy>
__global__ void __launch_bounds__(kThreads) col_reduce_kernel(Policy pol, int64_t rows, int c, ColGeom g,
                                                              float* scratch) {
  __shared__ float sm[kThreads * 16];
  __shared__ int is_last;
  const int tid = threadIdx.x;
  const int tx = tid % g.tx, ty = tid / g.tx;
  const int cv = blockIdx.y * g.tx + tx;
  const bool active = ty < g.ty && cv < g.cvecs;
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = b[i] = 0.0f;
  if (active) {
    const int64_t stride = (int64_t)gridDim.x * g.ty;
    int64_t r = (int64_t)blockIdx.x * g.ty + ty;
    // two rows per trip keeps two independent 16-byte loads per operand in flight
    for (; r + stride < rows; r += 2 * stride) {
      pol.load(r, cv, a, b);
      pol.load(r + stride, cv, a, b);
    }
    if (r < rows) pol.load(r, cv, a, b);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sm[(i)*kThreads + tid] = a[i];
    sm[(8 + i) * kThreads + tid] = b[i];
  }
  __syncthreads();
  float* partial = scratch + kScratchHeader + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * (size_t)(g.tx * 16);
  if (ty == 0 && cv < g.cvecs) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float s = 0.0f;
      for (int y = 0; y < g.ty; ++y) s += sm[i * kThreads + y * g.tx + tx];
      partial[i * g.tx + tx] = s;
    }
  }
  __threadfence();
  __syncthreads();
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch) + blockIdx.y;
  if (tid == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // last block of this channel chunk: fold the per-block partials in block order
  const float* base = scratch + kScratchHeader + (size_t)blockIdx.y * gridDim.x * (size_t)(g.tx * 16);
  for (int j = tid; j < g.tx * 8; j += kThreads) {
    int i = j / g.tx, x = j % g.tx;   // element i of vector x
    int cvv = blockIdx.y * g.tx + x;
    if (cvv >= g.cvecs) continue;
    float s1 = 0.0f, s2 = 0.0f;
    for (unsigned int bk = 0; bk < gridDim.x; ++bk) {
      const float* p = base + (size_t)bk * (g.tx * 16);
      s1 += __ldcg(p + i * g.tx + x);
      s2 += __ldcg(p + (8 + i) * g.tx + x);
    }
    pol.finalize(cvv * 8 + i, s1, s2);
  }
  if (tid == 0) *ticket = 0u;   // leave the counter armed for the next call
}

template <class Policy>
int launch_col_reduce(const Policy& pol, int64_t rows, int c, void* scratch, cudaStream_t st) {
  ColGeom g = col_geom(rows, c);
  dim3 grid(g.blocks, g.ychunks);
  col_reduce_kernel<Policy><<<grid, kThreads, 0, st>>>(pol, rows, c, g, static_cast<float*>(scratch));
  B200_LAUNCH_CHECK();
  return 0;
}

// ---- elementwise ----------------------------------------------------------------------------------

__global__ void __launch_bounds__(256) bn_apply_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ res,
                                                       b200_bf16* __restrict__ z, int64_t nvec, int cvecs,
                                                       const float* __restrict__ scale, const float* __restrict__ shift,
                                                       int relu) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    int ch = (int)(i % cvecs) * 8;
    float f[8];
    unpack8(ld8_stream(x + i * 8), f);
    float4 s0 = *reinterpret_cast<const float4*>(scale + ch), s1 = *reinterpret_cast<const float4*>(scale + ch + 4);
    float4 h0 = *reinterpret_cast<const float4*>(shift + ch), h1 = *reinterpret_cast<const float4*>(shift + ch + 4);
    float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]);
    if (res) {
      float r[8];
      unpack8(ld8_stream(res + i * 8), r);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += r[k];
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.0f);
    }
    st8(z + i * 8, pack8(f));
  }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const b200_bf16* __restrict__ dz, const b200_bf16* __restrict__ x,
                                                           const b200_bf16* __restrict__ z, b200_bf16* __restrict__ dx,
                                                           b200_bf16* __restrict__ dres, int64_t nvec, int cvecs,
                                                           float inv_rows, const float* __restrict__ mean,
                                                           const float* __restrict__ rstd, const float* __restrict__ scale,
                                                           const float* __restrict__ shift, const float* __restrict__ dgamma,
                                                           const float* __restrict__ dbeta, int relu) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    int ch = (int)(i % cvecs) * 8;
    float g[8], xv[8];
    unpack8(ld8_stream(dz + i * 8), g);
    unpack8(ld8_stream(x + i * 8), xv);
    if (relu) {
      if (z) {
        float zv[8];
        unpack8(ld8_stream(z + i * 8), zv);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = zv[k] > 0.0f ? g[k] : 0.0f;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = fmaf(scale[ch + k], xv[k], shift[ch + k]) > 0.0f ? g[k] : 0.0f;
      }
    }
    if (dres) st8(dres + i * 8, pack8(g));
    float o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float xhat = (xv[k] - mean[ch + k]) * rstd[ch + k];
      o[k] = scale[ch + k] * (g[k] - dbeta[ch + k] * inv_rows - xhat * dgamma[ch + k] * inv_rows);
    }
    st8(dx + i * 8, pack8(o));
  }
}

__global__ void __launch_bounds__(256) bias_act_kernel(const b200_bf16* __restrict__ x, const float* __restrict__ bias,
                                                       b200_bf16* __restrict__ z, int64_t nvec, int cvecs, int relu) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    int ch = (int)(i % cvecs) * 8;
    float f[8];
    unpack8(ld8_stream(x + i * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      f[k] += bias ? bias[ch + k] : 0.0f;
      if (relu) f[k] = fmaxf(f[k], 0.0f);
    }
    st8(z + i * 8, pack8(f));
  }
}

__global__ void bn_finalize_kernel(StatsPolicy pol, const float* sums, int c) {
  int ch = blockIdx.x * blockDim.x + threadIdx.x;
  if (ch < c) pol.bn_finish(ch, sums[ch], sums[c + ch]);
}

__global__ void __launch_bounds__(256) add_kernel(const b200_bf16* __restrict__ a, const b200_bf16* __restrict__ b,
                                                  b200_bf16* __restrict__ o, int64_t nvec) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    float x[8], y[8];
    unpack8(ld8_stream(a + i * 8), x);
    unpack8(ld8_stream(b + i * 8), y);
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] += y[k];
    st8(o + i * 8, pack8(x));
  }
}

__global__ void __launch_bounds__(256) cast_kernel(const float* __restrict__ x, b200_bf16* __restrict__ y, int64_t n) {
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 8;
  for (; i < n; i += stride) {
    if (i + 8 <= n) {
      float4 a = *reinterpret_cast<const float4*>(x + i), b = *reinterpret_cast<const float4*>(x + i + 4);
      float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      st8(y + i, pack8(f));
    } else {
      for (int64_t k = i; k < n; ++k) reinterpret_cast<__nv_bfloat16*>(y)[k] = __float2bfloat16(x[k]);
    }
  }
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int64_t b200_bn_stats_scratch_bytes(int64_t rows, int32_t c) {
  if (c <= 0 || c % 8 != 0) return -1;
  ColGeom g = col_geom(rows, c);
  return (int64_t)sizeof(float) * (kScratchHeader + (int64_t)g.ychunks * g.blocks * g.tx * 16);
}

int b200_bn_stats(const b200_bf16* x, int64_t rows, int32_t c, const float* gamma, const float* beta,
                  float* running_mean, float* running_var, float decay, float eps, float* mean, float* rstd,
                  float* scale, float* shift, void* scratch, void* stream) {
  B200_CHECK_ARG(x && mean && rstd && scale && shift && scratch, "bn_stats: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_stats: need rows>0 and c%%8==0 (c=%d)", c);
  StatsPolicy pol{x, gamma, beta, running_mean, running_var, decay, eps, mean, rstd, scale, shift, rows, c};
  return launch_col_reduce(pol, rows, c, scratch, static_cast<cudaStream_t>(stream));
}

int b200_bn_finalize(const float* sums, int64_t rows, int32_t c, const float* gamma, const float* beta,
                     float* running_mean, float* running_var, float decay, float eps, float* mean, float* rstd,
                     float* scale, float* shift, void* stream) {
  B200_CHECK_ARG(sums && mean && rstd && scale && shift, "bn_finalize: null pointer");
  StatsPolicy pol{nullptr, gamma, beta, running_mean, running_var, decay, eps, mean, rstd, scale, shift, rows, c};
  bn_finalize_kernel<<<(c + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(pol, sums, c);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_bn_apply(const b200_bf16* x, const b200_bf16* residual, b200_bf16* z, int64_t rows, int32_t c,
                  const float* scale, const float* shift, int32_t relu, void* stream) {
  B200_CHECK_ARG(x && z && scale && shift, "bn_apply: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_apply: need c%%8==0 (c=%d)", c);
  int64_t nvec = rows * (c / 8);
  bn_apply_kernel<<<grid_for(nvec, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, residual, z, nvec, c / 8,
                                                                                     scale, shift, relu);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_bn_bwd_reduce(const b200_bf16* dz, const b200_bf16* x, const b200_bf16* z, int64_t rows, int32_t c,
                       const float* mean, const float* rstd, const float* scale, const float* shift, int32_t relu,
                       float* dgamma, float* dbeta, void* scratch, void* stream) {
  B200_CHECK_ARG(dz && x && mean && rstd && scale && shift && dgamma && dbeta && scratch, "bn_bwd_reduce: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_bwd_reduce: need c%%8==0 (c=%d)", c);
  BwdReducePolicy pol{dz, x, z, mean, rstd, scale, shift, relu, dgamma, dbeta, c};
  return launch_col_reduce(pol, rows, c, scratch, static_cast<cudaStream_t>(stream));
}

int b200_bn_bwd_apply(const b200_bf16* dz, const b200_bf16* x, const b200_bf16* z, b200_bf16* dx, b200_bf16* dres,
                      int64_t rows, int32_t c, const float* mean, const float* rstd, const float* scale,
                      const float* shift, const float* dgamma, const float* dbeta, int32_t relu, void* stream) {
  B200_CHECK_ARG(dz && x && dx && mean && rstd && scale && shift && dgamma && dbeta, "bn_bwd_apply: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_bwd_apply: need c%%8==0 (c=%d)", c);
  int64_t nvec = rows * (c / 8);
  bn_bwd_apply_kernel<<<grid_for(nvec, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      dz, x, z, dx, dres, nvec, c / 8, 1.0f / (float)rows, mean, rstd, scale, shift, dgamma, dbeta, relu);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_bias_act(const b200_bf16* x, const float* bias, b200_bf16* z, int64_t rows, int32_t c, int32_t relu,
                  void* stream) {
  B200_CHECK_ARG(x && z, "bias_act: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bias_act: need c%%8==0 (c=%d)", c);
  int64_t nvec = rows * (c / 8);
  bias_act_kernel<<<grid_for(nvec, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, bias, z, nvec, c / 8, relu);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_bias_act_bwd(const b200_bf16* dz, const b200_bf16* z, b200_bf16* dx, float* dbias, int64_t rows,
                      int32_t c, int32_t relu, void* scratch, void* stream) {
  B200_CHECK_ARG(dz && dbias && scratch && (!relu || z), "bias_act_bwd: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bias_act_bwd: need c%%8==0 (c=%d)", c);
  BiasBwdPolicy pol{dz, z, dx, dbias, relu, c};
  return launch_col_reduce(pol, rows, c, scratch, static_cast<cudaStream_t>(stream));
}

int b200_add(const b200_bf16* a, const b200_bf16* b, b200_bf16* out, int64_t n, void* stream) {
  B200_CHECK_ARG(a && b && out && n % 8 == 0, "add: need n%%8==0");
  add_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, b, out, n / 8);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_cast_f32_bf16(const float* x, b200_bf16* y, int64_t n, void* stream) {
  B200_CHECK_ARG(x && y && n > 0, "cast: bad args");
  cast_kernel<<<grid_for((n + 7) / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, n);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of b200_bn_stats.
//
// Some callers are not allowed to obtain device memory: code running inside a framework's memory-planned region
// (every buffer was sized and placed ahead of time), plug-ins that are handed fixed input/output buffers and nothing
// else, or hosts whose allocator must not be re-entered from where this library is called. For them the reduction
// scratch lives in the library image itself: one fixed array per GPU, present from the moment the library is loaded
// on that GPU until the process exits.
namespace b200 {
namespace {

constexpr size_t kStaticScratchBytes = 2u << 20;   // 2 MiB per GPU that loads the library
static __device__ __align__(256) unsigned char g_bn_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
void* g_static_scratch_ptr[kMaxDevices] = {nullptr};

// Resolves the device address of the array for the current device. In host code `g_bn_static_scratch` is only a
// handle for the runtime; the address a kernel can use has to be asked for, per device, with that device current.
int static_scratch_for_current_device(void** out) {
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "bn_stats_noalloc: device index %d out of range", dev);
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_bn_static_scratch));
    g_static_scratch_ptr[dev] = p;
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}

}  // namespace
}  // namespace b200

extern "C" {

int64_t b200_bn_stats_noalloc_max_bytes(void) { return (int64_t)b200::kStaticScratchBytes; }

int b200_bn_stats_noalloc(const b200_bf16* x, int64_t rows, int32_t c, const float* gamma, const float* beta,
                          float* running_mean, float* running_var, float decay, float eps, float* mean, float* rstd,
                          float* scale, float* shift, void* stream) {
  B200_CHECK_ARG(x && mean && rstd && scale && shift, "bn_stats_noalloc: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_stats_noalloc: need rows>0 and c%%8==0 (c=%d)", c);
  const int64_t need = b200_bn_stats_scratch_bytes(rows, c);
  B200_CHECK_ARG(need > 0 && (size_t)need <= kStaticScratchBytes,
                 "bn_stats_noalloc: rows=%lld c=%d needs %lld bytes of scratch, the built-in array holds %lld; use "
                 "b200_bn_stats with caller-provided scratch for this size",
                 (long long)rows, c, (long long)need, (long long)kStaticScratchBytes);
  void* scratch = nullptr;
  if (int e = static_scratch_for_current_device(&scratch)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array keeps whatever the previous call (or an aborted one) left behind. Only the ticket counters at its
  // head have to start at zero; the partial sums behind them are fully overwritten before they are read.
  B200_CUDA(cudaMemsetAsync(scratch, 0, kScratchHeader * sizeof(float), st));
  StatsPolicy pol{x, gamma, beta, running_mean, running_var, decay, eps, mean, rstd, scale, shift, rows, c};
  return launch_col_reduce(pol, rows, c, scratch, st);
}

}  // extern "C"
