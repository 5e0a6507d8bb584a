// norm.cu — training-mode BatchNorm (+ReLU, +residual) forward/backward and bias(+ReLU) for NHWC bf16.
//
// Role in the reference stack: MKL-DNN's fused batch-normalisation and eltwise primitives that TensorFlow
// calls under --mkl=TRUE (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76); K4/K5/K6 in SURVEY.md §2.6.
//
// Column reductions (statistics, dgamma/dbeta, dbias) share one skeleton: a block covers 64 channels (8 threads x
// one 16-byte vector) x 32 rows per trip, four rows in flight per thread, per-channel constants held in registers;
// partial sums go to caller-provided scratch and the last block of each channel chunk (ticket counter in the
// scratch header) folds them in a fixed order, so results are bitwise reproducible run to run.
// Elementwise passes keep each thread on a fixed channel vector (grid stride is a multiple of C/8 whenever C/8
// divides 256), so scale/shift and the backward coefficients are loaded once per thread, not once per element.
#include "common.cuh"

#include <cstdlib>

namespace b200 {
namespace {

constexpr int kThreads = 256;
constexpr int kTargetBlocks = 8 * kNumSMs;   // elementwise passes: ~8 CTAs per SM
// Column reductions run as one wave of kOcc CTAs per SM (a short final fold). kOcc = 2 lets a thread keep 128 registers:
// the BN backward reduce holds the loads of four rows (8 vector + 4 mask loads) in flight and needs 126. With 3 CTAs per
// SM (85 registers) that batch spills; B200_REDUCE_OCC=3 selects it anyway for A/B runs (measured slower).
inline int reduce_occ() {
  static const int occ = [] {
    const char* e = getenv("B200_REDUCE_OCC");
    return (e && e[0] == '3') ? 3 : 2;
  }();
  return occ;
}
constexpr int kScratchHeader = 64;           // floats reserved in front of the partials (one ticket per channel chunk)

struct ColGeom {
  int tx;       // threads across channels (each 8 channels); 8 = 64 channels per block
  int ty;       // rows per block trip
  int cvecs;    // c / 8
  int ychunks;  // grid.y: channel chunks of tx vectors
  int blocks;   // grid.x
};

inline ColGeom col_geom(int64_t rows, int c) {
  ColGeom g;
  g.cvecs = c / 8;
  g.tx = g.cvecs < 8 ? g.cvecs : 8;
  g.ty = kThreads / g.tx;
  g.ychunks = (g.cvecs + g.tx - 1) / g.tx;
  int64_t b = (rows + (int64_t)g.ty * 8 - 1) / ((int64_t)g.ty * 8);
  int cap = reduce_occ() * kNumSMs / g.ychunks;
  if (cap < 1) cap = 1;
  g.blocks = (int)(b < 1 ? 1 : (b > cap ? cap : b));
  return g;
}

// ---- policies: prepare(cv) caches per-channel constants, load() accumulates one row ---------------------

struct StatsPolicy {
  const b200_bf16* x;
  const float *gamma, *beta;
  float *running_mean, *running_var;
  float decay, eps;
  float *mean, *rstd, *scale, *shift;
  int64_t rows;
  int c;

  static constexpr bool kBatched = false;
  __device__ __forceinline__ void prepare(int) {}
  __device__ __forceinline__ void load(int64_t row, int cv, float a[8], float b[8]) const {
    float f[8];
    unpack8(ld8_stream(x + row * c + cv * 8), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] += f[i];
      b[i] = fmaf(f[i], f[i], b[i]);
    }
  }
  __device__ __forceinline__ void finalize(int ch, float s1, float s2) const { bn_finish(ch, s1, s2); }
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
  const uint8_t* mask;     // optional: one byte per 8-channel vector, bit k = (output k of the forward ReLU was > 0)
  const float *mean, *rstd, *scale, *shift;
  int relu;
  float *dgamma, *dbeta;
  int c;
  float k_scale[8], k_shift[8];

  __device__ __forceinline__ void prepare(int cv) {
    const int ch = cv * 8;
    if (relu && !z && !mask) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        k_scale[i] = scale[ch + i];
        k_shift[i] = shift[ch + i];
      }
    }
  }
  // accumulates a = sum g and b = sum g*x; dgamma = rstd * (b - mean * a) is formed once per channel in finalize.
  // Split into fetch (loads only, no data-dependent branch) and accum so that the skeleton can put the loads of several
  // rows in flight before the first use: with the mode branches between the loads the compiler issued two loads per
  // row and waited, and the kernel ran at 58 % of copy bandwidth on load latency.
  static constexpr bool kBatched = true;
  static constexpr int kRows = 4;
  struct Row {
    bf16x8 g, x;
    unsigned m;
    int64_t off;     // only read in the (rare) mask-from-z mode, which loads z at accumulation time
  };
  __device__ __forceinline__ Row fetch(int64_t row, int cv) const {
    Row r;
    const int64_t off = row * c + cv * 8;
    r.g = ld8_stream(dz + off);
    r.x = ld8_stream(x + off);
    r.m = 0xffu;
    r.off = off;
    if (relu && mask) r.m = __ldg(mask + (off >> 3));
    return r;
  }
  __device__ __forceinline__ void accum(const Row& r, float a[8], float b[8]) const {
    float g[8], xv[8];
    unpack8(r.g, g);
    unpack8(r.x, xv);
    if (relu) {
      if (mask) {
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = (r.m >> i) & 1u ? g[i] : 0.0f;
      } else if (z) {
        float zv[8];
        unpack8(ld8_stream(z + r.off), zv);
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = zv[i] > 0.0f ? g[i] : 0.0f;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = fmaf(k_scale[i], xv[i], k_shift[i]) > 0.0f ? g[i] : 0.0f;
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] += g[i];
      b[i] = fmaf(g[i], xv[i], b[i]);
    }
  }
  __device__ __forceinline__ void load(int64_t row, int cv, float a[8], float b[8]) const { accum(fetch(row, cv), a, b); }
  __device__ __forceinline__ void finalize(int ch, float s1, float s2) const {
    dbeta[ch] = s1;
    dgamma[ch] = rstd[ch] * (s2 - mean[ch] * s1);
  }
};

struct BiasBwdPolicy {
  const b200_bf16 *dz, *z;
  b200_bf16* dx;
  float* dbias;
  int relu;
  int c;

  static constexpr bool kBatched = false;
  __device__ __forceinline__ void prepare(int) {}
  __device__ __forceinline__ void load(int64_t row, int cv, float a[8], float b[8]) const {
    float g[8];
    const int64_t off = row * c + cv * 8;
    unpack8(ld8_stream(dz + off), g);
    if (relu) {
      float zv[8];
      unpack8(ld8_stream(z + off), zv);
#pragma unroll
      for (int i = 0; i < 8; ++i) g[i] = zv[i] > 0.0f ? g[i] : 0.0f;
      if (dx) st8(dx + off, pack8(g));
    } else if (dx && dx != dz) {
      st8(dx + off, pack8(g));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] += g[i];
    (void)b;
  }
  __device__ __forceinline__ void finalize(int ch, float s1, float) const { dbias[ch] = s1; }
};

// ---- skeleton -------------------------------------------------------------------------------------

template <class Policy, int kOcc>
__global__ void __launch_bounds__(kThreads, kOcc) col_reduce_kernel(Policy pol, int64_t rows, int c, ColGeom g,
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
    pol.prepare(cv);
    const int64_t stride = (int64_t)gridDim.x * g.ty;
    int64_t r = (int64_t)blockIdx.x * g.ty + ty;
    if constexpr (Policy::kBatched) {
      // kRows rows per trip, every load of the trip issued before the first use
      constexpr int U = Policy::kRows;
      for (; r + (U - 1) * stride < rows; r += U * stride) {
        typename Policy::Row rw[U];
#pragma unroll
        for (int u = 0; u < U; ++u) rw[u] = pol.fetch(r + u * stride, cv);
#pragma unroll
        for (int u = 0; u < U; ++u) pol.accum(rw[u], a, b);
      }
    } else {
      // eight rows per trip: eight independent groups of 16-byte loads in flight per thread
      for (; r + 7 * stride < rows; r += 8 * stride) {
#pragma unroll
        for (int u = 0; u < 8; ++u) pol.load(r + u * stride, cv, a, b);
      }
    }
    for (; r < rows; r += stride) pol.load(r, cv, a, b);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sm[(i)*kThreads + tid] = a[i];
    sm[(8 + i) * kThreads + tid] = b[i];
  }
  __syncthreads();
  // fold the ty row-lanes: thread j < 16*tx sums column (j % tx) of quantity (j / tx)
  const int per_block = g.tx * 16;
  float* partial = scratch + kScratchHeader + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * (size_t)per_block;
  if (tid < per_block) {
    const int i = tid / g.tx, x = tid % g.tx;
    float s = 0.0f;
    for (int y = 0; y < g.ty; ++y) s += sm[i * kThreads + y * g.tx + x];
    partial[i * g.tx + x] = s;
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
  // last block of this channel chunk: fold the per-block partials. per_block (<= 128) outputs, 256 threads:
  // `seg` threads share one output, each summing a contiguous range of blocks; ranges are combined in order.
  const float* base = scratch + kScratchHeader + (size_t)blockIdx.y * gridDim.x * (size_t)per_block;
  const int seg = kThreads / per_block;              // >= 2
  const int j = tid % per_block, sg = tid / per_block;
  float acc = 0.0f;
  if (sg < seg) {
    const int nb = (int)gridDim.x;
    const int lo = (int)((int64_t)nb * sg / seg), hi = (int)((int64_t)nb * (sg + 1) / seg);
    int bk = lo;
    // sixteen loads in flight per trip: this tail runs on one CTA while the rest of the GPU idles, so its length is
    // the number of L2 round trips, not the bytes
    for (; bk + 16 <= hi; bk += 16) {
      float v[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) v[u] = __ldcg(base + (size_t)(bk + u) * per_block + j);
#pragma unroll
      for (int u = 0; u < 16; ++u) acc += v[u];
    }
    for (; bk + 4 <= hi; bk += 4) {
      const float v0 = __ldcg(base + (size_t)bk * per_block + j);
      const float v1 = __ldcg(base + (size_t)(bk + 1) * per_block + j);
      const float v2 = __ldcg(base + (size_t)(bk + 2) * per_block + j);
      const float v3 = __ldcg(base + (size_t)(bk + 3) * per_block + j);
      acc += v0;
      acc += v1;
      acc += v2;
      acc += v3;
    }
    for (; bk < hi; ++bk) acc += __ldcg(base + (size_t)bk * per_block + j);
  }
  __syncthreads();
  sm[tid] = acc;
  __syncthreads();
  if (tid < g.tx * 8) {
    const int i = tid / g.tx, x = tid % g.tx;      // quantity-1 element i of vector x; quantity 2 is at i + 8
    const int cvv = blockIdx.y * g.tx + x;
    if (cvv < g.cvecs) {
      float s1 = 0.0f, s2 = 0.0f;
      for (int q = 0; q < seg; ++q) {
        s1 += sm[q * per_block + i * g.tx + x];
        s2 += sm[q * per_block + (8 + i) * g.tx + x];
      }
      pol.finalize(cvv * 8 + i, s1, s2);
    }
  }
  if (tid == 0) *ticket = 0u;   // leave the counter armed for the next call
}

template <class Policy>
int launch_col_reduce(const Policy& pol, int64_t rows, int c, void* scratch, cudaStream_t st) {
  ColGeom g = col_geom(rows, c);
  B200_CHECK_ARG(g.ychunks <= kScratchHeader, "column reduce: too many channels (c=%d)", c);
  dim3 grid(g.blocks, g.ychunks);
  if (reduce_occ() == 3)
    col_reduce_kernel<Policy, 3><<<grid, kThreads, 0, st>>>(pol, rows, c, g, static_cast<float*>(scratch));
  else
    col_reduce_kernel<Policy, 2><<<grid, kThreads, 0, st>>>(pol, rows, c, g, static_cast<float*>(scratch));
  B200_LAUNCH_CHECK();
  return 0;
}

// ---- elementwise ----------------------------------------------------------------------------------
// Vector index v = row * cvecs + cv. When `fixed` (grid stride % cvecs == 0) a thread's cv never changes.

struct ElemGrid {
  int blocks;
  int fixed;
};

inline ElemGrid elem_grid(int64_t nvec, int cvecs) {
  ElemGrid e;
  int64_t b = (nvec + 256 * 4 - 1) / (256 * 4);
  if (b > kTargetBlocks) b = kTargetBlocks;
  if (b < 1) b = 1;
  e.fixed = (256 % cvecs == 0) ? 1 : 0;
  if (!e.fixed) {
    // make the stride a multiple of cvecs anyway when a nearby block count allows it
    for (int64_t t = b; t > 0 && t > b - cvecs; --t)
      if ((t * 256) % cvecs == 0) {
        b = t;
        e.fixed = 1;
        break;
      }
  }
  e.blocks = (int)b;
  return e;
}

__device__ __forceinline__ void load8f(const float* p, float o[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}

__global__ void __launch_bounds__(256) bn_apply_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ res,
                                                       b200_bf16* __restrict__ z, uint8_t* __restrict__ mask_out,
                                                       int64_t nvec, int cvecs,
                                                       const float* __restrict__ scale, const float* __restrict__ shift,
                                                       int relu, int fixed) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float sc[8], sh[8];
  const float hi = relu == 2 ? 6.0f : __int_as_float(0x7f800000);   // relu == 2: ReLU6 (the mask records 0 < f < 6)
  if (fixed && i < nvec) {
    const int ch = (int)(i % cvecs) * 8;
    load8f(scale + ch, sc);
    load8f(shift + ch, sh);
  }
  for (; i < nvec; i += 2 * stride) {
    const bool two = i + stride < nvec;
    bf16x8 v0 = ld8_stream(x + i * 8), v1, r0, r1;
    if (two) v1 = ld8_stream(x + (i + stride) * 8);
    if (res) {
      r0 = ld8_stream(res + i * 8);
      if (two) r1 = ld8_stream(res + (i + stride) * 8);
    }
#pragma unroll
    for (int k2 = 0; k2 < 2; ++k2) {
      if (k2 == 1 && !two) break;
      const int64_t idx = k2 ? i + stride : i;
      if (!fixed) {
        const int ch = (int)(idx % cvecs) * 8;
        load8f(scale + ch, sc);
        load8f(shift + ch, sh);
      }
      float f[8];
      unpack8(k2 ? v1 : v0, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]);
      if (res) {
        float r[8];
        unpack8(k2 ? r1 : r0, r);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] += r[k];
      }
      if (relu) {
        if (mask_out) {
          unsigned mb = 0;
#pragma unroll
          for (int k = 0; k < 8; ++k) mb |= (f[k] > 0.0f && f[k] < hi ? 1u : 0u) << k;
          mask_out[idx] = (uint8_t)mb;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fminf(fmaxf(f[k], 0.0f), hi);
      }
      st8(z + idx * 8, pack8(f));
    }
  }
}

// dx = A*g + B*x + C with g = dz masked by the ReLU; A = scale, B = -scale*dgamma*rstd/M, C = -scale*dbeta/M - B*mean
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const b200_bf16* __restrict__ dz, const b200_bf16* __restrict__ x,
                                                           const b200_bf16* __restrict__ z, const uint8_t* __restrict__ mask,
                                                           b200_bf16* __restrict__ dx,
                                                           b200_bf16* __restrict__ dres, int64_t nvec, int cvecs,
                                                           float inv_rows, const float* __restrict__ mean,
                                                           const float* __restrict__ rstd, const float* __restrict__ scale,
                                                           const float* __restrict__ shift, const float* __restrict__ dgamma,
                                                           const float* __restrict__ dbeta, int relu, int fixed) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float cA[8], cB[8], cC[8], cS[8];
  auto coeffs = [&](int ch) {
    float m[8], r[8], dg[8], db[8];
    load8f(scale + ch, cA);
    load8f(shift + ch, cS);
    load8f(mean + ch, m);
    load8f(rstd + ch, r);
    load8f(dgamma + ch, dg);
    load8f(dbeta + ch, db);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      cB[k] = -cA[k] * dg[k] * r[k] * inv_rows;
      cC[k] = -cA[k] * db[k] * inv_rows - cB[k] * m[k];
    }
  };
  if (fixed && i < nvec) coeffs((int)(i % cvecs) * 8);
  for (; i < nvec; i += 2 * stride) {
    const bool two = i + stride < nvec;
    bf16x8 g0 = ld8_stream(dz + i * 8), x0 = ld8_stream(x + i * 8), g1, x1, z0, z1;
    if (two) {
      g1 = ld8_stream(dz + (i + stride) * 8);
      x1 = ld8_stream(x + (i + stride) * 8);
    }
    unsigned m0 = 0, m1 = 0;
    if (relu && mask) {
      m0 = __ldg(mask + i);
      if (two) m1 = __ldg(mask + i + stride);
    } else if (relu && z) {
      z0 = ld8_stream(z + i * 8);
      if (two) z1 = ld8_stream(z + (i + stride) * 8);
    }
#pragma unroll
    for (int k2 = 0; k2 < 2; ++k2) {
      if (k2 == 1 && !two) break;
      const int64_t idx = k2 ? i + stride : i;
      if (!fixed) coeffs((int)(idx % cvecs) * 8);
      float g[8], xv[8];
      unpack8(k2 ? g1 : g0, g);
      unpack8(k2 ? x1 : x0, xv);
      if (relu) {
        if (mask) {
          const unsigned mb = k2 ? m1 : m0;
#pragma unroll
          for (int k = 0; k < 8; ++k) g[k] = (mb >> k) & 1u ? g[k] : 0.0f;
        } else if (z) {
          float zv[8];
          unpack8(k2 ? z1 : z0, zv);
#pragma unroll
          for (int k = 0; k < 8; ++k) g[k] = zv[k] > 0.0f ? g[k] : 0.0f;
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) g[k] = fmaf(cA[k], xv[k], cS[k]) > 0.0f ? g[k] : 0.0f;
        }
      }
      if (dres) st8(dres + idx * 8, pack8(g));
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = fmaf(cA[k], g[k], fmaf(cB[k], xv[k], cC[k]));
      st8(dx + idx * 8, pack8(o));
    }
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

// Fold [nrows][2][c] partial sums (conv epilogue) in a fixed order, then finalise. Block = 32 channels x 32 row slices:
// a slice is a handful of rows (nrows is one or two per SM), all of whose loads are issued before the first add - the
// kernel is a chain of L2 round trips, not bandwidth, so its length is the number of dependent trips (it was twenty with
// 8 slices and a rolled loop: 11 us per BN layer, 53 layers per step).
constexpr int kFinSlices = 32;
__global__ void __launch_bounds__(32 * kFinSlices) bn_finalize_partials_kernel(StatsPolicy pol, const float* __restrict__ part,
                                                                                int64_t nrows, int c) {
  __shared__ float sm[2][kFinSlices][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int ch = blockIdx.x * 32 + tx;
  float s1 = 0.0f, s2 = 0.0f;
  if (ch < c) {
    // contiguous row ranges per slice keep the summation order fixed
    const int64_t lo = nrows * ty / kFinSlices, hi = nrows * (ty + 1) / kFinSlices;
    int64_t r = lo;
    for (; r + 4 <= hi; r += 4) {
      float v1[4], v2[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float* p = part + (r + u) * 2 * c;
        v1[u] = __ldcg(p + ch);
        v2[u] = __ldcg(p + c + ch);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        s1 += v1[u];
        s2 += v2[u];
      }
    }
    for (; r < hi; ++r) {
      const float* p = part + r * 2 * c;
      s1 += __ldcg(p + ch);
      s2 += __ldcg(p + c + ch);
    }
  }
  sm[0][ty][tx] = s1;
  sm[1][ty][tx] = s2;
  __syncthreads();
  if (ty == 0 && ch < c) {
    float a = 0.0f, b = 0.0f;
#pragma unroll
    for (int y = 0; y < kFinSlices; ++y) {
      a += sm[0][y][tx];
      b += sm[1][y][tx];
    }
    pol.bn_finish(ch, a, b);
  }
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

// o = a + (mask bit ? b : 0): the gradient reaching a residual block's input is the main-path gradient plus the
// block-output gradient gated by the output ReLU, whose mask bn_apply saved as one bit per element (one byte per
// 8-channel vector). Doing the gate here means the gated copy is never written to memory.
__global__ void __launch_bounds__(256) add_masked_kernel(const b200_bf16* __restrict__ a, const b200_bf16* __restrict__ b,
                                                         const uint8_t* __restrict__ mask, b200_bf16* __restrict__ o,
                                                         int64_t nvec) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    float x[8], y[8];
    unpack8(ld8_stream(a + i * 8), x);
    unpack8(ld8_stream(b + i * 8), y);
    const unsigned mb = __ldg(mask + i);
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] += (mb >> k) & 1u ? y[k] : 0.0f;
    st8(o + i * 8, pack8(x));
  }
}

// o = relu(a + b): the standalone residual sum of the CIFAR ResNets and pre-activation models (blocks whose sum is not
// behind a batch-norm, so it cannot ride in bn_apply); relu == 0 gives the plain sum.
__global__ void __launch_bounds__(256) add_act_kernel(const b200_bf16* __restrict__ a, const b200_bf16* __restrict__ b,
                                                      b200_bf16* __restrict__ o, int64_t nvec, int relu) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    float x[8], y[8];
    unpack8(ld8_stream(a + i * 8), x);
    unpack8(ld8_stream(b + i * 8), y);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      x[k] += y[k];
      if (relu) x[k] = fmaxf(x[k], 0.0f);
    }
    st8(o + i * 8, pack8(x));
  }
}

// dx = z > 0 ? dz : 0 (backward of the ReLU above, gated by its output)
__global__ void __launch_bounds__(256) relu_bwd_kernel(const b200_bf16* __restrict__ dz, const b200_bf16* __restrict__ z,
                                                       b200_bf16* __restrict__ dx, int64_t nvec) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < nvec; i += stride) {
    float g[8], v[8];
    unpack8(ld8_stream(dz + i * 8), g);
    unpack8(ld8_stream(z + i * 8), v);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = v[k] > 0.0f ? g[k] : 0.0f;
    st8(dx + i * 8, pack8(g));
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

int b200_bn_finalize_partials(const float* partials, int64_t nrows, int64_t rows, int32_t c, const float* gamma,
                              const float* beta, float* running_mean, float* running_var, float decay, float eps,
                              float* mean, float* rstd, float* scale, float* shift, void* stream) {
  B200_CHECK_ARG(partials && mean && rstd && scale && shift && nrows > 0, "bn_finalize_partials: bad args");
  StatsPolicy pol{nullptr, gamma, beta, running_mean, running_var, decay, eps, mean, rstd, scale, shift, rows, c};
  bn_finalize_partials_kernel<<<(c + 31) / 32, 32 * kFinSlices, 0, static_cast<cudaStream_t>(stream)>>>(pol, partials, nrows, c);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_bn_apply(const b200_bf16* x, const b200_bf16* residual, b200_bf16* z, uint8_t* relu_mask, int64_t rows,
                  int32_t c, const float* scale, const float* shift, int32_t relu, void* stream) {
  B200_CHECK_ARG(x && z && scale && shift, "bn_apply: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_apply: need c%%8==0 (c=%d)", c);
  int64_t nvec = rows * (c / 8);
  ElemGrid e = elem_grid(nvec, c / 8);
  bn_apply_kernel<<<e.blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, residual, z, relu_mask, nvec, c / 8, scale,
                                                                          shift, relu, e.fixed);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_bn_bwd_reduce(const b200_bf16* dz, const b200_bf16* x, const b200_bf16* z, const uint8_t* relu_mask,
                       int64_t rows, int32_t c, const float* mean, const float* rstd, const float* scale, const float* shift, int32_t relu,
                       float* dgamma, float* dbeta, void* scratch, void* stream) {
  B200_CHECK_ARG(dz && x && mean && rstd && scale && shift && dgamma && dbeta && scratch, "bn_bwd_reduce: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_bwd_reduce: need c%%8==0 (c=%d)", c);
  BwdReducePolicy pol{dz, x, z, relu_mask, mean, rstd, scale, shift, relu, dgamma, dbeta, c};
  return launch_col_reduce(pol, rows, c, scratch, static_cast<cudaStream_t>(stream));
}

int b200_bn_bwd_apply(const b200_bf16* dz, const b200_bf16* x, const b200_bf16* z, const uint8_t* relu_mask,
                      b200_bf16* dx, b200_bf16* dres, int64_t rows, int32_t c, const float* mean, const float* rstd, const float* scale,
                      const float* shift, const float* dgamma, const float* dbeta, int32_t relu, void* stream) {
  B200_CHECK_ARG(dz && x && dx && mean && rstd && scale && shift && dgamma && dbeta, "bn_bwd_apply: null pointer");
  B200_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "bn_bwd_apply: need c%%8==0 (c=%d)", c);
  int64_t nvec = rows * (c / 8);
  ElemGrid e = elem_grid(nvec, c / 8);
  bn_bwd_apply_kernel<<<e.blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      dz, x, z, relu_mask, dx, dres, nvec, c / 8, 1.0f / (float)rows, mean, rstd, scale, shift, dgamma, dbeta, relu, e.fixed);
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

int b200_add_masked(const b200_bf16* a, const b200_bf16* b, const uint8_t* mask, b200_bf16* out, int64_t n,
                    void* stream) {
  B200_CHECK_ARG(a && b && mask && out && n % 8 == 0, "add_masked: need n%%8==0");
  add_masked_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, b, mask, out, n / 8);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_add_act(const b200_bf16* a, const b200_bf16* b, b200_bf16* out, int64_t n, int32_t relu, void* stream) {
  B200_CHECK_ARG(a && b && out && n > 0 && n % 8 == 0, "add_act: need n%%8==0");
  add_act_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, b, out, n / 8, relu);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_relu_bwd(const b200_bf16* dz, const b200_bf16* z, b200_bf16* dx, int64_t n, void* stream) {
  B200_CHECK_ARG(dz && z && dx && n > 0 && n % 8 == 0, "relu_bwd: need n%%8==0");
  relu_bwd_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dz, z, dx, n / 8);
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
