// depthwise.cu — depthwise convolution (groups == channels) forward, backward-data and backward-weights, NHWC bf16.
//
// Role in the reference stack: MKL-DNN's depthwise convolution primitive under --mkl=TRUE
// (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76), reached when --model (:66) selects MobileNet. A depthwise layer
// does R*S multiply-adds per output element and reads every input once: it is bandwidth-bound, there is no GEMM in it,
// so these are plain vector kernels (8 channels = one 16-byte vector per thread, filter taps in registers), not
// tensor-core kernels. Filters are [C][R][S] (the KRSC layout of a [C,1,R,S] weight), R*S <= 9.
#include "b200conv.h"
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kMaxTaps = 9;

struct DwGeom {
  int n, h, w, c, r, s, p, q, sh, sw, pt, pl;
};

// thread (tx, ty): tx walks the 8-channel vectors, ty the pixels; a thread's channels are fixed for its lifetime when
// cvecs <= blockDim.x (every MobileNet layer), so the taps are loaded into registers once
__device__ __forceinline__ void load_taps(const b200_bf16* __restrict__ w, int c0, int taps, float wt[kMaxTaps][8]) {
#pragma unroll
  for (int t = 0; t < kMaxTaps; ++t)
#pragma unroll
    for (int k = 0; k < 8; ++k)
      wt[t][k] = t < taps ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(w)[(size_t)(c0 + k) * taps + t]) : 0.0f;
}

__global__ void __launch_bounds__(256) dw_fwd_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ w,
                                                     b200_bf16* __restrict__ y, DwGeom g) {
  const int cvecs = g.c / 8, taps = g.r * g.s;
  const int64_t pixels = (int64_t)g.n * g.p * g.q;
  for (int cv = threadIdx.x; cv < cvecs; cv += blockDim.x) {
    float wt[kMaxTaps][8];
    load_taps(w, cv * 8, taps, wt);
    for (int64_t pix = (int64_t)blockIdx.x * blockDim.y + threadIdx.y; pix < pixels; pix += (int64_t)gridDim.x * blockDim.y) {
      const int q = (int)(pix % g.q);
      const int p = (int)((pix / g.q) % g.p);
      const int n = (int)(pix / ((int64_t)g.q * g.p));
      float acc[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
#pragma unroll
      for (int t = 0; t < kMaxTaps; ++t) {
        if (t >= taps) break;
        const int r = t / g.s, s = t - r * g.s;
        const int ih = p * g.sh - g.pt + r, iw = q * g.sw - g.pl + s;
        if (ih < 0 || ih >= g.h || iw < 0 || iw >= g.w) continue;
        float f[8];
        unpack8(ld8(x + (((size_t)n * g.h + ih) * g.w + iw) * g.c + cv * 8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(f[k], wt[t][k], acc[k]);
      }
      st8(y + (size_t)pix * g.c + cv * 8, pack8(acc));
    }
  }
}

// dx[n,h,w,c] = sum over taps (r,s) with (h + pt - r) % sh == 0 and (w + pl - s) % sw == 0 of dy[n, ., ., c] * w[c,r,s]
__global__ void __launch_bounds__(256) dw_dgrad_kernel(const b200_bf16* __restrict__ dy, const b200_bf16* __restrict__ w,
                                                       b200_bf16* __restrict__ dx, DwGeom g) {
  const int cvecs = g.c / 8, taps = g.r * g.s;
  const int64_t pixels = (int64_t)g.n * g.h * g.w;
  for (int cv = threadIdx.x; cv < cvecs; cv += blockDim.x) {
    float wt[kMaxTaps][8];
    load_taps(w, cv * 8, taps, wt);
    for (int64_t pix = (int64_t)blockIdx.x * blockDim.y + threadIdx.y; pix < pixels; pix += (int64_t)gridDim.x * blockDim.y) {
      const int iw = (int)(pix % g.w);
      const int ih = (int)((pix / g.w) % g.h);
      const int n = (int)(pix / ((int64_t)g.w * g.h));
      float acc[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
#pragma unroll
      for (int t = 0; t < kMaxTaps; ++t) {
        if (t >= taps) break;
        const int r = t / g.s, s = t - r * g.s;
        const int a = ih + g.pt - r, b = iw + g.pl - s;
        if (a < 0 || b < 0 || a % g.sh || b % g.sw) continue;
        const int p = a / g.sh, q = b / g.sw;
        if (p >= g.p || q >= g.q) continue;
        float f[8];
        unpack8(ld8(dy + (((size_t)n * g.p + p) * g.q + q) * g.c + cv * 8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(f[k], wt[t][k], acc[k]);
      }
      st8(dx + (size_t)pix * g.c + cv * 8, pack8(acc));
    }
  }
}

// dw[c,r,s] = sum over output pixels of dy[n,p,q,c] * x[n, p*sh - pt + r, q*sw - pl + s, c]. Each thread owns 8 channels
// and a strided set of pixels and keeps taps x 8 fp32 sums in registers; the block folds its pixel lanes (threadIdx.y)
// through shared memory in lane order and writes one row of `partials` ([gridDim.x][taps][c]); dw_fold_kernel sums the
// rows in index order, so the result is bitwise reproducible.
__global__ void __launch_bounds__(256) dw_wgrad_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ dy,
                                                       float* __restrict__ partials, DwGeom g) {
  __shared__ float red[256 * 8];
  const int cvecs = g.c / 8, taps = g.r * g.s;
  const int64_t pixels = (int64_t)g.n * g.p * g.q;
  const int tx = threadIdx.x, ty = threadIdx.y, nx = blockDim.x, ny = blockDim.y;
  for (int cv0 = 0; cv0 < cvecs; cv0 += nx) {
    const int cv = cv0 + tx;
    const bool active = cv < cvecs;
    float acc[kMaxTaps][8];
#pragma unroll
    for (int t = 0; t < kMaxTaps; ++t)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[t][k] = 0.0f;
    if (active) {
      for (int64_t pix = (int64_t)blockIdx.x * ny + ty; pix < pixels; pix += (int64_t)gridDim.x * ny) {
        const int q = (int)(pix % g.q);
        const int p = (int)((pix / g.q) % g.p);
        const int n = (int)(pix / ((int64_t)g.q * g.p));
        float d[8];
        unpack8(ld8(dy + (size_t)pix * g.c + cv * 8), d);
#pragma unroll
        for (int t = 0; t < kMaxTaps; ++t) {
          if (t >= taps) break;
          const int r = t / g.s, s = t - r * g.s;
          const int ih = p * g.sh - g.pt + r, iw = q * g.sw - g.pl + s;
          if (ih < 0 || ih >= g.h || iw < 0 || iw >= g.w) continue;
          float f[8];
          unpack8(ld8(x + (((size_t)n * g.h + ih) * g.w + iw) * g.c + cv * 8), f);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[t][k] = fmaf(f[k], d[k], acc[t][k]);
        }
      }
    }
#pragma unroll
    for (int t = 0; t < kMaxTaps; ++t) {
      if (t >= taps) break;
      __syncthreads();
#pragma unroll
      for (int k = 0; k < 8; ++k) red[(ty * nx + tx) * 8 + k] = acc[t][k];
      __syncthreads();
      if (active)
        for (int k = ty; k < 8; k += ny) {
          float sum = 0.0f;
          for (int j = 0; j < ny; ++j) sum += red[(j * nx + tx) * 8 + k];
          partials[((size_t)blockIdx.x * taps + t) * g.c + cv * 8 + k] = sum;
        }
    }
  }
}

// block (32 columns, 8 row lanes): lane j sums rows j, j+8, ...; the 8 lane sums are added in lane order
__global__ void __launch_bounds__(256) dw_fold_kernel(const float* __restrict__ partials, float* __restrict__ dw, int rows,
                                                      int taps, int c) {
  __shared__ float red[8][33];
  const int i = blockIdx.x * 32 + threadIdx.x;
  const int cols = taps * c;
  float a = 0.0f;
  if (i < cols)
    for (int r = threadIdx.y; r < rows; r += 8) a += partials[(size_t)r * cols + i];
  red[threadIdx.y][threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.y == 0 && i < cols) {
    float sum = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) sum += red[j][threadIdx.x];
    const int t = i / c, ch = i - t * c;
    dw[(size_t)ch * taps + t] = sum;
  }
}

struct DwLaunch {
  dim3 block;
  int grid;
};

// block = (channel vectors up to 128, pixels filling 256 threads); `pixels` caps the grid for small layers
DwLaunch dw_launch(int c, int64_t pixels, int blocks_per_sm) {
  const int cvecs = c / 8;
  int tx = cvecs < 128 ? cvecs : 128;
  int ty = 256 / tx;
  if (ty < 1) ty = 1;
  int64_t want = (pixels + ty - 1) / ty;
  int64_t cap = (int64_t)kNumSMs * blocks_per_sm;
  return {dim3(tx, ty), (int)(want < cap ? want : cap)};
}

int check(const b200_conv_geom* g, const char* what) {
  B200_CHECK_ARG(g, "%s: null geometry", what);
  B200_CHECK_ARG(g->k == g->c && g->c % 8 == 0, "%s: depthwise needs k == c and c %% 8 == 0 (k=%d c=%d)", what, g->k, g->c);
  B200_CHECK_ARG(g->r * g->s <= kMaxTaps && g->r > 0 && g->s > 0, "%s: filter %dx%d has more than %d taps", what, g->r, g->s, kMaxTaps);
  B200_CHECK_ARG(g->stride_h > 0 && g->stride_w > 0 && g->n > 0 && g->p > 0 && g->q > 0, "%s: bad geometry", what);
  return 0;
}

DwGeom to_dw(const b200_conv_geom* g) {
  return {g->n, g->h, g->w, g->c, g->r, g->s, g->p, g->q, g->stride_h, g->stride_w, g->pad_t, g->pad_l};
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_dwconv_supported(const b200_conv_geom* g) {
  return g && g->k == g->c && g->c % 8 == 0 && g->r > 0 && g->s > 0 && g->r * g->s <= kMaxTaps;
}

int b200_dwconv_fprop(const b200_bf16* x, const b200_bf16* w, b200_bf16* y, const b200_conv_geom* g, void* stream) {
  B200_CHECK_ARG(x && w && y, "dwconv_fprop: null pointer");
  if (int e = check(g, "dwconv_fprop")) return e;
  DwLaunch l = dw_launch(g->c, (int64_t)g->n * g->p * g->q, 8);
  dw_fwd_kernel<<<l.grid, l.block, 0, static_cast<cudaStream_t>(stream)>>>(x, w, y, to_dw(g));
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_dwconv_dgrad(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* stream) {
  B200_CHECK_ARG(dy && w && dx, "dwconv_dgrad: null pointer");
  if (int e = check(g, "dwconv_dgrad")) return e;
  DwLaunch l = dw_launch(g->c, (int64_t)g->n * g->h * g->w, 8);
  dw_dgrad_kernel<<<l.grid, l.block, 0, static_cast<cudaStream_t>(stream)>>>(dy, w, dx, to_dw(g));
  B200_LAUNCH_CHECK();
  return 0;
}

static int64_t dw_wgrad_rows(const b200_conv_geom* g) {
  return dw_launch(g->c, (int64_t)g->n * g->p * g->q, 2).grid;
}

int64_t b200_dwconv_wgrad_workspace_bytes(const b200_conv_geom* g) {
  if (!b200_dwconv_supported(g)) return 0;
  return dw_wgrad_rows(g) * g->r * g->s * g->c * (int64_t)sizeof(float);
}

int b200_dwconv_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, void* workspace,
                      void* stream) {
  B200_CHECK_ARG(x && dy && dw && workspace, "dwconv_wgrad: null pointer");
  if (int e = check(g, "dwconv_wgrad")) return e;
  DwLaunch l = dw_launch(g->c, (int64_t)g->n * g->p * g->q, 2);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dw_wgrad_kernel<<<l.grid, l.block, 0, st>>>(x, dy, static_cast<float*>(workspace), to_dw(g));
  B200_LAUNCH_CHECK();
  const int taps = g->r * g->s;
  dw_fold_kernel<<<(taps * g->c + 31) / 32, dim3(32, 8), 0, st>>>(static_cast<const float*>(workspace), dw, l.grid, taps,
                                                                  g->c);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
