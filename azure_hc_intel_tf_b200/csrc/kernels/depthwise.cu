// depthwise.cu — depthwise convolution (groups == channels) forward, backward-data and backward-weights, NHWC bf16.
//
// Role in the reference stack: MKL-DNN's depthwise convolution primitive under --mkl=TRUE
// (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76), reached when --model (:66) selects MobileNet. A depthwise layer
// does R*S multiply-adds per output element and reads every input once: it is bandwidth-bound, there is no GEMM in it,
// so these are plain vector kernels (8 channels = one 16-byte vector per thread, filter taps in registers), not
// tensor-core kernels. Filters are [C][R][S] (the KRSC layout of a [C,1,R,S] weight), up to 7x7; 3x3 at stride 1 / 2
// (every MobileNet layer) has a fast path, everything else (NASNet's 5x5 / 7x7) the generic kernels.
#include "b200conv.h"
#include "common.cuh"

#include <climits>
#include <cstdlib>

namespace b200 {
namespace {

constexpr int kMaxTaps = 49, kMaxS = 7;      // generic kernels: filters up to 7x7
constexpr int kGenCvecs = 32;                 // 8-channel vectors per block of the generic kernels

struct DwGeom {
  int n, h, w, c, r, s, p, q, sh, sw, pt, pl;
};

// acc += a.half[ha] * b.half[hb]: mixed-precision FMA with the bf16 operands taken in place from the packed registers
// (PTX fma.rn.f32.bf16 = SASS FHFMA.BF16 with .H0/.H1 operand selectors on sm_100a) - exact product, one rounding, so
// it equals an fp32 FMA of the converted values without spending instructions on the conversions. ha / hb are
// compile-time constants at every call (fully unrolled loops), so one of the four forms survives.
__device__ __forceinline__ void fma_bf16(float& acc, uint32_t a, int ha, uint32_t b, int hb) {
  if (ha == 0 && hb == 0)
    asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.bf16 %0, al, bl, %0; }" : "+f"(acc) : "r"(a), "r"(b));
  else if (ha == 0)
    asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.bf16 %0, al, bh, %0; }" : "+f"(acc) : "r"(a), "r"(b));
  else if (hb == 0)
    asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.bf16 %0, ah, bl, %0; }" : "+f"(acc) : "r"(a), "r"(b));
  else
    asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.bf16 %0, ah, bh, %0; }" : "+f"(acc) : "r"(a), "r"(b));
}

// ---------------------------------------------------------------------------------------------------------------
// Generic kernels (any filter up to 7x7, any stride: NASNet's 5x5 / 7x7 separable convolutions, 2x2 filters, ...).
// Block = (up to 32 channel vectors, pixels); gridDim.y walks the channel chunks. The block's filter taps are staged
// once in shared memory as [tap][channel vector] 16-byte words, so a thread reads one input vector and one tap vector
// per tap and feeds both packed words to FHFMA.BF16.

// taps of channels [c0, c0 + 8 * nvec) -> smem [tap][vec] (uint4 = 8 bf16); all threads of the block take part
__device__ __forceinline__ void stage_taps(const b200_bf16* __restrict__ w, int c0, int nvec, int taps, uint4* wsm) {
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, nthr = blockDim.x * blockDim.y;
  for (int e = tid; e < taps * nvec; e += nthr) {
    const int t = e / nvec, v = e - t * nvec;
    uint32_t pk[4];
#pragma unroll
    for (int k2 = 0; k2 < 4; ++k2) {
      const uint32_t lo = w[(size_t)(c0 + v * 8 + 2 * k2) * taps + t], hi = w[(size_t)(c0 + v * 8 + 2 * k2 + 1) * taps + t];
      pk[k2] = lo | (hi << 16);
    }
    wsm[t * kGenCvecs + v] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
  __syncthreads();
}

__device__ __forceinline__ void fma_vec8(float acc[8], const uint4& a, const uint4& b) {
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
  for (int k = 0; k < 8; ++k) fma_bf16(acc[k], aw[k >> 1], k & 1, bw[k >> 1], k & 1);
}

__global__ void __launch_bounds__(256) dw_fwd_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ w,
                                                     b200_bf16* __restrict__ y, DwGeom g) {
  __shared__ uint4 wsm[kMaxTaps * kGenCvecs];
  const int cvecs = g.c / 8, taps = g.r * g.s;
  const int cv0 = blockIdx.y * kGenCvecs, nvec = min(kGenCvecs, cvecs - cv0);
  stage_taps(w, cv0 * 8, nvec, taps, wsm);
  if ((int)threadIdx.x >= nvec) return;
  const int cv = cv0 + threadIdx.x;
  const int64_t pixels = (int64_t)g.n * g.p * g.q;
  for (int64_t pix = (int64_t)blockIdx.x * blockDim.y + threadIdx.y; pix < pixels; pix += (int64_t)gridDim.x * blockDim.y) {
    const int q = (int)(pix % g.q);
    const int p = (int)((pix / g.q) % g.p);
    const int n = (int)(pix / ((int64_t)g.q * g.p));
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    for (int r = 0; r < g.r; ++r) {
      const int ih = p * g.sh - g.pt + r;
      if (ih < 0 || ih >= g.h) continue;
      const b200_bf16* xr = x + ((size_t)(n * g.h + ih) * g.w) * g.c + cv * 8;
#pragma unroll
      for (int s = 0; s < kMaxS; ++s) {
        const int iw = q * g.sw - g.pl + s;
        if (s < g.s && iw >= 0 && iw < g.w)
          fma_vec8(acc, __ldg(reinterpret_cast<const uint4*>(xr + (size_t)iw * g.c)), wsm[(r * g.s + s) * kGenCvecs + threadIdx.x]);
      }
    }
    st8(y + (size_t)pix * g.c + cv * 8, pack8(acc));
  }
}

// dx[n,h,w,c] = sum over taps (r,s) with (h + pt - r) % sh == 0 and (w + pl - s) % sw == 0 of dy[n, ., ., c] * w[c,r,s]
__global__ void __launch_bounds__(256) dw_dgrad_kernel(const b200_bf16* __restrict__ dy, const b200_bf16* __restrict__ w,
                                                       b200_bf16* __restrict__ dx, DwGeom g) {
  __shared__ uint4 wsm[kMaxTaps * kGenCvecs];
  const int cvecs = g.c / 8, taps = g.r * g.s;
  const int cv0 = blockIdx.y * kGenCvecs, nvec = min(kGenCvecs, cvecs - cv0);
  stage_taps(w, cv0 * 8, nvec, taps, wsm);
  if ((int)threadIdx.x >= nvec) return;
  const int cv = cv0 + threadIdx.x;
  const int64_t pixels = (int64_t)g.n * g.h * g.w;
  for (int64_t pix = (int64_t)blockIdx.x * blockDim.y + threadIdx.y; pix < pixels; pix += (int64_t)gridDim.x * blockDim.y) {
    const int iw = (int)(pix % g.w);
    const int ih = (int)((pix / g.w) % g.h);
    const int n = (int)(pix / ((int64_t)g.w * g.h));
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    for (int r = 0; r < g.r; ++r) {
      const int a = ih + g.pt - r;
      if (a < 0 || a % g.sh || a / g.sh >= g.p) continue;
      const b200_bf16* yr = dy + ((size_t)(n * g.p + a / g.sh) * g.q) * g.c + cv * 8;
#pragma unroll
      for (int s = 0; s < kMaxS; ++s) {
        const int b = iw + g.pl - s;
        if (s < g.s && b >= 0 && b % g.sw == 0 && b / g.sw < g.q)
          fma_vec8(acc, __ldg(reinterpret_cast<const uint4*>(yr + (size_t)(b / g.sw) * g.c)), wsm[(r * g.s + s) * kGenCvecs + threadIdx.x]);
      }
    }
    st8(dx + (size_t)pix * g.c + cv * 8, pack8(acc));
  }
}

// dw[c,r,s] = sum over output pixels of dy[n,p,q,c] * x[n, p*sh - pt + r, q*sw - pl + s, c]. gridDim.y = channel chunks
// x filter rows: a thread owns 8 channels and ONE filter row (up to 7 x 8 fp32 sums in registers) over a strided set of
// pixels; the block folds its pixel lanes (threadIdx.y) through shared memory in lane order and writes one row of
// `partials` ([gridDim.x][taps][c]); dw_fold_kernel sums the rows in index order (bitwise reproducible).
__global__ void __launch_bounds__(256) dw_wgrad_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ dy,
                                                       float* __restrict__ partials, DwGeom g) {
  __shared__ float red[256 * 8];
  const int cvecs = g.c / 8, taps = g.r * g.s;
  const int chunk = blockIdx.y / g.r, r = blockIdx.y - chunk * g.r;
  const int tx = threadIdx.x, ty = threadIdx.y, nx = blockDim.x, ny = blockDim.y;
  const int cv = chunk * kGenCvecs + tx;
  const bool active = tx < kGenCvecs && cv < cvecs;
  const int64_t pixels = (int64_t)g.n * g.p * g.q;
  float acc[kMaxS][8];
#pragma unroll
  for (int s = 0; s < kMaxS; ++s)
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[s][k] = 0.0f;
  if (active) {
    for (int64_t pix = (int64_t)blockIdx.x * ny + ty; pix < pixels; pix += (int64_t)gridDim.x * ny) {
      const int q = (int)(pix % g.q);
      const int p = (int)((pix / g.q) % g.p);
      const int n = (int)(pix / ((int64_t)g.q * g.p));
      const int ih = p * g.sh - g.pt + r;
      if (ih < 0 || ih >= g.h) continue;
      const uint4 d = __ldg(reinterpret_cast<const uint4*>(dy + (size_t)pix * g.c + cv * 8));
      const b200_bf16* xr = x + ((size_t)(n * g.h + ih) * g.w) * g.c + cv * 8;
#pragma unroll
      for (int s = 0; s < kMaxS; ++s) {
        const int iw = q * g.sw - g.pl + s;
        if (s < g.s && iw >= 0 && iw < g.w) fma_vec8(acc[s], __ldg(reinterpret_cast<const uint4*>(xr + (size_t)iw * g.c)), d);
      }
    }
  }
#pragma unroll
  for (int s = 0; s < kMaxS; ++s) {
    if (s >= g.s) break;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; ++k) red[(ty * nx + tx) * 8 + k] = acc[s][k];
    __syncthreads();
    if (active)
      for (int k = ty; k < 8; k += ny) {
        float sum = 0.0f;
        for (int j = 0; j < ny; ++j) sum += red[(j * nx + tx) * 8 + k];
        partials[((size_t)blockIdx.x * taps + r * g.s + s) * g.c + cv * 8 + k] = sum;
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

// ---------------------------------------------------------------------------------------------------------------
// 3x3 fast path (every depthwise layer of MobileNet): a thread produces kT consecutive outputs of one row for its V
// channels (V = 4 or 8: one 8- or 16-byte vector). The input columns those outputs share are loaded once per filter
// row - (kT-1)*stride + 3 vector loads, issued ahead of the FMAs and predicated (never branched) on the image bounds -
// instead of 9 per output; the filter taps stay packed as bf16 pairs in registers and feed FHFMA.BF16 directly (fma_bf16);
// every index computation is 32-bit.
// V = 4 halves the registers per thread (twice the resident warps to hide the load latency); see dw_vec().

constexpr int kT = 4;

template <int V>
struct Vec {                       // V bf16 values
  uint32_t w[V / 2];
};

template <int V>
__device__ __forceinline__ Vec<V> ldv(const b200_bf16* p, bool ok) {
  Vec<V> v;
#pragma unroll
  for (int i = 0; i < V / 2; ++i) v.w[i] = 0u;
  if (ok) {
    if (V == 8) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
      v.w[0] = u.x, v.w[1] = u.y, v.w[V / 2 - 2] = u.z, v.w[V / 2 - 1] = u.w;
    } else {
      const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
      v.w[0] = u.x, v.w[1] = u.y;
    }
  }
  return v;
}

template <int V>
__device__ __forceinline__ void st_acc(b200_bf16* p, const float a[V]) {
  uint32_t o[V / 2];
#pragma unroll
  for (int i = 0; i < V / 2; ++i) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
    o[i] = *reinterpret_cast<const uint32_t*>(&h);
  }
  if (V == 8)
    *reinterpret_cast<uint4*>(p) = make_uint4(o[0], o[1], o[V / 2 - 2], o[V / 2 - 1]);
  else
    *reinterpret_cast<uint2*>(p) = make_uint2(o[0], o[1]);
}

template <int V>
struct Taps3 {
  uint32_t w[V * 9 / 2];   // [V channels][9 taps] bf16, exactly as stored in the filter tensor (contiguous bytes)
};

template <int V>
__device__ __forceinline__ void load_taps3(const b200_bf16* __restrict__ w, int c0, Taps3<V>& t) {
  const uint2* p = reinterpret_cast<const uint2*>(w + (size_t)c0 * 9);      // c0 % 4 == 0: 72-byte multiples
#pragma unroll
  for (int i = 0; i < V * 9 / 4; ++i) {
    const uint2 v = __ldg(p + i);
    t.w[2 * i] = v.x;
    t.w[2 * i + 1] = v.y;
  }
}

// acc += (element k of a V-vector) * (tap tp of channel k)
template <int V>
__device__ __forceinline__ void fma_tap(float& acc, const Vec<V>& v, const Taps3<V>& t, int k, int tp) {
  const int idx = k * 9 + tp;
  fma_bf16(acc, v.w[k >> 1], k & 1, t.w[idx >> 1], idx & 1);
}

// A thread walks its runs (row, chunk-of-kT-outputs) with the grid stride; the stride is decomposed on the host into
// (images, rows, chunks) so that the walk needs no division. All element offsets are 32-bit (fast3() checks the sizes).
struct RunStep {
  int dc, dr, dn;
};
struct RunIter {
  int chunk, row, n;   // row within the image
};

__device__ __forceinline__ RunIter first_run(int run, int chunks, int rows) {
  RunIter it;
  it.chunk = run % chunks;
  const int t = run / chunks;
  it.row = t % rows;
  it.n = t / rows;
  return it;
}

__device__ __forceinline__ void next_run(RunIter& it, const RunStep& st, int chunks, int rows) {
  it.chunk += st.dc;
  if (it.chunk >= chunks) it.chunk -= chunks, ++it.row;
  it.row += st.dr;
  if (it.row >= rows) it.row -= rows, ++it.n;
  it.n += st.dn;
}

template <int ST, int V>
__global__ void __launch_bounds__(256, V == 4 ? 3 : 1)
dw3_fwd_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ w, b200_bf16* __restrict__ y, DwGeom g,
               int chunks, RunStep step) {
  constexpr int NC = (kT - 1) * ST + 3;
  const int cvecs = g.c / V;
  for (int cv = threadIdx.x; cv < cvecs; cv += blockDim.x) {
    Taps3<V> t;
    load_taps3<V>(w, cv * V, t);
    for (RunIter it = first_run(blockIdx.x * blockDim.y + threadIdx.y, chunks, g.p); it.n < g.n; next_run(it, step, chunks, g.p)) {
      const int q0 = it.chunk * kT, iw0 = q0 * ST - g.pl, ih0 = it.row * ST - g.pt;
      bool cok[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c) cok[c] = (unsigned)(iw0 + c) < (unsigned)g.w;
      float acc[kT][V];
#pragma unroll
      for (int j = 0; j < kT; ++j)
#pragma unroll
        for (int k = 0; k < V; ++k) acc[j][k] = 0.0f;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const bool rok = (unsigned)(ih0 + r) < (unsigned)g.h;
        const int off = ((it.n * g.h + ih0 + r) * g.w + iw0) * g.c + cv * V;
        Vec<V> L[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) L[c] = ldv<V>(x + (off + c * g.c), rok && cok[c]);
#pragma unroll
        for (int j = 0; j < kT; ++j)
#pragma unroll
          for (int s = 0; s < 3; ++s)
#pragma unroll
            for (int k = 0; k < V; ++k) fma_tap<V>(acc[j][k], L[j * ST + s], t, k, r * 3 + s);
      }
      const int yoff = ((it.n * g.p + it.row) * g.q + q0) * g.c + cv * V;
#pragma unroll
      for (int j = 0; j < kT; ++j)
        if (q0 + j < g.q) st_acc<V>(y + (yoff + j * g.c), acc[j]);
    }
  }
}

// taps s = S0, S0 + ST, ... of filter row r: output j (pixel w0 + j*ST) reads dy column qb - m + j = L[j + MM - m]
template <int ST, int S0, int V>
__device__ __forceinline__ void dgrad_row(float acc[kT][V], const Vec<V>* L, const Taps3<V>& t, int r) {
  constexpr int MM = 2 / ST;
#pragma unroll
  for (int j = 0; j < kT; ++j)
#pragma unroll
    for (int m = 0; m <= MM; ++m) {
      const int s = S0 + m * ST;
      if (s < 3) {
#pragma unroll
        for (int k = 0; k < V; ++k) fma_tap<V>(acc[j][k], L[j + MM - m], t, k, r * 3 + s);
      }
    }
}

// dx: a thread produces the kT pixels w0, w0 + ST, ... of one stride phase of a row, so that all of them use the same
// filter columns (s == (w + pl) mod ST) and consecutive dy columns. chunks = ST * cpp, phase-major.
template <int ST, int V>
__global__ void __launch_bounds__(256, V == 4 ? 3 : 2)
dw3_dgrad_kernel(const b200_bf16* __restrict__ dy, const b200_bf16* __restrict__ w, b200_bf16* __restrict__ dx, DwGeom g,
                 int cpp, RunStep step) {
  constexpr int MM = 2 / ST, NL = kT + MM;
  const int cvecs = g.c / V;
  const int chunks = cpp * ST;
  for (int cv = threadIdx.x; cv < cvecs; cv += blockDim.x) {
    Taps3<V> t;
    load_taps3<V>(w, cv * V, t);
    for (RunIter it = first_run(blockIdx.x * blockDim.y + threadIdx.y, chunks, g.h); it.n < g.n; next_run(it, step, chunks, g.h)) {
      const int phase = (ST == 1 || it.chunk < cpp) ? 0 : 1, ci = it.chunk - phase * cpp;
      const int w0 = ci * kT * ST + phase;
      const int s0 = (w0 + g.pl) % ST, qb = (w0 + g.pl - s0) / ST - MM;
      bool cok[NL];
#pragma unroll
      for (int c = 0; c < NL; ++c) cok[c] = (unsigned)(qb + c) < (unsigned)g.q;
      float acc[kT][V];
#pragma unroll
      for (int j = 0; j < kT; ++j)
#pragma unroll
        for (int k = 0; k < V; ++k) acc[j][k] = 0.0f;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int a = it.row + g.pt - r;
        const int pr = a / ST;
        if (a < 0 || a % ST != 0 || pr >= g.p) continue;
        const int off = ((it.n * g.p + pr) * g.q + qb) * g.c + cv * V;
        Vec<V> L[NL];
#pragma unroll
        for (int c = 0; c < NL; ++c) L[c] = ldv<V>(dy + (off + c * g.c), cok[c]);
        if (ST == 1 || s0 == 0)
          dgrad_row<ST, 0, V>(acc, L, t, r);
        else
          dgrad_row<ST, 1, V>(acc, L, t, r);
      }
      const int xoff = ((it.n * g.h + it.row) * g.w + w0) * g.c + cv * V;
#pragma unroll
      for (int j = 0; j < kT; ++j)
        if (w0 + j * ST < g.w) st_acc<V>(dx + (xoff + j * ST * g.c), acc[j]);
    }
  }
}

// dw: same loads as the forward kernel plus the kT dy vectors; 9*V fp32 sums per thread, folded like dw_wgrad_kernel
template <int ST, int V>
__global__ void __launch_bounds__(256, V == 4 ? 2 : 1)
dw3_wgrad_kernel(const b200_bf16* __restrict__ x, const b200_bf16* __restrict__ dy, float* __restrict__ partials, DwGeom g,
                 int chunks, RunStep step) {
  constexpr int NC = (kT - 1) * ST + 3;
  __shared__ float red[256 * V];
  const int cvecs = g.c / V;
  const int tx = threadIdx.x, ty = threadIdx.y, nx = blockDim.x, ny = blockDim.y;
  for (int cv0 = 0; cv0 < cvecs; cv0 += nx) {
    const int cv = cv0 + tx;
    const bool active = cv < cvecs;
    float acc[9][V];
#pragma unroll
    for (int tp = 0; tp < 9; ++tp)
#pragma unroll
      for (int k = 0; k < V; ++k) acc[tp][k] = 0.0f;
    if (active) {
      for (RunIter it = first_run(blockIdx.x * ny + ty, chunks, g.p); it.n < g.n; next_run(it, step, chunks, g.p)) {
        const int q0 = it.chunk * kT, iw0 = q0 * ST - g.pl, ih0 = it.row * ST - g.pt;
        const int yoff = ((it.n * g.p + it.row) * g.q + q0) * g.c + cv * V;
        Vec<V> d[kT];
#pragma unroll
        for (int j = 0; j < kT; ++j) d[j] = ldv<V>(dy + (yoff + j * g.c), q0 + j < g.q);
        bool cok[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) cok[c] = (unsigned)(iw0 + c) < (unsigned)g.w;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const bool rok = (unsigned)(ih0 + r) < (unsigned)g.h;
          const int off = ((it.n * g.h + ih0 + r) * g.w + iw0) * g.c + cv * V;
          Vec<V> L[NC];
#pragma unroll
          for (int c = 0; c < NC; ++c) L[c] = ldv<V>(x + (off + c * g.c), rok && cok[c]);
#pragma unroll
          for (int j = 0; j < kT; ++j)
#pragma unroll
            for (int s = 0; s < 3; ++s)
#pragma unroll
              for (int k = 0; k < V; ++k) fma_bf16(acc[r * 3 + s][k], L[j * ST + s].w[k >> 1], k & 1, d[j].w[k >> 1], k & 1);
        }
      }
    }
#pragma unroll
    for (int tp = 0; tp < 9; ++tp) {
      __syncthreads();
#pragma unroll
      for (int k = 0; k < V; ++k) red[(ty * nx + tx) * V + k] = acc[tp][k];
      __syncthreads();
      if (active)
        for (int k = ty; k < V; k += ny) {
          float sum = 0.0f;
          for (int j = 0; j < ny; ++j) sum += red[(j * nx + tx) * V + k];
          partials[((size_t)blockIdx.x * 9 + tp) * g.c + cv * V + k] = sum;
        }
    }
  }
}

// channels per thread of the 3x3 kernels: measured best (profiles/depthwise_bench.md) is 4 for the forward kernel (twice
// the resident warps) and 8 for the two backward kernels (fewer instructions per byte); B200_DW_VEC=4|8 forces one.
enum { kOpFwd = 0, kOpDgrad = 1, kOpWgrad = 2 };
int dw_vec(int op) {
  static const int forced = [] {
    const char* e = getenv("B200_DW_VEC");
    const int v = e ? atoi(e) : 0;
    return (v == 4 || v == 8) ? v : 0;
  }();
  return forced ? forced : (op == kOpFwd ? 4 : 8);
}

bool fast3(const b200_conv_geom* g) {
  const int64_t in = (int64_t)g->n * g->h * g->w * g->c, out = (int64_t)g->n * g->p * g->q * g->c;
  return g->r == 3 && g->s == 3 && g->stride_h == g->stride_w && (g->stride_h == 1 || g->stride_h == 2) &&
         g->c % 4 == 0 && in + 16 * (int64_t)g->c < INT32_MAX && out + 16 * (int64_t)g->c < INT32_MAX;   // 32-bit element offsets
}

struct DwLaunch {
  dim3 block;
  int grid;
};

// block = (channel vectors up to 256, pixels filling 256 threads); `pixels` caps the grid for small layers
DwLaunch dw_launch(int c, int64_t pixels, int blocks_per_sm, int vec = 8) {
  const int cvecs = c / vec;
  int tx = cvecs < 256 ? cvecs : 256;
  int ty = 256 / tx;
  if (ty < 1) ty = 1;
  int64_t want = (pixels + ty - 1) / ty;
  int64_t cap = (int64_t)kNumSMs * blocks_per_sm;
  return {dim3(tx, ty), (int)(want < cap ? want : cap)};
}

// generic kernels: block = (up to 32 channel vectors, pixels), grid = (pixel blocks, channel chunks [x filter rows])
struct GenLaunch {
  dim3 block, grid;
};
GenLaunch gen_launch(int c, int64_t pixels, int blocks_per_sm, int y_mult = 1) {
  const int cvecs = c / 8;
  const int tx = cvecs < kGenCvecs ? cvecs : kGenCvecs, ty = 256 / tx;
  const int gy = (cvecs + kGenCvecs - 1) / kGenCvecs * y_mult;
  int64_t want = (pixels + ty - 1) / ty;
  int64_t cap = (int64_t)kNumSMs * blocks_per_sm / gy;
  if (cap < 1) cap = 1;
  return {dim3(tx, ty), dim3((unsigned)(want < cap ? want : cap), gy)};
}

// grid stride (in runs) as (images, rows, chunks)
RunStep run_step(const DwLaunch& l, int chunks, int rows) {
  const int stride = l.grid * (int)l.block.y;
  return {stride % chunks, (stride / chunks) % rows, (stride / chunks) / rows};
}

int check(const b200_conv_geom* g, const char* what) {
  B200_CHECK_ARG(g, "%s: null geometry", what);
  B200_CHECK_ARG(g->k == g->c && g->c % 8 == 0, "%s: depthwise needs k == c and c %% 8 == 0 (k=%d c=%d)", what, g->k, g->c);
  B200_CHECK_ARG(g->r > 0 && g->s > 0 && g->r <= kMaxS && g->s <= kMaxS, "%s: filter %dx%d is larger than %dx%d", what, g->r, g->s, kMaxS, kMaxS);
  B200_CHECK_ARG(g->stride_h > 0 && g->stride_w > 0 && g->n > 0 && g->p > 0 && g->q > 0, "%s: bad geometry", what);
  return 0;
}

DwGeom to_dw(const b200_conv_geom* g) {
  return {g->n, g->h, g->w, g->c, g->r, g->s, g->p, g->q, g->stride_h, g->stride_w, g->pad_t, g->pad_l};
}

}  // namespace
}  // namespace b200

using namespace b200;

// instantiate on (stride, channels per thread)
#define DW3_LAUNCH(KERNEL, vec, stride, l, st, ...)                                     \
  do {                                                                                  \
    if ((vec) == 4) {                                                                \
      if ((stride) == 1) KERNEL<1, 4><<<(l).grid, (l).block, 0, st>>>(__VA_ARGS__);     \
      else KERNEL<2, 4><<<(l).grid, (l).block, 0, st>>>(__VA_ARGS__);                   \
    } else {                                                                            \
      if ((stride) == 1) KERNEL<1, 8><<<(l).grid, (l).block, 0, st>>>(__VA_ARGS__);     \
      else KERNEL<2, 8><<<(l).grid, (l).block, 0, st>>>(__VA_ARGS__);                   \
    }                                                                                   \
  } while (0)

extern "C" {

int b200_dwconv_supported(const b200_conv_geom* g) {
  return g && g->k == g->c && g->c % 8 == 0 && g->r > 0 && g->s > 0 && g->r <= kMaxS && g->s <= kMaxS;
}

int b200_dwconv_fprop(const b200_bf16* x, const b200_bf16* w, b200_bf16* y, const b200_conv_geom* g, void* stream) {
  B200_CHECK_ARG(x && w && y, "dwconv_fprop: null pointer");
  if (int e = check(g, "dwconv_fprop")) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (fast3(g)) {
    const int chunks = (g->q + kT - 1) / kT, runs = g->n * g->p * chunks;
    DwLaunch l = dw_launch(g->c, runs, 8, dw_vec(kOpFwd));
    DW3_LAUNCH(dw3_fwd_kernel, dw_vec(kOpFwd), g->stride_h, l, st, x, w, y, to_dw(g), chunks, run_step(l, chunks, g->p));
  } else {
    GenLaunch l = gen_launch(g->c, (int64_t)g->n * g->p * g->q, 8);
    dw_fwd_kernel<<<l.grid, l.block, 0, st>>>(x, w, y, to_dw(g));
  }
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_dwconv_dgrad(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* stream) {
  B200_CHECK_ARG(dy && w && dx, "dwconv_dgrad: null pointer");
  if (int e = check(g, "dwconv_dgrad")) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (fast3(g)) {
    const int sw = g->stride_w;
    const int cpp = ((g->w + sw - 1) / sw + kT - 1) / kT;      // chunks per stride phase of a row
    const int runs = g->n * g->h * cpp * sw;
    DwLaunch l = dw_launch(g->c, runs, 8, dw_vec(kOpDgrad));
    DW3_LAUNCH(dw3_dgrad_kernel, dw_vec(kOpDgrad), sw, l, st, dy, w, dx, to_dw(g), cpp, run_step(l, cpp * sw, g->h));
  } else {
    GenLaunch l = gen_launch(g->c, (int64_t)g->n * g->h * g->w, 8);
    dw_dgrad_kernel<<<l.grid, l.block, 0, st>>>(dy, w, dx, to_dw(g));
  }
  B200_LAUNCH_CHECK();
  return 0;
}

static DwLaunch dw3_wgrad_launch(const b200_conv_geom* g) {
  return dw_launch(g->c, (int64_t)g->n * g->p * ((g->q + kT - 1) / kT), dw_vec(kOpWgrad) == 4 ? 4 : 2, dw_vec(kOpWgrad));
}
static GenLaunch gen_wgrad_launch(const b200_conv_geom* g) {
  return gen_launch(g->c, (int64_t)g->n * g->p * g->q, 4, g->r);
}

int64_t b200_dwconv_wgrad_workspace_bytes(const b200_conv_geom* g) {
  if (!b200_dwconv_supported(g)) return 0;
  const int64_t rows = fast3(g) ? dw3_wgrad_launch(g).grid : gen_wgrad_launch(g).grid.x;
  return rows * g->r * g->s * g->c * (int64_t)sizeof(float);
}

int b200_dwconv_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, void* workspace,
                      void* stream) {
  B200_CHECK_ARG(x && dy && dw && workspace, "dwconv_wgrad: null pointer");
  if (int e = check(g, "dwconv_wgrad")) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* part = static_cast<float*>(workspace);
  int rows;
  if (fast3(g)) {
    DwLaunch l = dw3_wgrad_launch(g);
    const int chunks = (g->q + kT - 1) / kT;
    DW3_LAUNCH(dw3_wgrad_kernel, dw_vec(kOpWgrad), g->stride_h, l, st, x, dy, part, to_dw(g), chunks, run_step(l, chunks, g->p));
    rows = l.grid;
  } else {
    GenLaunch l = gen_wgrad_launch(g);
    dw_wgrad_kernel<<<l.grid, l.block, 0, st>>>(x, dy, part, to_dw(g));
    rows = (int)l.grid.x;
  }
  B200_LAUNCH_CHECK();
  const int taps = g->r * g->s;
  dw_fold_kernel<<<(taps * g->c + 31) / 32, dim3(32, 8), 0, st>>>(part, dw, rows, taps, g->c);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
