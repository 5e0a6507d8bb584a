// pool.cu — max / average / global-average pooling forward and backward for NHWC bf16.
//
// Role in the reference stack: MKL-DNN pooling primitives (K7 in SURVEY.md §2.6): ResNet's 3x3/2 max-pool
// and spatial_mean, VGG's 2x2/2 max-pools, Inception's 3x3/1 'SAME' average pools.
// One thread handles 8 channels (one 16-byte vector) of one output (forward) or input (backward) pixel, so
// every access is a full-width coalesced vector and backward needs no atomics.
#include "common.cuh"

namespace b200 {
namespace {

// Idx = uint32_t whenever the element count fits: 64-bit integer division in the index decode would otherwise cost
// more than the memory traffic of these kernels.
template <typename Idx>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const b200_bf16* __restrict__ x, b200_bf16* __restrict__ y,
                                                          uint8_t* __restrict__ idx, b200_pool_geom g) {
  const int cv = g.c / 8;
  const Idx total = (Idx)g.n * g.p * g.q * cv;
  for (Idx i = (Idx)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (Idx)gridDim.x * blockDim.x) {
    int c8 = (int)(i % cv);
    Idx t = i / cv;
    int q = (int)(t % g.q);
    t /= g.q;
    int p = (int)(t % g.p);
    int n = (int)(t / g.p);
    float best[8];
    int arg[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      best[k] = -INFINITY;
      arg[k] = 0;
    }
    const int h0 = p * g.sh - g.pad_t, w0 = q * g.sw - g.pad_l;
    for (int r = 0; r < g.kh; ++r) {
      int h = h0 + r;
      if (h < 0 || h >= g.h) continue;
      for (int s = 0; s < g.kw; ++s) {
        int w = w0 + s;
        if (w < 0 || w >= g.w) continue;
        float f[8];
        unpack8(ld8(x + (((int64_t)n * g.h + h) * g.w + w) * g.c + c8 * 8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (f[k] > best[k]) {
            best[k] = f[k];
            arg[k] = r * g.kw + s;
          }
        }
      }
    }
    st8(y + (int64_t)i * 8, pack8(best));
    uint2 pk;
    pk.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    pk.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    *reinterpret_cast<uint2*>(idx + (int64_t)i * 8) = pk;
  }
}

// Forward with a compile-time window: all KH*KW vector loads of an output are issued before the first compare, so a
// thread has the whole window in flight instead of one load per loop trip (ResNet 3x3/2: 0.23 ms with the loop).
template <int KH, int KW>
__global__ void __launch_bounds__(256) maxpool_fwd_fixed_kernel(const b200_bf16* __restrict__ x, b200_bf16* __restrict__ y,
                                                                uint8_t* __restrict__ idx, b200_pool_geom g) {
  const int cv = g.c / 8;
  const uint32_t total = (uint32_t)g.n * g.p * g.q * cv;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cv);
    uint32_t t = i / cv;
    const int q = (int)(t % g.q);
    t /= g.q;
    const int p = (int)(t % g.p);
    const int n = (int)(t / g.p);
    const int h0 = p * g.sh - g.pad_t, w0 = q * g.sw - g.pad_l;
    bf16x8 v[KH * KW];
    bool ok[KH * KW];
#pragma unroll
    for (int r = 0; r < KH; ++r)
#pragma unroll
      for (int s = 0; s < KW; ++s) {
        const int h = h0 + r, w = w0 + s;
        ok[r * KW + s] = h >= 0 && h < g.h && w >= 0 && w < g.w;
        if (ok[r * KW + s]) v[r * KW + s] = ld8(x + (((n * g.h + h) * g.w + w) * g.c + c8 * 8));   // 32-bit offsets (host-checked)
      }
    // running maximum and its tap index, both kept packed (two bf16 / two 16-bit indices per register): one packed
    // compare and two bit-selects per channel pair and tap, no conversions. Strict '>' keeps the first maximum.
    uint32_t best[4], arg[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      best[k] = 0xff80ff80u;      // (-inf, -inf)
      arg[k] = 0u;
    }
#pragma unroll
    for (int j = 0; j < KH * KW; ++j) {
      if (!ok[j]) continue;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t xv = *reinterpret_cast<const uint32_t*>(&v[j].v[k]);
        const uint32_t gt = __hgt2_mask(v[j].v[k], *reinterpret_cast<const __nv_bfloat162*>(&best[k]));   // 0xffff per lane
        best[k] = (xv & gt) | (best[k] & ~gt);
        arg[k] = ((uint32_t)j * 0x00010001u & gt) | (arg[k] & ~gt);
      }
    }
    *reinterpret_cast<uint4*>(y + (int64_t)i * 8) = make_uint4(best[0], best[1], best[2], best[3]);
    uint2 pk;                                          // index bytes 0 and 2 of each register pair
    pk.x = __byte_perm(arg[0], arg[1], 0x6420);
    pk.y = __byte_perm(arg[2], arg[3], 0x6420);
    *reinterpret_cast<uint2*>(idx + (int64_t)i * 8) = pk;
  }
}

template <typename Idx>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const b200_bf16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                          b200_bf16* __restrict__ dx, b200_pool_geom g) {
  const int cv = g.c / 8;
  const Idx total = (Idx)g.n * g.h * g.w * cv;
  for (Idx i = (Idx)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (Idx)gridDim.x * blockDim.x) {
    int c8 = (int)(i % cv);
    Idx t = i / cv;
    int w = (int)(t % g.w);
    t /= g.w;
    int h = (int)(t % g.h);
    int n = (int)(t / g.h);
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    // outputs whose window covers (h, w): p*sh - pad_t <= h < p*sh - pad_t + kh
    int p_lo = (h + g.pad_t - g.kh + g.sh) / g.sh;
    if (h + g.pad_t - g.kh + 1 <= 0) p_lo = 0;
    int p_hi = (h + g.pad_t) / g.sh;
    if (p_hi > g.p - 1) p_hi = g.p - 1;
    int q_lo = (w + g.pad_l - g.kw + g.sw) / g.sw;
    if (w + g.pad_l - g.kw + 1 <= 0) q_lo = 0;
    int q_hi = (w + g.pad_l) / g.sw;
    if (q_hi > g.q - 1) q_hi = g.q - 1;
    for (int p = p_lo; p <= p_hi; ++p) {
      int r = h - (p * g.sh - g.pad_t);
      if (r < 0 || r >= g.kh) continue;
      for (int q = q_lo; q <= q_hi; ++q) {
        int s = w - (q * g.sw - g.pad_l);
        if (s < 0 || s >= g.kw) continue;
        int64_t o = ((((int64_t)n * g.p + p) * g.q + q) * cv + c8) * 8;
        uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
        float f[8];
        unpack8(ld8(dy + o), f);
        int me = r * g.kw + s;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          int a = ((k < 4 ? pk.x : pk.y) >> (8 * (k & 3))) & 0xff;
          if (a == me) acc[k] += f[k];
        }
      }
    }
    st8(dx + (int64_t)i * 8, pack8(acc));
  }
}

// Backward of max pooling for stride SH x SW (compile-time): one thread owns an SH x SW block of input pixels (8
// channels) and visits every output whose window can touch the block once, routing each channel's gradient to the
// input position its saved argmax names. Against the one-input-pixel-per-thread kernel above this reads each
// (dy, argmax) pair SH*SW times less often (ResNet's 3x3/2 pool: 0.49 ms -> L2 traffic cut 4x).
template <int SH, int SW>
__global__ void __launch_bounds__(256) maxpool_bwd_block_kernel(const b200_bf16* __restrict__ dy,
                                                                const uint8_t* __restrict__ idx,
                                                                b200_bf16* __restrict__ dx, b200_pool_geom g) {
  const int cv = g.c / 8;
  const int hb = (g.h + SH - 1) / SH, wb = (g.w + SW - 1) / SW;
  const uint32_t total = (uint32_t)g.n * hb * wb * cv;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cv);
    uint32_t t = i / cv;
    const int bw = (int)(t % wb);
    t /= wb;
    const int bh = (int)(t % hb);
    const int n = (int)(t / hb);
    const int h0 = bh * SH, w0 = bw * SW;
    float acc[SH][SW][8];
#pragma unroll
    for (int a = 0; a < SH; ++a)
#pragma unroll
      for (int b = 0; b < SW; ++b)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[a][b][k] = 0.0f;
    // outputs p with p*SH - pad_t <= h0 + SH - 1 and p*SH - pad_t + kh - 1 >= h0
    int p_lo = (h0 + g.pad_t - g.kh + SH) / SH;
    if (h0 + g.pad_t - g.kh + 1 <= 0) p_lo = 0;
    int p_hi = (h0 + SH - 1 + g.pad_t) / SH;
    if (p_hi > g.p - 1) p_hi = g.p - 1;
    int q_lo = (w0 + g.pad_l - g.kw + SW) / SW;
    if (w0 + g.pad_l - g.kw + 1 <= 0) q_lo = 0;
    int q_hi = (w0 + SW - 1 + g.pad_l) / SW;
    if (q_hi > g.q - 1) q_hi = g.q - 1;
    for (int p = p_lo; p <= p_hi; ++p) {
      const int dh0 = p * SH - g.pad_t - h0;           // block-relative row of the window's first tap
      for (int q = q_lo; q <= q_hi; ++q) {
        const int dw0 = q * SW - g.pad_l - w0;
        const int64_t o = ((((int64_t)n * g.p + p) * g.q + q) * cv + c8) * 8;
        const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
        float f[8];
        unpack8(ld8(dy + o), f);
        // the tap of this output's window that lands on each pixel of the block (thread-uniform scalars; 255 = none),
        // so that per channel only byte compares and predicated adds remain
        int tgt[SH][SW];
#pragma unroll
        for (int aa = 0; aa < SH; ++aa)
#pragma unroll
          for (int bb = 0; bb < SW; ++bb) {
            const int r = aa - dh0, s2 = bb - dw0;
            tgt[aa][bb] = (r >= 0 && r < g.kh && s2 >= 0 && s2 < g.kw) ? r * g.kw + s2 : 255;
          }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int a = ((k < 4 ? pk.x : pk.y) >> (8 * (k & 3))) & 0xff;
#pragma unroll
          for (int aa = 0; aa < SH; ++aa)
#pragma unroll
            for (int bb = 0; bb < SW; ++bb)
              if (a == tgt[aa][bb]) acc[aa][bb][k] += f[k];
        }
      }
    }
#pragma unroll
    for (int aa = 0; aa < SH; ++aa)
#pragma unroll
      for (int bb = 0; bb < SW; ++bb) {
        const int h = h0 + aa, w = w0 + bb;
        if (h < g.h && w < g.w) st8(dx + ((((int64_t)n * g.h + h) * g.w + w) * cv + c8) * 8, pack8(acc[aa][bb]));
      }
  }
}

__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const b200_bf16* __restrict__ x, b200_bf16* __restrict__ y,
                                                          b200_pool_geom g) {
  const int cv = g.c / 8;
  const int64_t total = (int64_t)g.n * g.p * g.q * cv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c8 = (int)(i % cv);
    int64_t t = i / cv;
    int q = (int)(t % g.q);
    t /= g.q;
    int p = (int)(t % g.p);
    int n = (int)(t / g.p);
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    int cnt = 0;
    const int h0 = p * g.sh - g.pad_t, w0 = q * g.sw - g.pad_l;
    for (int r = 0; r < g.kh; ++r) {
      int h = h0 + r;
      if (h < 0 || h >= g.h) continue;
      for (int s = 0; s < g.kw; ++s) {
        int w = w0 + s;
        if (w < 0 || w >= g.w) continue;
        float f[8];
        unpack8(ld8(x + (((int64_t)n * g.h + h) * g.w + w) * g.c + c8 * 8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
        ++cnt;
      }
    }
    float inv = 1.0f / (float)(cnt > 0 ? cnt : 1);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] *= inv;
    st8(y + i * 8, pack8(acc));
  }
}

__device__ __forceinline__ int taps_inside(int o, int stride, int pad, int k, int size) {
  int lo = o * stride - pad, hi = lo + k;
  if (lo < 0) lo = 0;
  if (hi > size) hi = size;
  return hi - lo;
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const b200_bf16* __restrict__ dy, b200_bf16* __restrict__ dx,
                                                          b200_pool_geom g) {
  const int cv = g.c / 8;
  const int64_t total = (int64_t)g.n * g.h * g.w * cv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c8 = (int)(i % cv);
    int64_t t = i / cv;
    int w = (int)(t % g.w);
    t /= g.w;
    int h = (int)(t % g.h);
    int n = (int)(t / g.h);
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    int p_hi = (h + g.pad_t) / g.sh;
    if (p_hi > g.p - 1) p_hi = g.p - 1;
    int q_hi = (w + g.pad_l) / g.sw;
    if (q_hi > g.q - 1) q_hi = g.q - 1;
    for (int p = p_hi; p >= 0; --p) {
      int r = h - (p * g.sh - g.pad_t);
      if (r >= g.kh) break;
      int ch = taps_inside(p, g.sh, g.pad_t, g.kh, g.h);
      for (int q = q_hi; q >= 0; --q) {
        int s = w - (q * g.sw - g.pad_l);
        if (s >= g.kw) break;
        int cw = taps_inside(q, g.sw, g.pad_l, g.kw, g.w);
        float f[8];
        unpack8(ld8(dy + ((((int64_t)n * g.p + p) * g.q + q) * cv + c8) * 8), f);
        float inv = 1.0f / (float)(ch * cw);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k] * inv;
      }
    }
    st8(dx + i * 8, pack8(acc));
  }
}

// x[n, hw, c] -> y[n, c]: block = (c/8 vectors) x (row lanes); rows folded through shared memory
__global__ void __launch_bounds__(256) spatial_mean_fwd_kernel(const b200_bf16* __restrict__ x, b200_bf16* __restrict__ y,
                                                               int hw, int c) {
  __shared__ float sm[256 * 8];
  const int cv = c / 8;
  const int tx = cv < 256 ? cv : 256;
  const int ty = 256 / tx;
  const int lx = threadIdx.x % tx, ly = threadIdx.x / tx;
  const int n = blockIdx.x;
  const int v = blockIdx.y * tx + lx;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
  if (ly < ty && v < cv) {
    for (int r = ly; r < hw; r += ty) {
      float f[8];
      unpack8(ld8_stream(x + ((int64_t)n * hw + r) * c + v * 8), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) sm[k * 256 + threadIdx.x] = acc[k];
  __syncthreads();
  if (ly == 0 && v < cv) {
    float o[8];
    float inv = 1.0f / (float)hw;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float s = 0.0f;
      for (int yy = 0; yy < ty; ++yy) s += sm[k * 256 + yy * tx + lx];
      o[k] = s * inv;
    }
    st8(y + (int64_t)n * c + v * 8, pack8(o));
  }
}

__global__ void __launch_bounds__(256) spatial_mean_bwd_kernel(const b200_bf16* __restrict__ dy, b200_bf16* __restrict__ dx,
                                                               int64_t nvec, int hw, int cv) {
  const float inv = 1.0f / (float)hw;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    int c8 = (int)(i % cv);
    int64_t n = i / cv / hw;
    float f[8];
    unpack8(ld8(dy + (n * cv + c8) * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] *= inv;
    st8(dx + i * 8, pack8(f));
  }
}

int check_geom(const b200_pool_geom* g, const char* who) {
  B200_CHECK_ARG(g != nullptr, "%s: null geometry", who);
  B200_CHECK_ARG(g->c > 0 && g->c % 8 == 0, "%s: need c%%8==0 (c=%d)", who, g->c);
  B200_CHECK_ARG(g->n > 0 && g->h > 0 && g->w > 0 && g->p > 0 && g->q > 0 && g->kh > 0 && g->kw > 0 && g->sh > 0 &&
                     g->sw > 0 && g->kh * g->kw <= 255,
                 "%s: bad geometry", who);
  return 0;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_maxpool_fwd(const b200_bf16* x, b200_bf16* y, uint8_t* idx, const b200_pool_geom* g, void* stream) {
  if (int e = check_geom(g, "maxpool_fwd")) return e;
  B200_CHECK_ARG(x && y && idx, "maxpool_fwd: null pointer");
  int64_t total = (int64_t)g->n * g->p * g->q * (g->c / 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool small = total < (1ll << 31) && (int64_t)g->n * g->h * g->w * g->c < (1ll << 31);   // 32-bit element offsets
  if (small && g->kh == 3 && g->kw == 3) {
    maxpool_fwd_fixed_kernel<3, 3><<<grid_for(total, 256), 256, 0, st>>>(x, y, idx, *g);
    B200_LAUNCH_CHECK();
    return 0;
  }
  if (small && g->kh == 2 && g->kw == 2) {
    maxpool_fwd_fixed_kernel<2, 2><<<grid_for(total, 256), 256, 0, st>>>(x, y, idx, *g);
    B200_LAUNCH_CHECK();
    return 0;
  }
  if (total < (1ll << 31))
    maxpool_fwd_kernel<uint32_t><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, idx, *g);
  else
    maxpool_fwd_kernel<int64_t><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, idx, *g);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_maxpool_bwd(const b200_bf16* dy, const uint8_t* idx, b200_bf16* dx, const b200_pool_geom* g, void* stream) {
  if (int e = check_geom(g, "maxpool_bwd")) return e;
  B200_CHECK_ARG(dy && dx && idx, "maxpool_bwd: null pointer");
  int64_t total = (int64_t)g->n * g->h * g->w * (g->c / 8);
  if (g->sh == 2 && g->sw == 2 && total < (1ll << 31) && g->kh * g->kw <= 255) {
    const int64_t blocks = (int64_t)g->n * ((g->h + 1) / 2) * ((g->w + 1) / 2) * (g->c / 8);
    maxpool_bwd_block_kernel<2, 2><<<grid_for(blocks, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, idx, dx, *g);
    B200_LAUNCH_CHECK();
    return 0;
  }
  if (total < (1ll << 31))
    maxpool_bwd_kernel<uint32_t><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, idx, dx, *g);
  else
    maxpool_bwd_kernel<int64_t><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, idx, dx, *g);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_avgpool_fwd(const b200_bf16* x, b200_bf16* y, const b200_pool_geom* g, void* stream) {
  if (int e = check_geom(g, "avgpool_fwd")) return e;
  B200_CHECK_ARG(x && y, "avgpool_fwd: null pointer");
  int64_t total = (int64_t)g->n * g->p * g->q * (g->c / 8);
  avgpool_fwd_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, *g);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_avgpool_bwd(const b200_bf16* dy, b200_bf16* dx, const b200_pool_geom* g, void* stream) {
  if (int e = check_geom(g, "avgpool_bwd")) return e;
  B200_CHECK_ARG(dy && dx, "avgpool_bwd: null pointer");
  int64_t total = (int64_t)g->n * g->h * g->w * (g->c / 8);
  avgpool_bwd_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, dx, *g);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_spatial_mean_fwd(const b200_bf16* x, b200_bf16* y, int32_t n, int32_t hw, int32_t c, void* stream) {
  B200_CHECK_ARG(x && y && n > 0 && hw > 0 && c > 0 && c % 8 == 0, "spatial_mean_fwd: bad args (c=%d)", c);
  int cv = c / 8, tx = cv < 256 ? cv : 256;
  dim3 grid(n, (cv + tx - 1) / tx);
  spatial_mean_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, hw, c);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_spatial_mean_bwd(const b200_bf16* dy, b200_bf16* dx, int32_t n, int32_t hw, int32_t c, void* stream) {
  B200_CHECK_ARG(dy && dx && n > 0 && hw > 0 && c > 0 && c % 8 == 0, "spatial_mean_bwd: bad args (c=%d)", c);
  int64_t nvec = (int64_t)n * hw * (c / 8);
  spatial_mean_bwd_kernel<<<grid_for(nvec, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, dx, nvec, hw, c / 8);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
