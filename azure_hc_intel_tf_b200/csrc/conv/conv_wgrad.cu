// conv_wgrad.cu — convolution backward-weights (and dense-layer weight gradient) on tcgen05 tensor cores.
//
// Role in the reference stack: MKL-DNN convolution backward-weights (K3 in SURVEY.md §2.6); its output is the
// gradient tensor Horovod enqueues for the fused allreduce, so this kernel writes fp32 straight into the
// fusion-buffer view of the parameter's gradient (KRSC layout).
//
// GEMM view (the reduction runs over output pixels):
//     D[(tap, c), k] = sum_pix  X_im2col[pix, (tap, c)] * dY[pix, k]
//   A = im2col(x): one TMA im2col box = 64 pixels x 64 channels of one filter tap  -> MN-major A operand
//   B = dy       : one TMA tiled  box = 64 pixels x 64 output channels            -> MN-major B operand
// Both operands are MN-major (the 128-byte swizzled row holds 64 M/N elements, rows are K = pixels), which is
// exactly how NHWC activations sit in memory, so no transposed copies are made.
//
// Work split: a CTA owns a slice of the pixel range and a group of up to T M-tiles (128 (tap,c) rows each) times
// one N-tile (<= 256 output channels); T*N <= 512 TMEM columns hold all its fp32 accumulators. Partial sums of the
// pixel slices are combined with red.global.add.f32 into dw (zeroed first when more than one slice exists), or
// stored directly when a single slice covers all pixels. Stores are coalesced: a warp's 32 lanes are 32
// consecutive (tap,c) positions of one output channel.
#include "b200conv.h"
#include "common.cuh"
#include "conv/tensormap.h"
#include "conv/umma.cuh"

namespace b200 {
namespace {

using namespace umma;

constexpr int kThreads = 320;        // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kBlockK = 64;          // pixels per pipeline stage
constexpr int kBox = 64 * 128;       // one 64 x 64 bf16 TMA box, bytes
constexpr int kTmemCols = 512;
constexpr int kMaxStages = 6;
constexpr int kHaloCols = 16, kHaloRows = 18;
constexpr int kHaloBytes = kHaloRows * kHaloCols * 128;   // one 64-channel halo plane
constexpr int kDyTileBytes = 128 * 128;                   // 16 x 8 pixels x 64 output channels

struct WgradParams {
  int rows_total;        // R*S*C  (valid (tap,c) rows, = row length of dw)
  int C, S, RS;
  int Cp;                // C rounded up to 64: the M index space is (tap, c') with c' < Cp; c' >= C is zero padding
  int kout;              // output channels
  int block_n;           // N tile (multiple of 16, <= 256)
  int n_tiles;           // ceil(kout / block_n)
  int m_tiles_total;     // ceil(rows_total / 128)
  int T;                 // M tiles per CTA
  int m_groups;          // ceil(m_tiles_total / T)
  int slices;            // pixel slices
  int64_t pix_total;
  int kblocks_total;     // ceil(pix_total / 64)
  int kblocks_per_slice;
  int PQ, Q, stride_h, stride_w, base_h, base_w;
  int stages;
  int dy_one_shot;       // dy planes of a stage come with one 3-D TMA instruction instead of one per 64 channels
  // halo mode (3x3, stride 1, pad 1): a k-block is a 16 x 8 tile of output pixels; the input halo of each 64-channel
  // plane is loaded once and the taps are shifted windows of it. M index space = (channel plane, tap) in 64-row units.
  int halo, tiles_x, tiles_y, cchunks;
  int atomic;            // 1: red.add into dw, 0: plain stores (to dw, or to this slice's workspace partial)
  float* dw;
  float* ws;             // when non-null: partial of slice s goes to ws + s * out_elems, a second kernel folds them
  int64_t out_elems;     // kout * rows_total
};

struct __align__(8) WBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                  const WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nb = (p.block_n + 63) / 64;                 // dy boxes per stage
  // halo mode: at most two channel planes are touched by the 2T consecutive (plane, tap) positions of one CTA
  const int a_bytes = p.halo ? 2 * kHaloBytes : p.T * 2 * kBox;
  const int stage_bytes = a_bytes + nb * (p.halo ? kDyTileBytes : kBox);
  WBarriers* bars = reinterpret_cast<WBarriers*>(smem + (size_t)p.stages * stage_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // blockIdx -> (slice, n tile, m group); groups of one slice are adjacent so they share dy / x through L2
  const int per_slice = p.m_groups * p.n_tiles;
  const int slice = blockIdx.x / per_slice;
  const int rem = blockIdx.x - slice * per_slice;
  const int nt = rem / p.m_groups, mg = rem - nt * p.m_groups;
  const int n0 = nt * p.block_n;
  const int tile0 = mg * p.T;
  int tiles_here = p.m_tiles_total - tile0;
  if (tiles_here > p.T) tiles_here = p.T;
  const int kb0 = slice * p.kblocks_per_slice;
  int kb1 = kb0 + p.kblocks_per_slice;
  if (kb1 > p.kblocks_total) kb1 = p.kblocks_total;
  const int nkb = kb1 - kb0;        // >= 1 by construction
  const bool dual = p.block_n <= 128;                        // two issuing warps (see issue_halo / issue_generic)
  const int t_split = dual ? (tiles_here + 1) / 2 : tiles_here;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_x);
    prefetch_tensormap(&map_dy);
    for (int i = 0; i < p.stages; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], dual ? 2 : 1);
    }
    mbar_init(&bars->tmem_full, dual ? 2 : 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  // halo mode: channel planes covered by this CTA's positions [2*tile0, 2*(tile0+tiles_here))
  const int pos_lo = 2 * tile0;
  int pos_hi = 2 * (tile0 + tiles_here) - 1;
  if (pos_hi > 9 * p.cchunks - 1) pos_hi = 9 * p.cchunks - 1;
  const int plane_lo = p.halo ? pos_lo / 9 : 0;
  const int nplanes = p.halo ? pos_hi / 9 - plane_lo + 1 : 0;

  // MMA issue. Tiles of at most 128 columns are bound by the issue rate of one thread (~80 cycles per tcgen05
  // instruction against 32-64 cycles of math, scripts/micro/mma_rate.cu), so for those a second warp issues too: the
  // M tiles of this CTA (each with its own accumulator columns) are split between warp 1 and warp 2; both wait on the
  // same full barriers and both commit to empty / tmem_full (count 2). Warp 2 joins the epilogue afterwards.
  auto issue_halo = [&](int t_lo, int t_hi) {
    const uint32_t idesc = make_idesc_bf16(128, (uint32_t)p.block_n, 1, 1);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < nkb; ++kb) {
      mbar_wait(&bars->full[stage], phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = smem_u32(smem + (size_t)stage * stage_bytes);
        const uint32_t sb = sa + a_bytes;
        for (int t = t_lo; t < t_hi; ++t) {
          // the two 64-row halves of M tile t: positions 2t and 2t+1 of the (plane, tap) space
          const int p0 = 2 * (tile0 + t);
          int p1 = p0 + 1;
          if (p1 > 9 * p.cchunks - 1) p1 = p0;          // odd tail: second half is masked in the epilogue
          const int pl0 = p0 / 9, tap0 = p0 - pl0 * 9, pl1 = p1 / 9, tap1 = p1 - pl1 * 9;
          const uint32_t off0 = (uint32_t)((pl0 - plane_lo) * kHaloBytes + ((tap0 / 3) * kHaloCols + tap0 % 3) * 128);
          const uint32_t off1 = (uint32_t)((pl1 - plane_lo) * kHaloBytes + ((tap1 / 3) * kHaloCols + tap1 % 3) * 128);
          const uint32_t lbo = off1 > off0 ? off1 - off0 : 16;
          // descriptor low words stepped with adds (the issue rate of this thread bounds the 64-wide tiles);
          // shifted windows need no descriptor fix-up: the 128-byte swizzle follows the smem address (measured)
          const uint32_t a_lo = sdesc_lo(sa + off0, lbo), a_hi = sdesc_hi_sw128(kHaloCols * 128);
          const uint32_t b_lo = sdesc_lo(sb, kDyTileBytes), b_hi = sdesc_hi_sw128(1024);
          const uint32_t d_t = tmem_base + (uint32_t)(t * p.block_n);
#pragma unroll
          for (int k = 0; k < 8; ++k)                   // 16 pixels per step = output rows 2k, 2k+1 of the tile
            mma_bf16_lohi(d_t, a_lo + k * ((2 * kHaloCols * 128) >> 4), a_hi, b_lo + k * (2048 >> 4), b_hi, idesc,
                          (kb | k) != 0 ? 1u : 0u);
        }
        mma_commit(&bars->empty[stage]);
        if (kb == nkb - 1) mma_commit(&bars->tmem_full);
      }
      __syncwarp();
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
    }
  };
  auto issue_generic = [&](int t_lo, int t_hi) {
    const uint32_t idesc = make_idesc_bf16(128, (uint32_t)p.block_n, 1, 1);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < nkb; ++kb) {
      mbar_wait(&bars->full[stage], phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = smem_u32(smem + (size_t)stage * stage_bytes);
        const uint32_t sb = sa + a_bytes;
        const uint32_t a_lo = sdesc_lo(sa, kBox), b_lo = sdesc_lo(sb, kBox), d_hi = sdesc_hi_sw128(1024);
#pragma unroll
        for (int k = 0; k < kBlockK / 16; ++k) {
          for (int t = t_lo; t < t_hi; ++t)
            mma_bf16_lohi(tmem_base + (uint32_t)(t * p.block_n), a_lo + (uint32_t)((t * 2 * kBox + k * 2048) >> 4), d_hi,
                          b_lo + k * (2048 >> 4), d_hi, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        mma_commit(&bars->empty[stage]);
        if (kb == nkb - 1) mma_commit(&bars->tmem_full);
      }
      __syncwarp();
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
    }
  };

  if (warp == 0 && p.halo) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = (uint32_t)(nplanes * kHaloBytes + nb * kDyTileBytes);
      const int per_img = p.tiles_x * p.tiles_y;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&bars->empty[stage], phase ^ 1);
        uint8_t* sa = smem + (size_t)stage * stage_bytes;
        uint8_t* sb = sa + a_bytes;
        mbar_expect_tx(&bars->full[stage], tx_bytes);
        const int img = kb / per_img, r2 = kb - img * per_img;
        const int ty = r2 / p.tiles_x, tx = r2 - ty * p.tiles_x;
        for (int pl = 0; pl < nplanes; ++pl)
          tma_load_4d(sa + pl * kHaloBytes, &map_x, &bars->full[stage], (plane_lo + pl) * 64, tx * 8 - 1, ty * 16 - 1, img);
        for (int j = 0; j < nb; ++j)
          tma_load_4d(sb + j * kDyTileBytes, &map_dy, &bars->full[stage], n0 + j * 64, tx * 8, ty * 16, img);
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = (uint32_t)(tiles_here * 2 * kBox + nb * kBox);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&bars->empty[stage], phase ^ 1);
        uint8_t* sa = smem + (size_t)stage * stage_bytes;
        uint8_t* sb = sa + a_bytes;
        mbar_expect_tx(&bars->full[stage], tx_bytes);
        const int64_t pix = (int64_t)kb * kBlockK;
        const int img = (int)(pix / p.PQ);
        const int r2 = (int)(pix - (int64_t)img * p.PQ);
        const int pp = r2 / p.Q, qq = r2 - pp * p.Q;
        const int hh = pp * p.stride_h + p.base_h, ww = qq * p.stride_w + p.base_w;
        for (int t = 0; t < tiles_here; ++t) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            int idx = (tile0 + t) * 128 + half * 64;      // first (tap,c') row of this 64-row chunk
            int tap = idx / p.Cp;
            int c0 = idx - tap * p.Cp;
            if (tap >= p.RS) {                            // past the last tap: load something valid, it is masked
              tap = p.RS - 1;
              c0 = p.C;                                   // fully out of bounds in C -> zero fill
            }
            const int fr = tap / p.S, fs = tap - fr * p.S;
            tma_load_im2col_4d(sa + (t * 2 + half) * kBox, &map_x, &bars->full[stage], c0, ww, hh, img, (uint16_t)fs,
                               (uint16_t)fr);
          }
        }
        if (p.dy_one_shot) {
          tma_load_3d(sb, &map_dy, &bars->full[stage], 0, (int32_t)pix, n0 / 64);
        } else {
          for (int j = 0; j < nb; ++j)
            tma_load_2d(sb + j * kBox, &map_dy, &bars->full[stage], n0 + j * 64, (int32_t)pix);
        }
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (p.halo) issue_halo(0, t_split);
    else issue_generic(0, t_split);
  } else {
    if (dual && warp == 2) {
      if (p.halo) issue_halo(t_split, tiles_here);
      else issue_generic(t_split, tiles_here);
    }
    const int lane_group = warp & 3;            // TMEM lanes [32*lane_group, +32) are visible to this warp
    const int col_half = (warp - 2) >> 2;       // eight epilogue warps: two per lane group
    mbar_wait_backoff(&bars->tmem_full, 0);
    tc_fence_after();
    const size_t ld = (size_t)p.rows_total;       // dw[k][rows_total]
    float* const outp = p.ws ? p.ws + (size_t)slice * p.out_elems : p.dw;
    for (int t = 0; t < tiles_here; ++t) {
      const int pidx = (tile0 + t) * 128 + lane_group * 32 + lane;   // position in the padded (tap, c') space
      int ptap = pidx / p.Cp, pc = pidx - ptap * p.Cp;
      bool valid = ptap < p.RS && pc < p.C;
      if (p.halo) {
        // halo mode orders the M space (plane, tap): 64-row unit u = plane*9 + tap
        const int u = pidx >> 6;
        const int pl = u / 9;
        ptap = u - pl * 9;
        pc = pl * 64 + (pidx & 63);
        valid = u < 9 * p.cchunks;
      }
      const int idx = ptap * p.C + pc;                                // position in a dw row
      const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)(t * p.block_n);
      // the two warps of a lane group take alternate 16-column chunks
      for (int c = col_half * 16; c < p.block_n; c += 32) {
        uint32_t r[16];
        tmem_ld_x16(t_row + (uint32_t)c, r);
        tmem_ld_wait();
        if (!valid) continue;
        float* d = outp + (size_t)(n0 + c) * ld + idx;
        if (!p.atomic && n0 + c + 16 <= p.kout) {
          // fast path: 16 plain stores, one pointer bump each (a warp's store covers one 128-byte line)
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            *d = __uint_as_float(r[j]);
            d += ld;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (n0 + c + j < p.kout) {
              if (p.atomic)
                asm volatile("red.global.add.f32 [%0], %1;" ::"l"(d), "f"(__uint_as_float(r[j])) : "memory");
              else
                *d = __uint_as_float(r[j]);
            }
            d += ld;
          }
        }
      }
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// dw[e] = sum over slices of ws[s][e] in a fixed order (bitwise reproducible): a block is 32 float4 columns x 8 slice
// lanes; lane y adds slices y, y+8, ... in order, then the eight lane sums are added in lane order. The slice lanes
// keep enough loads in flight that small filters (few columns, many slices) are not latency-bound.
__global__ void __launch_bounds__(256) wgrad_fold_kernel(const float* __restrict__ ws, float* __restrict__ dw,
                                                         int64_t n, int slices) {
  __shared__ float4 part[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t i = ((int64_t)blockIdx.x * 32 + tx) * 4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i + 4 <= n) {
#pragma unroll 4
    for (int s2 = ty; s2 < slices; s2 += 8) {
      const float4 b = __ldcs(reinterpret_cast<const float4*>(ws + (size_t)s2 * n + i));
      a.x += b.x;
      a.y += b.y;
      a.z += b.z;
      a.w += b.w;
    }
  } else if (i < n) {
    float* av = &a.x;
    for (int j = 0; j < (int)(n - i); ++j)
      for (int s2 = ty; s2 < slices; s2 += 8) av[j] += ws[(size_t)s2 * n + i + j];
  }
  part[ty][tx] = a;
  __syncthreads();
  if (ty == 0 && i < n) {
#pragma unroll
    for (int y = 1; y < 8; ++y) {
      const float4 b = part[y][tx];
      a.x += b.x;
      a.y += b.y;
      a.z += b.z;
      a.w += b.w;
    }
    if (i + 4 <= n) {
      *reinterpret_cast<float4*>(dw + i) = a;
    } else {
      const float* av = &a.x;
      for (int j = 0; j < (int)(n - i); ++j) dw[i + j] = av[j];
    }
  }
}

constexpr int64_t kWorkspaceCap = 96ll << 20;

bool wgrad_halo_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_HALO");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

int plan(const b200_conv_geom* g, WgradParams& p) {
  p.halo = (wgrad_halo_enabled() && g->r == 3 && g->s == 3 && g->stride_h == 1 && g->stride_w == 1 && g->pad_t == 1 &&
            g->pad_l == 1 && g->pad_b == 1 && g->pad_r == 1 && g->c % 64 == 0 && g->c <= 128 && g->k % 64 == 0 &&
            g->k <= 128 && g->h >= 16 && g->w >= 8)
               ? 1
               : 0;
  p.cchunks = g->c / 64;
  p.tiles_y = (g->h + 15) / 16;
  p.tiles_x = (g->w + 7) / 8;
  p.C = g->c;
  p.Cp = (g->c + 63) / 64 * 64;
  p.S = g->s;
  p.RS = g->r * g->s;
  p.rows_total = p.RS * g->c;
  p.kout = g->k;
  p.block_n = g->k >= 256 ? 256 : (g->k + 15) / 16 * 16;
  p.n_tiles = (g->k + p.block_n - 1) / p.block_n;
  p.m_tiles_total = (p.RS * p.Cp + 127) / 128;   // same count in halo mode: 9 * cchunks 64-row units
  const int nb = (p.block_n + 63) / 64;
  int T = kTmemCols / p.block_n;
  // two pipeline stages must fit in ~200 KiB: (T*16 KiB + nb*8 KiB) * 2
  while (!p.halo && T > 1 && (T * 2 * kBox + nb * kBox) * 2 > 200 * 1024) --T;
  if (T > p.m_tiles_total) T = p.m_tiles_total;
  p.T = T;
  p.m_groups = (p.m_tiles_total + T - 1) / T;
  p.pix_total = (int64_t)g->n * g->p * g->q;
  p.kblocks_total = p.halo ? g->n * p.tiles_y * p.tiles_x : (int)((p.pix_total + kBlockK - 1) / kBlockK);
  const int groups = p.m_groups * p.n_tiles;
  p.out_elems = (int64_t)g->k * p.rows_total;
  // one wave of CTAs: as many pixel slices as fit next to the output groups, bounded by the work per slice and by
  // the size of the partial-sum workspace
  int slices = kNumSMs / groups;
  int max_slices = p.kblocks_total / 4;
  if (max_slices < 1) max_slices = 1;
  if (slices > max_slices) slices = max_slices;
  const int64_t by_ws = kWorkspaceCap / (p.out_elems * 4);
  if (slices > by_ws) slices = (int)by_ws;
  if (slices < 1) slices = 1;
  p.kblocks_per_slice = (p.kblocks_total + slices - 1) / slices;
  p.slices = (p.kblocks_total + p.kblocks_per_slice - 1) / p.kblocks_per_slice;
  p.atomic = p.slices > 1;
  const int stage_bytes = p.halo ? 2 * kHaloBytes + nb * kDyTileBytes : T * 2 * kBox + nb * kBox;
  int stages = (200 * 1024) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) stages = 2;
  p.stages = stages;
  p.PQ = g->p * g->q;
  p.Q = g->q;
  p.stride_h = g->stride_h;
  p.stride_w = g->stride_w;
  p.base_h = -g->pad_t;
  p.base_w = -g->pad_l;
  return 0;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_conv_min_channels();

int b200_conv_wgrad_supported(const b200_conv_geom* g) {
  if (!g) return 0;
  return g->c >= b200_conv_min_channels() && g->c % 8 == 0 && g->k % 8 == 0 && g->r <= 15 && g->s <= 15 && g->n > 0;
}

int64_t b200_conv_wgrad_workspace_bytes(const b200_conv_geom* g) {
  if (!g || g->c < b200_conv_min_channels() || g->c % 8 != 0) return 0;
  WgradParams p{};
  plan(g, p);
  return p.slices > 1 ? (int64_t)p.slices * p.out_elems * 4 : 0;
}

static int wgrad_impl(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, int dy_ld,
                      void* workspace, void* stream);

int b200_conv_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, void* workspace,
                    void* stream) {
  B200_CHECK_ARG(g, "conv_wgrad: null geometry");
  return wgrad_impl(x, dy, dw, g, g->k, workspace, stream);
}

int64_t b200_dense_wgrad_workspace_bytes(int32_t batch, int32_t in_features, int32_t out_features) {
  b200_conv_geom g{};
  g.n = batch;
  g.h = g.w = g.p = g.q = 1;
  g.c = in_features;
  g.k = out_features;
  g.r = g.s = 1;
  g.stride_h = g.stride_w = 1;
  return b200_conv_wgrad_workspace_bytes(&g);
}

int b200_dense_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, int32_t batch, int32_t in_features,
                     int32_t out_features, int32_t dy_ld, void* workspace, void* stream) {
  B200_CHECK_ARG(dy_ld >= out_features && dy_ld % 8 == 0, "dense_wgrad: dy_ld must be >= out and a multiple of 8");
  b200_conv_geom g{};
  g.n = batch;
  g.h = g.w = g.p = g.q = 1;
  g.c = in_features;
  g.k = out_features;
  g.r = g.s = 1;
  g.stride_h = g.stride_w = 1;
  return wgrad_impl(x, dy, dw, &g, dy_ld, workspace, stream);
}

static int wgrad_impl(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, int dy_ld,
                      void* workspace, void* stream) {
  B200_CHECK_ARG(x && dy && dw && g, "conv_wgrad: null pointer");
  B200_CHECK_ARG(g->c >= b200_conv_min_channels() && g->c % 8 == 0 && dy_ld % 8 == 0 && g->r <= 15 && g->s <= 15 && g->n > 0,
                 "conv_wgrad: unsupported geometry (need Cin%%8==0, dy row stride %%8==0)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  WgradParams p{};
  plan(g, p);
  p.dw = dw;
  CUtensorMap mx, mdy;
  const int upper_w = g->pad_r - (g->s - 1), upper_h = g->pad_b - (g->r - 1);
  if (p.halo) {
    if (tmap_nhwc_tile(&mx, x, g->n, g->h, g->w, g->c, 64, kHaloCols, kHaloRows)) return -1;
    if (tmap_nhwc_tile(&mdy, dy, g->n, g->p, g->q, dy_ld, 64, 8, 16)) return -1;
  } else if (tmap_im2col_nhwc(&mx, x, g->n, g->h, g->w, g->c, -g->pad_l, -g->pad_t, upper_w, upper_h, g->stride_w,
                              g->stride_h, 64, kBlockK))
    return -1;
  const int nb_host = (p.block_n + 63) / 64;
  p.dy_one_shot = 0;
  if (p.halo) {
    // maps were built above
  } else if (dy_ld % 64 == 0 && g->k % 64 == 0 && nb_host > 1 &&
      tmap_chunked_rows(&mdy, dy, (uint64_t)p.pix_total, (uint64_t)dy_ld * 2, (uint32_t)(g->k / 64), kBlockK, nb_host) == 0) {
    p.dy_one_shot = 1;
  } else if (tmap_tiled_2d(&mdy, dy, 2, (uint64_t)dy_ld, (uint64_t)p.pix_total, (uint64_t)dy_ld * 2, 64, kBlockK, true)) {
    return -1;
  }
  // several pixel slices: partials go to the workspace and are folded in a fixed order; without a workspace they
  // are combined with float atomics directly in dw (not bitwise reproducible)
  if (p.slices > 1 && workspace) {
    p.ws = static_cast<float*>(workspace);
    p.atomic = 0;
  } else if (p.atomic) {
    B200_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)p.out_elems, st));
  }
  const int nb = (p.block_n + 63) / 64;
  const size_t smem = (size_t)p.stages * (p.halo ? 2 * kHaloBytes + nb * kDyTileBytes : p.T * 2 * kBox + nb * kBox) +
                      sizeof(WBarriers) + 1024;
  static std::atomic<uint64_t> attr_set{0};
  if (first_use_on_device(attr_set))
    B200_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxConvSmem));
  const int grid = p.slices * p.m_groups * p.n_tiles;
  conv_wgrad_kernel<<<grid, kThreads, smem, st>>>(mx, mdy, p);
  B200_LAUNCH_CHECK();
  if (p.ws) {
    const int64_t vec = (p.out_elems + 3) / 4;
    wgrad_fold_kernel<<<(unsigned)((vec + 31) / 32), 256, 0, st>>>(p.ws, dw, p.out_elems, p.slices);
    B200_LAUNCH_CHECK();
  }
  return 0;
}

}  // extern "C"
