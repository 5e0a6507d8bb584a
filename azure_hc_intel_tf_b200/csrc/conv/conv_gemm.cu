// conv_gemm.cu — implicit-GEMM convolution (forward, backward-data) and dense GEMM on the 5th-generation tensor
// cores: tcgen05.mma with TMEM accumulators, operands staged by TMA (im2col mode for the activation operand), a
// persistent warp-specialised CTA per SM.
//
// Role in the reference stack: MKL-DNN convolution forward / backward-data and inner-product primitives that
// TensorFlow dispatches under --mkl=TRUE (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76); K1, K2, K8 in SURVEY.md.
//
// GEMM view:  D[M, N] = A[M, K] * B[N, K]^T
//   fprop : M = N*P*Q pixels, N = Cout,  K = R*S*Cin;  A = im2col(x)  (TMA im2col),  B = w[Cout][R*S*Cin]  (K-major)
//   dgrad : M = N*H*W pixels, N = Cin,   K = R*S*Cout; A = im2col(dy) (TMA im2col),  B = w viewed [Cout][(r,s),Cin],
//           read as an MN-major operand with the taps flipped - no transposed weight copy is made
//   gemm  : A[M,K] row-major, B[N,K] row-major (both TMA tiled)
//
// CTA layout (576 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected lane),
// warps 2..17 = four epilogue groups of four warps (TMEM -> registers -> bf16 -> swizzled smem -> TMA bulk store, or
// a cooperative scatter for strided dgrad). Pipelines: smem ring (full/empty mbarriers) between TMA and MMA; two TMEM
// accumulator stages (tmem_full/tmem_empty). Groups 2a and 2a+1 drain stage a together, each taking half of the
// tile's 64-column boxes, so the epilogues of tiles i and i+1 overlap each other and the main loop of tile i+2 and
// every SM sub-partition has four epilogue warps to hide the TMEM-load / shared-memory latencies behind each other
// (with two groups the memory-bound 1x1 layers were epilogue-latency-bound at ~40% of DRAM bandwidth).
#include "b200conv.h"
#include "common.cuh"
#include "conv/tensormap.h"
#include "conv/umma.cuh"

namespace b200 {
namespace {

using namespace umma;

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                    // 64 bf16 = 128 bytes = one swizzle row
constexpr int kMaxBlockN = 256;
constexpr int kThreads = 576;                  // warp 0 TMA, warp 1 MMA, warps 2-17: four epilogue groups
constexpr int kABytes = kBlockM * kBlockK * 2;  // 16 KiB
constexpr int kAccStages = 2;
constexpr int kTmemCols = 512;

struct GemmParams {
  int M, N;               // output rows / columns
  int block_n;            // tile width, multiple of 16, <= 256
  int num_m_tiles, num_n_tiles, num_k_blocks;
  int stages;             // smem ring depth
  int a_im2col;           // 1: A through the im2col map, 0: A through a tiled 2-D map
  int PQ, Q;              // output spatial extent, to decompose the first row of a tile into (n, p, q)
  int stride_h, stride_w; // traversal strides
  int base_h, base_w;     // bounding-box lower corner
  int R, S, cchunks;      // filter taps and K/64 chunks per tap
  int b_mn_major;         // dgrad: B is [k rows][n] with the taps flipped
  int b_row_len;          // dgrad: elements between taps in a weight row = Cin
  void* out;              // bf16 (or fp32 when out_fp32) [M][ldo]
  int ldo;
  int out_fp32;
  // strided 1x1 dgrad: GEMM row (n,p,q) lands at (n, p*sy, q*sx) of an H x W image; the other pixels get zeros
  int scatter, sy, sx, H, W;
  // phase form of the scatter (general strided dgrad, one launch per stride phase): the row lands at
  // (n, p*sy + oy, q*sx + ox) and the other pixels of the stride cell are left alone (another phase writes them)
  int oy, ox, scatter_fill;
  float* col_stats;       // optional [gridDim.x*4][2][N]: per (CTA, 32-row group) running column sums / sums of squares
  int tma_store;          // epilogue stages bf16 tiles in shared memory and writes them with TMA bulk stores
  // halo mode (3x3, stride 1, 64 reduction channels): an M tile is a 16 x 8 block of output pixels of one image; its
  // 18 x 16 input halo is loaded once and the nine taps are shifted windows of it, the filter stays resident in smem
  int halo, tiles_x, tiles_y, Himg, Wimg;
  // resident-filter mode (generic path): the whole [K][block_n] filter slice of this CTA's N tile stays in shared
  // memory for the life of the CTA, so the pipeline only streams the activation operand
  int b_resident;
  // debug (B200_CONV_TRACE=1): CTA 0 accumulates the cycles its roles spend waiting on each barrier class
  long long* trace;
  int ctas;               // 1, or 2: CTA pairs (cta_group::2) share 256-row tiles; set by plan_ctas() before map_b is built
};

constexpr int kHaloRows = 18, kHaloCols = 16;
constexpr int kHaloBytes = kHaloRows * kHaloCols * 128;     // 36 KiB: one 64-channel halo tile

constexpr int kStageOutBytes = kBlockM * 64 * 2;   // one 128 x 64 bf16 output box, 128-byte swizzled
constexpr int kOutBufs = 4;                        // one staging box per epilogue group

struct __align__(8) Barriers {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t tmem_full[kAccStages];
  uint64_t tmem_empty[kAccStages];
  uint64_t bres;           // halo mode: the resident filter has landed
  uint32_t tmem_base;
};

__device__ __forceinline__ void store_row_bf16(__nv_bfloat16* dst, const uint32_t r[16]) {
  uint4 v0, v1;
  __nv_bfloat162 t;
#define PK(a, b) (t = __floats2bfloat162_rn(__uint_as_float(a), __uint_as_float(b)), *reinterpret_cast<uint32_t*>(&t))
  v0.x = PK(r[0], r[1]);
  v0.y = PK(r[2], r[3]);
  v0.z = PK(r[4], r[5]);
  v0.w = PK(r[6], r[7]);
  v1.x = PK(r[8], r[9]);
  v1.y = PK(r[10], r[11]);
  v1.z = PK(r[12], r[13]);
  v1.w = PK(r[14], r[15]);
#undef PK
  reinterpret_cast<uint4*>(dst)[0] = v0;
  reinterpret_cast<uint4*>(dst)[1] = v1;
}

__device__ long long g_trace[8];

// timed barrier wait for the trace: slot 0 producer<-empty, 1 MMA<-tmem_empty, 2 MMA<-full, 3 epilogue<-tmem_full
#define TRACED_WAIT(slot, call)                                   \
  do {                                                            \
    if (tr) {                                                     \
      const long long t0_ = clock64();                            \
      call;                                                       \
      if (lane == 0) tr_acc[slot] += clock64() - t0_;             \
    } else {                                                      \
      call;                                                       \
    }                                                             \
  } while (0)

// kCtas == 2: the CTAs of a cluster pair share one 256-row tile (cta_group::2): CTA r stages rows [128 r, +128) of A
// and rows/columns [r * block_n / 2, +block_n / 2) of the B tile; the leader (rank 0) owns the full / tmem_empty
// barriers and issues the MMAs, both CTAs drain their own 128 TMEM lanes.
template <int kCtas, bool kHalo>
__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_out, const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle atoms
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int bn_cta = p.block_n / kCtas;          // B rows (K-major) or columns (MN-major) staged by this CTA
  const int b_bytes = bn_cta * kBlockK * 2;
  // halo mode: [resident filter: 9 taps x b_bytes][stages x halo tile]; otherwise [stages x (A tile, B tile)]
  const int stage_bytes = kHalo ? kHaloBytes : (p.b_resident ? kABytes : kABytes + b_bytes);
  uint8_t* const bres = smem;
  if (kHalo) smem += 9 * b_bytes;
  if (p.b_resident) smem += p.num_k_blocks * b_bytes;
  uint8_t* out_stage = smem + (size_t)p.stages * stage_bytes;      // kOutBufs x 16 KiB when tma_store
  Barriers* bars = reinterpret_cast<Barriers*>(out_stage + (p.tma_store ? kOutBufs * kStageOutBytes : 0));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cr = kCtas == 2 ? cluster_ctarank() : 0u;
  // tiles are (M tile group of kCtas x 128 rows, N tile); a cluster walks them with the stride of the cluster count
  const int num_tiles = ((p.num_m_tiles + kCtas - 1) / kCtas) * p.num_n_tiles;
  const int tile_first = blockIdx.x / kCtas, tile_step = gridDim.x / kCtas;
  const bool tr = p.trace != nullptr && blockIdx.x == 0;
  long long tr_acc[4] = {0, 0, 0, 0};
  const long long tr_start = tr ? clock64() : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_a);
    prefetch_tensormap(&map_b);
    if (p.tma_store) prefetch_tensormap(&map_out);
    for (int i = 0; i < p.stages; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], kHalo ? 2 : 1);      // halo mode has two MMA-issuing warps (see halo_issue)
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&bars->tmem_full[i], kHalo ? 2 : 1);
      // one arrive per epilogue warp that drains the stage: both groups of the pair when the tile has >= 2 boxes
      mbar_init(&bars->tmem_empty[i], ((p.tma_store && p.block_n >= 128) ? 8 : 4) * kCtas);
    }
    mbar_init(&bars->bres, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    if constexpr (kCtas == 2) tmem_alloc_pair<kTmemCols>(&bars->tmem_base);
    else tmem_alloc<kTmemCols>(&bars->tmem_base);
  }
  tc_fence_before();
  // the peer's barriers must be initialised before this CTA's TMA / commits / arrives can touch them
  if constexpr (kCtas == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  // 2-D tiled load; in a pair the bytes are counted on the leader's barrier
  auto load_2d = [](void* dst, const CUtensorMap* map, uint64_t* bar, int32_t c0, int32_t c1) {
    if constexpr (kCtas == 2) tma_load_2d_pair(dst, map, bar, c0, c1);
    else tma_load_2d(dst, map, bar, c0, c1);
  };

  // Halo mode, MMA issue: the 36 MMAs of a tile are 64 columns wide (32 cycles of math each) and one thread cannot
  // issue them faster than ~78 cycles apiece, so TWO warps issue: warp 1 takes taps [0, 5) into accumulator columns
  // [0, 64) of the stage, warp 6 (an epilogue warp that has no box to drain in this mode) takes taps [5, 9) into
  // columns [64, 128); the epilogue adds the two partial accumulators. Both wait on the same full / tmem_empty
  // barriers and both commit to empty / tmem_full (count 2).
  auto halo_issue = [&](int tap_lo, int tap_hi, uint32_t col_off) {
    const uint32_t idesc = make_idesc_bf16(kBlockM, (uint32_t)p.block_n, 0, (uint32_t)p.b_mn_major);
    const uint32_t d_hi = sdesc_hi_sw128(1024), a_hi = sdesc_hi_sw128(kHaloCols * 128);
    const uint32_t b_lbo = p.b_mn_major ? (uint32_t)(kBlockK * 128) : 16u;
    const uint32_t b_kstep = p.b_mn_major ? (2048u >> 4) : (32u >> 4);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    mbar_wait(&bars->bres, 0);
    for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
      mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * kMaxBlockN) + col_off;
      mbar_wait(&bars->full[stage], phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sh = smem_u32(smem + (size_t)stage * stage_bytes);
        const uint32_t sbr = smem_u32(bres);
        for (int tap = tap_lo; tap < tap_hi; ++tap) {
          const int r = tap / 3, s3 = tap - r * 3;
          // A rows of this tap: halo pixel (ty + r, tx + s): 8-pixel groups 16 halo pixels (2 KiB) apart. The
          // 128-byte swizzle is a function of the shared-memory address itself (measured: a window that starts s rows
          // into a TMA-written tile reads correctly with base offset 0), so a shifted start needs no fix-up
          const uint32_t a_lo = sdesc_lo(sh + (uint32_t)((r * kHaloCols + s3) * 128), 16);
          const uint32_t b_lo = sdesc_lo(sbr + (uint32_t)(tap * b_bytes), b_lbo);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            mma_bf16_lohi(d_tmem, a_lo + k * 2, a_hi, b_lo + k * b_kstep, d_hi, idesc, (tap != tap_lo || k != 0) ? 1u : 0u);
        }
        mma_commit(&bars->empty[stage]);
        mma_commit(&bars->tmem_full[acc]);
      }
      __syncwarp();
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
      if (++acc == kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  };

  if (warp == 0) {
    // ================================================================ TMA producer
    if (lane == 0 && kHalo) {
      // filter of this CTA's N tile (gridDim.x is a multiple of num_n_tiles, so the N tile never changes): resident
      const int nt0 = blockIdx.x % p.num_n_tiles, n00 = nt0 * p.block_n;
      mbar_expect_tx(&bars->bres, (uint32_t)(9 * b_bytes));
      for (int tap = 0; tap < 9; ++tap) {
        if (!p.b_mn_major) {
          tma_load_2d(bres + tap * b_bytes, &map_b, &bars->bres, tap * p.b_row_len, n00);
        } else {
          const int wtap = 8 - tap;                     // flipped taps for the backward-data direction
          for (int j = 0; j < p.block_n / 64; ++j)
            tma_load_2d(bres + tap * b_bytes + j * (kBlockK * 128), &map_b, &bars->bres, wtap * p.b_row_len + n00 + j * 64, 0);
        }
      }
      int stage = 0;
      uint32_t phase = 0;
      const int per_img = p.tiles_x * p.tiles_y;
      for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
        const int mt = tile / p.num_n_tiles;
        const int img = mt / per_img, rem = mt - img * per_img;
        const int ty = rem / p.tiles_x, tx = rem - ty * p.tiles_x;
        mbar_wait(&bars->empty[stage], phase ^ 1);
        mbar_expect_tx(&bars->full[stage], (uint32_t)kHaloBytes);
        tma_load_4d(smem + (size_t)stage * stage_bytes, &map_a, &bars->full[stage], 0, tx * 8 - 1, ty * 16 - 1, img);
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    } else if (lane == 0) {
      if (p.b_resident) {
        // gridDim.x is a multiple of num_n_tiles here, so this CTA only ever sees one N tile
        const int n00 = (tile_first % p.num_n_tiles) * p.block_n + (int)cr * bn_cta;
        if (cr == 0) mbar_expect_tx(&bars->bres, (uint32_t)(p.num_k_blocks * b_bytes * kCtas));
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          const int tap = kb / p.cchunks, cc = kb - tap * p.cchunks;
          uint8_t* sb = bres + kb * b_bytes;
          if (!p.b_mn_major) {
            load_2d(sb, &map_b, &bars->bres, p.a_im2col ? tap * p.b_row_len + cc * kBlockK : kb * kBlockK, n00);
          } else {
            const int r = tap / p.S, s2 = tap - r * p.S;
            const int wtap = (p.R - 1 - r) * p.S + (p.S - 1 - s2);
            for (int j = 0; j < bn_cta / 64; ++j)
              load_2d(sb + j * (kBlockK * 128), &map_b, &bars->bres, wtap * p.b_row_len + n00 + j * 64, cc * kBlockK);
          }
        }
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
        // an odd tile count leaves the last pair's second CTA without rows: it re-loads the last real tile (its
        // accumulators are never stored) so the pair's byte count and MMA shape stay uniform
        int mt = (tile / p.num_n_tiles) * kCtas + (int)cr;
        if (mt >= p.num_m_tiles) mt = p.num_m_tiles - 1;
        const int nt = tile % p.num_n_tiles;
        const int m0 = mt * kBlockM, n0 = nt * p.block_n + (int)cr * bn_cta;
        int img = 0, ph = 0, qw = 0;
        if (p.a_im2col) {
          img = m0 / p.PQ;
          int rem = m0 - img * p.PQ;
          int pp = rem / p.Q;
          ph = pp * p.stride_h + p.base_h;
          qw = (rem - pp * p.Q) * p.stride_w + p.base_w;
        }
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          TRACED_WAIT(0, mbar_wait(&bars->empty[stage], phase ^ 1));
          uint8_t* sa = smem + (size_t)stage * stage_bytes;
          uint8_t* sb = sa + kABytes;
          if (cr == 0) mbar_expect_tx(&bars->full[stage], (uint32_t)(stage_bytes * kCtas));
          const int tap = kb / p.cchunks, cc = kb - tap * p.cchunks;
          const int r = tap / p.S, s = tap - r * p.S;
          if (p.a_im2col) {
            if constexpr (kCtas == 2)
              tma_load_im2col_4d_pair(sa, &map_a, &bars->full[stage], cc * kBlockK, qw, ph, img, (uint16_t)s, (uint16_t)r);
            else
              tma_load_im2col_4d(sa, &map_a, &bars->full[stage], cc * kBlockK, qw, ph, img, (uint16_t)s, (uint16_t)r);
          } else {
            load_2d(sa, &map_a, &bars->full[stage], kb * kBlockK, m0);
          }
          if (p.b_resident) {
            // filter already in shared memory
          } else if (!p.b_mn_major) {
            // K-major weights [Cout][(r,s),Cin]: chunk cc of tap (r,s). With Cin % 64 != 0 the last chunk of a tap
            // runs into the next tap's channels; the matching A columns were zero-filled by TMA, so they add nothing.
            load_2d(sb, &map_b, &bars->full[stage], p.a_im2col ? tap * p.b_row_len + cc * kBlockK : kb * kBlockK, n0);
          } else {
            // weight row = one output channel (GEMM K index); columns = ((r,s), cin). Taps are flipped.
            const int wtap = (p.R - 1 - r) * p.S + (p.S - 1 - s);
            for (int j = 0; j < bn_cta / 64; ++j)
              load_2d(sb + j * (kBlockK * 128), &map_b, &bars->full[stage], wtap * p.b_row_len + n0 + j * 64,
                      cc * kBlockK);
          }
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1 && cr == 0) {
    // ================================================================ MMA issuer (the pair's leader only)
    if (kHalo) halo_issue(0, 5, 0);
    const uint32_t idesc = make_idesc_bf16(kBlockM * kCtas, (uint32_t)p.block_n, 0, (uint32_t)p.b_mn_major);
    // descriptors: constant high word, low word stepped with adds (the issue rate of this one thread bounds tiles
    // narrower than 256 columns)
    const uint32_t d_hi = sdesc_hi_sw128(1024);
    auto mma = [d_hi](uint32_t d, uint32_t a_lo, uint32_t b_lo, uint32_t id, uint32_t accum) {
      if constexpr (kCtas == 2) mma_bf16_pair_lohi(d, a_lo, d_hi, b_lo, d_hi, id, accum);
      else mma_bf16_lohi(d, a_lo, d_hi, b_lo, d_hi, id, accum);
    };
    const uint32_t b_lbo = p.b_mn_major ? (uint32_t)(kBlockK * 128) : 16u;   // MN-major: distance between 64-column chunks
    const uint32_t b_kstep = p.b_mn_major ? (2048u >> 4) : (32u >> 4);        // descriptor step per 16 K elements
    auto commit = [](uint64_t* bar) {
      if constexpr (kCtas == 2) mma_commit_pair(bar);
      else mma_commit(bar);
    };
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    if (p.b_resident) mbar_wait(&bars->bres, 0);
    for (int tile = tile_first; tile < (kHalo ? 0 : num_tiles); tile += tile_step) {
      TRACED_WAIT(1, mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1));
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * kMaxBlockN);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        TRACED_WAIT(2, mbar_wait(&bars->full[stage], phase));
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + (size_t)stage * stage_bytes);
          const uint32_t sb = p.b_resident ? smem_u32(bres) + (uint32_t)(kb * b_bytes) : sa + kABytes;
          const uint32_t a_lo = sdesc_lo(sa, 16), b_lo = sdesc_lo(sb, b_lbo);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            mma(d_tmem, a_lo + k * 2, b_lo + k * b_kstep, idesc, (kb | k) != 0 ? 1u : 0u);
          commit(&bars->empty[stage]);                           // frees the smem slot (in both CTAs of a pair)
          if (kb == p.num_k_blocks - 1) commit(&bars->tmem_full[acc]);
        }
        __syncwarp();
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (++acc == kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  } else if (warp >= 2) {
    // ================================================================ epilogue (warps 2..17 = groups 0..3)
    // handing a drained accumulator stage back: the MMA issuer waits on the leader's barrier
    auto release_acc = [&](uint64_t* bar) {
      if (kCtas == 2 && cr != 0) mbar_arrive_cluster(bar, 0);
      else mbar_arrive(bar);
    };
    const int lane_group = warp & 3;            // TMEM lanes [32*lane_group, +32) are visible to this warp
    const int grp = (warp - 2) >> 2;            // 0..3
    const int acc = grp >> 1;                   // pair a = groups 2a, 2a+1 owns accumulator stage a: tiles a, a+2, ...
    const int half = grp & 1;                   // which half of the tile's boxes this group drains
    uint32_t acc_phase = 0;
    if (p.col_stats) {
      // running statistics of this (CTA, pair): four 32-row-group rows of [2][N] floats, zeroed here (no memset launch)
      const int eg = threadIdx.x - 64 - acc * 256;    // 0..255 within the pair
      float4* z = reinterpret_cast<float4*>(p.col_stats + ((size_t)blockIdx.x * 2 + acc) * 4 * 2 * p.N);
      const int n4 = 4 * 2 * p.N / 4;
      for (int i = eg; i < n4; i += 256) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      __threadfence_block();
      asm volatile("bar.sync %0, 256;" ::"r"(6 + acc) : "memory");
    }
    if (kHalo && warp == 6) halo_issue(5, 9, 64);      // second MMA-issuing warp of halo mode (its group has no box)
    if (p.tma_store) {
      // ---- staged path: TMEM -> registers -> bf16 -> swizzled shared memory -> TMA bulk store (or cooperative
      //      scatter); BatchNorm statistics are taken from the staged (stored) values with conflict-free reads.
      const int et = threadIdx.x - 64 - grp * 128;   // 0..127 within the group
      const int nboxes = p.block_n / 64;
      // boxes [bx0, bx1) of every tile of this pair are this group's; a one-box tile leaves the second group idle
      const int bx0 = half ? (nboxes + 1) / 2 : 0, bx1 = half ? nboxes : (nboxes + 1) / 2;
      const int r_in_tile = lane_group * 32 + lane;
      const int cp = et & 31;                   // column pair handled for statistics
      const int qtr = et >> 5;                  // 32-row group handled for statistics
      uint8_t* const buf = out_stage + grp * kStageOutBytes;   // one staging box per group
      const int bar_id = 1 + grp;
      bool store_pending = false;
      int tile_i = 0;
      for (int tile = tile_first; tile < num_tiles && bx0 < bx1; tile += tile_step, ++tile_i) {
        if ((tile_i & 1) != acc) continue;
        const int mt = (tile / p.num_n_tiles) * kCtas + (int)cr, nt = tile % p.num_n_tiles;
        const int n0 = nt * p.block_n;
        if (kCtas == 2 && mt >= p.num_m_tiles) {
          // rowless second half of the last pair: nothing to store, just hand the stage back
          mbar_wait_backoff(&bars->tmem_full[acc], acc_phase);
          tc_fence_after();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(&bars->tmem_empty[acc]);
          acc_phase ^= 1;
          continue;
        }
        int h_img = 0, h_y0 = 0, h_x0 = 0;           // halo mode: image and top-left output pixel of this tile
        if (kHalo) {
          const int per_img = p.tiles_x * p.tiles_y;
          h_img = mt / per_img;
          const int rem = mt - h_img * per_img;
          h_y0 = (rem / p.tiles_x) * 16;
          h_x0 = (rem % p.tiles_x) * 8;
        }
        TRACED_WAIT(3, mbar_wait_backoff(&bars->tmem_full[acc], acc_phase));
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)(acc * kMaxBlockN);
        for (int bx = bx0; bx < bx1; ++bx) {
          // this CTA-group's running column statistics for (row group qtr, column pair cp) of this box: fetched now,
          // the latency hides behind the TMEM loads below; only this thread ever touches these addresses
          float* const st_dst = (p.col_stats && n0 + bx * 64 + cp * 2 < p.N)
                                    ? p.col_stats + (((size_t)blockIdx.x * 2 + acc) * 4 + qtr) * 2 * p.N + n0 + bx * 64 + cp * 2
                                    : nullptr;
          float2 old1 = make_float2(0.f, 0.f), old2 = make_float2(0.f, 0.f);
          if (st_dst) {
            old1 = __ldcg(reinterpret_cast<const float2*>(st_dst));
            old2 = __ldcg(reinterpret_cast<const float2*>(st_dst + p.N));
          }
          // TMEM -> registers -> packed bf16 first; only then wait for the staging buffer, so the previous bulk
          // store's read of the buffer overlaps these loads
          uint4 pk[8];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t r[32];
            if (kHalo) {
              // two partial accumulators (one per issuing warp): columns [0, 64) and [64, 128) of the stage
#pragma unroll
              for (int q = 0; q < 2; ++q) {
                uint32_t ra[16], rb[16];
                tmem_ld_x16(t_row + (uint32_t)(half * 32 + q * 16), ra);
                tmem_ld_x16(t_row + (uint32_t)(64 + half * 32 + q * 16), rb);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; ++j) r[q * 16 + j] = __float_as_uint(__uint_as_float(ra[j]) + __uint_as_float(rb[j]));
              }
            } else {
              tmem_ld_x32(t_row + (uint32_t)(bx * 64 + half * 32), r);
              tmem_ld_wait();
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {       // four 16-byte chunks of 8 bf16
              __nv_bfloat162 t;
#define PK2(a, b) (t = __floats2bfloat162_rn(__uint_as_float(a), __uint_as_float(b)), *reinterpret_cast<uint32_t*>(&t))
              pk[half * 4 + j].x = PK2(r[8 * j + 0], r[8 * j + 1]);
              pk[half * 4 + j].y = PK2(r[8 * j + 2], r[8 * j + 3]);
              pk[half * 4 + j].z = PK2(r[8 * j + 4], r[8 * j + 5]);
              pk[half * 4 + j].w = PK2(r[8 * j + 6], r[8 * j + 7]);
#undef PK2
            }
          }
          // the bulk store that last read this buffer must have finished reading it
          if (et == 0 && store_pending) tma_store_wait_read<0>();
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
#pragma unroll
          for (int chunk = 0; chunk < 8; ++chunk)
            *reinterpret_cast<uint4*>(buf + r_in_tile * 128 + ((chunk ^ (r_in_tile & 7)) << 4)) = pk[chunk];
          if (bx == bx1 - 1) {
            // all of this tile's accumulators are out of TMEM: hand the stage back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) release_acc(&bars->tmem_empty[acc]);
          }
          fence_proxy_async();
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
          if (!p.scatter) {
            if (et == 0) {
              if (kHalo)
                tma_store_4d(&map_out, buf, n0 + bx * 64, h_x0, h_y0, h_img);   // 16 x 8 pixels; clipped at the edges
              else
                tma_store_2d(&map_out, buf, n0 + bx * 64, mt * kBlockM);
              tma_store_commit();
              store_pending = true;
            }
          } else {
            // strided 1x1 dgrad: GEMM row (n,p,q) -> pixel (n, p*sy, q*sx); the skipped pixels get zeros. Eight
            // threads write one 128-byte row segment, so every store instruction covers whole cache lines.
            const int chunk = et & 7;
            const int col = n0 + bx * 64 + chunk * 8;
            if (col < p.N) {
#pragma unroll 2
              for (int pass = 0; pass < 8; ++pass) {
                const int rr = pass * 16 + (et >> 3);
                const int row = mt * kBlockM + rr;
                if (row >= p.M) break;
                const int img = row / p.PQ;
                const int rem = row - img * p.PQ;
                const int pp = rem / p.Q, qq = rem - pp * p.Q;
                const size_t opix = ((size_t)img * p.H + (size_t)pp * p.sy + p.oy) * p.W + (size_t)qq * p.sx + p.ox;
                const uint4 v = *reinterpret_cast<const uint4*>(buf + rr * 128 + ((chunk ^ (rr & 7)) << 4));
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + opix * p.ldo + col;
                const uint4 z = make_uint4(0, 0, 0, 0);
                if (!p.scatter_fill) {
                  *reinterpret_cast<uint4*>(o) = v;
                } else {
                  for (int dy = 0; dy < p.sy; ++dy)
                    for (int dx = 0; dx < p.sx; ++dx)
                      *reinterpret_cast<uint4*>(o + ((size_t)dy * p.W + dx) * p.ldo) = (dy | dx) ? z : v;
                }
              }
            }
          }
          if (st_dst) {
            float s1a = 0.f, s1b = 0.f, s2a = 0.f, s2b = 0.f;
            const int chunk = cp >> 2, word = cp & 3;
#pragma unroll 8
            for (int i = 0; i < 32; ++i) {
              const int row = qtr * 32 + i;
              // halo tiles may hang over the image edge: those rows hold values that are never stored
              if (kHalo && (h_y0 + (row >> 3) >= p.Himg || h_x0 + (row & 7) >= p.Wimg)) continue;
              const uint32_t w = *reinterpret_cast<const uint32_t*>(buf + row * 128 + ((chunk ^ (row & 7)) << 4) + word * 4);
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w));
              s1a += f.x;
              s1b += f.y;
              s2a = fmaf(f.x, f.x, s2a);
              s2b = fmaf(f.y, f.y, s2b);
            }
            __stcg(reinterpret_cast<float2*>(st_dst), make_float2(old1.x + s1a, old1.y + s1b));
            __stcg(reinterpret_cast<float2*>(st_dst + p.N), make_float2(old2.x + s2a, old2.y + s2b));
          }
        }
        acc_phase ^= 1;
      }
      if (et == 0 && store_pending) tma_store_wait<0>();
    } else {
      int tile_i = 0;
      // narrow / fp32 outputs: the first group of each pair drains the whole tile
      for (int tile = tile_first; tile < num_tiles && half == 0; tile += tile_step, ++tile_i) {
        if ((tile_i & 1) != acc) continue;
        const int mt = (tile / p.num_n_tiles) * kCtas + (int)cr, nt = tile % p.num_n_tiles;
        const int row = mt * kBlockM + lane_group * 32 + lane;
        const int n0 = nt * p.block_n;
        mbar_wait_backoff(&bars->tmem_full[acc], acc_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)(acc * kMaxBlockN);
        // the rowless second half of the last pair holds a copy of the last real tile: no stores, no statistics
        const bool ghost = kCtas == 2 && mt >= p.num_m_tiles;
        const bool valid = row < p.M && !ghost;
        size_t orow = (size_t)row;
        if (p.scatter && valid) {
          const int img = row / p.PQ;
          const int rem = row - img * p.PQ;
          const int pp = rem / p.Q, qq = rem - pp * p.Q;
          orow = ((size_t)img * p.H + (size_t)pp * p.sy + p.oy) * p.W + (size_t)qq * p.sx + p.ox;
        }
        for (int c = 0; c < p.block_n; c += 16) {
          uint32_t r[16];
          tmem_ld_x16(t_row + (uint32_t)c, r);
          tmem_ld_wait();
          if (p.col_stats && !ghost) {
            // per-column sum / sum of squares over this warp's 32 rows (rows past M contribute zeros: their A rows
            // were zero-filled by TMA), via a halving butterfly: 16+8+4+2+1 shuffles per statistic
            float s1[16], s2[16];
  #pragma unroll
            for (int i = 0; i < 16; ++i) {
              // statistics are taken of the bf16-rounded value, i.e. of what is stored
              float v = __bfloat162float(__float2bfloat16(__uint_as_float(r[i])));
              s1[i] = v;
              s2[i] = v * v;
            }
  #pragma unroll
            for (int i = 0; i < 16; ++i) {
              s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 16);
              s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 16);
            }
  #pragma unroll
            for (int w = 8; w >= 1; w >>= 1) {
  #pragma unroll
              for (int i = 0; i < w; ++i) {
                const bool up = (lane & w) != 0;
                float a1 = up ? s1[i] : s1[i + w], b1 = up ? s1[i + w] : s1[i];
                float a2 = up ? s2[i] : s2[i + w], b2 = up ? s2[i + w] : s2[i];
                s1[i] = b1 + __shfl_xor_sync(0xffffffffu, a1, w);
                s2[i] = b2 + __shfl_xor_sync(0xffffffffu, a2, w);
              }
            }
            // lane l now holds column (l & 15) of this chunk (both 16-lane halves hold the same totals)
            if (lane < 16 && n0 + c + lane < p.N) {
              const int col = lane;
              float* dst = p.col_stats + (((size_t)blockIdx.x * 2 + acc) * 4 + lane_group) * 2 * p.N + n0 + c + col;
              dst[0] += s1[0];
              dst[p.N] += s2[0];
            }
          }
          if (valid && n0 + c < p.N) {
            if (p.out_fp32) {
              float* o = reinterpret_cast<float*>(p.out) + orow * p.ldo + n0 + c;
              if (n0 + c + 16 <= p.N && (p.ldo & 3) == 0) {
  #pragma unroll
                for (int i = 0; i < 4; ++i)
                  reinterpret_cast<uint4*>(o)[i] = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
              } else {
                for (int i = 0; i < 16 && n0 + c + i < p.N; ++i) o[i] = __uint_as_float(r[i]);
              }
            } else {
              __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + orow * p.ldo + n0 + c;
              if (n0 + c + 16 <= p.N && (p.ldo & 7) == 0) {
                store_row_bf16(o, r);
                if (p.scatter && p.scatter_fill) {
                  // the pixels the strided forward conv never read receive a zero gradient
                  const uint4 z = make_uint4(0, 0, 0, 0);
                  for (int dy = 0; dy < p.sy; ++dy)
                    for (int dx = 0; dx < p.sx; ++dx) {
                      if (dy == 0 && dx == 0) continue;
                      __nv_bfloat16* oz = o + ((size_t)dy * p.W + dx) * p.ldo;
                      reinterpret_cast<uint4*>(oz)[0] = z;
                      reinterpret_cast<uint4*>(oz)[1] = z;
                    }
                }
              } else {
                for (int i = 0; i < 16 && n0 + c + i < p.N; ++i) o[i] = __float2bfloat16(__uint_as_float(r[i]));
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(&bars->tmem_empty[acc]);
        acc_phase ^= 1;
      }
    }
    if (p.col_stats) {
      // fold this CTA's 8 running rows (2 pairs x 4 row groups) into one row of the compact result region that
      // follows the scratch rows, in a fixed order; bn_finalize_partials then reads one row per CTA
      __threadfence_block();
      asm volatile("bar.sync 5, 512;" ::: "memory");
      const int e8 = threadIdx.x - 64;          // 0..511
      const float* src = p.col_stats + (size_t)blockIdx.x * 8 * 2 * p.N;
      float* dst = p.col_stats + (size_t)gridDim.x * 8 * 2 * p.N + (size_t)blockIdx.x * 2 * p.N;
      for (int i = e8; i < 2 * p.N; i += 512) {
        float a = 0.0f;
#pragma unroll
        for (int rrow = 0; rrow < 8; ++rrow) a += __ldcg(src + (size_t)rrow * 2 * p.N + i);
        dst[i] = a;
      }
    }
  }

  tc_fence_before();
  // a pair leaves together: the peer may still read this CTA's operands or signal its barriers
  if constexpr (kCtas == 2) cluster_sync_all();
  else __syncthreads();
  if (tr && lane == 0) {
    // slots 0..2 have one writer each; slot 3 = epilogue warp 2 (group 0, lane group 2); slot 4 = CTA lifetime
    if (warp == 0) p.trace[0] = tr_acc[0];
    if (warp == 1) {
      p.trace[1] = tr_acc[1];
      p.trace[2] = tr_acc[2];
    }
    if (warp == 2) p.trace[3] = tr_acc[3];
    if (warp == 3) p.trace[4] = clock64() - tr_start;
  }
  if (warp == 1) {
    tc_fence_after();
    if constexpr (kCtas == 2) tmem_dealloc_pair<kTmemCols>(tmem_base);
    else tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------- host side

int pick_block_n(int n) {
  // N >= 64: a multiple of 64 (staged TMA-store epilogue); columns past N are zero-filled on load, clipped on store
  // and skipped in the statistics. Prefer the widest tile that wastes at most a quarter of its columns.
  if (n >= 64) {
    const int cand[4] = {256, 192, 128, 64};
    for (int i = 0; i < 4; ++i) {
      const int bn = cand[i];
      const int tiles = (n + bn - 1) / bn;
      if ((tiles * bn - n) * 4 <= tiles * bn) return bn;
    }
    return 64;
  }
  return (n + 15) / 16 * 16;   // narrow outputs: one tile, direct-store epilogue masks the tail
}

// CTA pairs need a tile width that splits into two 64-column-aligned halves and more than one M tile. They pay off
// where the main loop is long (measured on ResNet-50 layers: >= 8 k-blocks; 5-13% for 256-wide tiles and for the
// MN-major weight operand of dgrad); short-K memory-bound layers lose a little to the cross-CTA hand-offs per tile.
int plan_ctas(int block_n, int num_m_tiles, int num_k_blocks, bool b_mn_major) {
  static const int mode = getenv("B200_CONV_2CTA") ? atoi(getenv("B200_CONV_2CTA")) : 1;   // 0 off, 1 auto, 2 always
  if (mode == 0 || block_n % 128 != 0 || num_m_tiles < 2 || num_k_blocks <= 0) return 1;
  if (mode == 2) return 2;
  return (num_k_blocks >= 8 && (block_n == 256 || b_mn_major)) ? 2 : 1;
}

// persistent grid: one CTA (or CTA pair) per SM (pair), never more than there are tiles
int plan_grid(int num_m_tiles, int num_n_tiles, int ctas) {
  const int groups = (num_m_tiles + ctas - 1) / ctas * num_n_tiles;
  const int max_groups = kNumSMs / ctas;
  return (groups < max_groups ? groups : max_groups) * ctas;
}

template <int kCtas, bool kHalo>
cudaError_t launch_kernel(int grid, size_t smem, cudaStream_t st, const CUtensorMap& ma, const CUtensorMap& mb,
                          const CUtensorMap& mo, const GemmParams& p) {
  static std::atomic<uint64_t> attr_set{0};
  if (first_use_on_device(attr_set)) {
    cudaError_t e = cudaFuncSetAttribute(conv_gemm_kernel<kCtas, kHalo>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxConvSmem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = kCtas > 1 ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, conv_gemm_kernel<kCtas, kHalo>, ma, mb, mo, p);
}

int launch(const CUtensorMap& ma, const CUtensorMap& mb, GemmParams& p, cudaStream_t st) {
  if (p.ctas < 1) p.ctas = 1;
  if (p.halo) p.ctas = 1;
  const int bn_cta = p.block_n / p.ctas;
  const int stage_bytes = kABytes + bn_cta * kBlockK * 2;
  // bf16 row-major outputs whose tile width is a multiple of 64 go out through shared memory + TMA bulk stores
  CUtensorMap mo;
  p.tma_store = (!p.out_fp32 && p.block_n % 64 == 0 && p.N % 8 == 0 && p.ldo % 8 == 0) ? 1 : 0;
  if (p.halo) {
    p.tma_store = 1;
    if (tmap_nhwc_tile(&mo, p.out, p.M / (p.Himg * p.Wimg), p.Himg, p.Wimg, p.N, 64, 8, 16)) return -1;
  } else if (p.tma_store && !p.scatter) {
    if (tmap_tiled_2d(&mo, p.out, 2, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ldo * 2, 64, kBlockM, true)) return -1;
  } else {
    mo = ma;   // unused
  }
  int fixed = (int)sizeof(Barriers) + 1024 + (p.tma_store ? kOutBufs * kStageOutBytes : 0);
  int stage_bytes_eff = stage_bytes;
  const int tiles_all = p.num_m_tiles * p.num_n_tiles;
  if (p.halo) {
    fixed += 9 * p.block_n * kBlockK * 2;      // resident filter
    stage_bytes_eff = kHaloBytes;
  } else {
    // keep the filter slice resident when it is small and every CTA sticks to one N tile (memory-bound 1x1 layers:
    // the per-tile filter reload would otherwise be most of the L2 traffic)
    static const bool res_ok = !(getenv("B200_CONV_BRES") && getenv("B200_CONV_BRES")[0] == '0');
    const int64_t bres_bytes = (int64_t)p.num_k_blocks * bn_cta * kBlockK * 2;
    const int grid0 = plan_grid(p.num_m_tiles, p.num_n_tiles, p.ctas) / p.ctas;
    static const int res_max_kb = getenv("B200_CONV_BRES_KB") ? atoi(getenv("B200_CONV_BRES_KB")) : 96;
    // at least two 16 KiB activation stages must still fit next to the resident slice
    const bool fits = fixed + bres_bytes + 2 * kABytes <= kMaxConvSmem;
    p.b_resident = (res_ok && fits && bres_bytes <= (int64_t)res_max_kb * 1024 && p.num_m_tiles >= 2 * kNumSMs &&
                    grid0 % p.num_n_tiles == 0) ? 1 : 0;
    if (p.b_resident) {
      fixed += (int)bres_bytes;
      stage_bytes_eff = kABytes;
    }
  }
  int stages = (kMaxConvSmem - fixed) / stage_bytes_eff;
  if (stages > 8) stages = 8;
  if (stages < 2) stages = 2;
  p.stages = stages;
  const size_t smem = (size_t)stages * stage_bytes_eff + fixed;
  int grid = plan_grid(p.num_m_tiles, p.num_n_tiles, p.ctas);
  if (p.halo) grid = grid / p.num_n_tiles * p.num_n_tiles;   // a CTA keeps one N tile's filter resident
  static const bool trace_on = getenv("B200_CONV_TRACE") && getenv("B200_CONV_TRACE")[0] == '1';
  static long long* trace_buf = nullptr;
  if (trace_on) {
    // debug only: synchronous; never on in a benchmark or under graph capture
    if (!trace_buf) B200_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&trace_buf), g_trace));
    B200_CUDA(cudaMemsetAsync(trace_buf, 0, 8 * sizeof(long long), st));
    p.trace = trace_buf;
  }
  if (p.halo) B200_CUDA((launch_kernel<1, true>(grid, smem, st, ma, mb, mo, p)));
  else if (p.ctas == 2) B200_CUDA((launch_kernel<2, false>(grid, smem, st, ma, mb, mo, p)));
  else B200_CUDA((launch_kernel<1, false>(grid, smem, st, ma, mb, mo, p)));
  B200_LAUNCH_CHECK();
  if (trace_on) {
    long long h[8];
    B200_CUDA(cudaStreamSynchronize(st));
    B200_CUDA(cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost));
    fprintf(stderr,
            "[conv trace] M=%d N=%d kblocks=%d block_n=%d stages=%d bres=%d halo=%d mnB=%d grid=%d ctas=%d | CTA0 cycles %lld: "
            "producer<-empty %lld, mma<-tmem_empty %lld, mma<-full %lld, epilogue<-tmem_full %lld\n",
            p.M, p.N, p.num_k_blocks, p.block_n, p.stages, p.b_resident, p.halo, p.b_mn_major, grid, p.ctas, h[4], h[0], h[1],
            h[2], h[3]);
  }
  return 0;
}

}  // namespace
}  // namespace b200

using namespace b200;

static bool halo_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_HALO");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

// 3x3 / stride 1 / pad 1 with exactly 64 reduction channels and 64 output columns: the halo kernel applies (with wider
// outputs the generic im2col path, which covers all columns in one pass over the input, measured faster)
static bool halo_shape(const b200_conv_geom* g, int red_ch, int out_ch) {
  return halo_enabled() && g->r == 3 && g->s == 3 && g->stride_h == 1 && g->stride_w == 1 && g->pad_t == 1 &&
         g->pad_l == 1 && g->pad_b == 1 && g->pad_r == 1 && red_ch == 64 && out_ch == 64 &&
         g->h >= 16 && g->w >= 8;
}

static int64_t halo_grid(const b200_conv_geom* g, int out_ch) {
  const int nt = out_ch / 64;
  int64_t tiles = (int64_t)g->n * ((g->h + 15) / 16) * ((g->w + 7) / 8) * nt;
  int64_t grid = tiles < kNumSMs ? tiles : kNumSMs;
  return grid / nt * nt;
}


extern "C" int b200_conv_wgrad_supported(const b200_conv_geom* g);

// smallest INPUT channel count the TMA paths take. A 64-channel box over fewer channels is zero-filled past the edge,
// so thin layers (Inception-v3 stem 32x149x149, GoogLeNet 5x5 branches) compute correctly on the same kernels with half
// of every k-block idle. The library measured faster on exactly those layers (Inception-v3 at batch 128: 22.4 ms/step
// with them on cuDNN, 24.3 ms with everything here), but the product definition is "no cuDNN on the hot ops", so the
// default takes everything with Cin % 8 == 0; B200_CONV_MIN_C=64 restores the library hand-off for thin inputs.
extern "C" int b200_conv_min_channels() {
  static const int v = getenv("B200_CONV_MIN_C") ? atoi(getenv("B200_CONV_MIN_C")) : 8;
  return v;
}
static int min_channels() { return b200_conv_min_channels(); }

extern "C" {

int b200_conv_supported(const b200_conv_geom* g, int op) {
  if (!g) return 0;
  if (g->n <= 0 || g->p <= 0 || g->q <= 0) return 0;
  if (op == B200_CONV_WGRAD) return b200_conv_wgrad_supported(g);
  if (op == B200_CONV_FPROP) {
    // channel chunks past Cin are zero-filled by TMA, so Cin only has to keep rows 16-byte aligned
    return g->c >= min_channels() && g->c % 8 == 0 && g->k % 8 == 0 && g->r <= 15 && g->s <= 15;
  }
  if (op == B200_CONV_DGRAD) {
    // the reduction runs over Cout here, so a thin Cout only leaves part of a 64-channel k-block idle; a thin Cin
    // would be a thin output tile, which the library handles better
    if (g->c < min_channels() || g->k % 8 != 0 || g->c % 8 != 0) return 0;
    if (g->stride_h == 1 && g->stride_w == 1) return g->r <= 15 && g->s <= 15;
    // strided 1x1 without padding (ResNet v1 down-sampling convolutions): one launch with a scattering epilogue
    if (g->r == 1 && g->s == 1 && g->pad_t == 0 && g->pad_l == 0 && g->pad_b == 0 && g->pad_r == 0 &&
        g->p * g->stride_h == g->h && g->q * g->stride_w == g->w)
      return 1;
    // any other strided filter: one stride-1 launch per stride phase (b200_conv_dgrad_strided) -> answer 2
    return (g->stride_h <= 4 && g->stride_w <= 4 && g->r <= 15 && g->s <= 15) ? 2 : 0;
  }
  return 0;
}

int b200_conv_fprop(const b200_bf16* x, const b200_bf16* w, b200_bf16* y, const b200_conv_geom* g, float* col_stats,
                    void* stream) {
  B200_CHECK_ARG(x && w && y && g, "conv_fprop: null pointer");
  B200_CHECK_ARG(b200_conv_supported(g, B200_CONV_FPROP), "conv_fprop: unsupported geometry (need Cin>=64, Cin%%8==0, Cout%%8==0)");
  CUtensorMap ma, mb;
  if (halo_shape(g, g->c, g->k)) {
    GemmParams p{};
    p.halo = 1;
    p.Himg = g->h;
    p.Wimg = g->w;
    p.tiles_y = (g->h + 15) / 16;
    p.tiles_x = (g->w + 7) / 8;
    p.M = g->n * g->h * g->w;
    p.N = g->k;
    p.block_n = 64;
    p.num_m_tiles = g->n * p.tiles_y * p.tiles_x;
    p.num_n_tiles = g->k / 64;
    p.num_k_blocks = 1;
    p.cchunks = 1;
    p.R = p.S = 3;
    p.b_row_len = g->c;
    p.out = y;
    p.ldo = g->k;
    p.col_stats = col_stats;
    if (tmap_nhwc_tile(&ma, x, g->n, g->h, g->w, g->c, 64, kHaloCols, kHaloRows)) return -1;
    const int kd = 9 * g->c;
    if (tmap_tiled_2d(&mb, w, 2, (uint64_t)kd, (uint64_t)g->k, (uint64_t)kd * 2, kBlockK, 64, true)) return -1;
    return launch(ma, mb, p, static_cast<cudaStream_t>(stream));
  }
  const int upper_w = g->pad_r - (g->s - 1), upper_h = g->pad_b - (g->r - 1);
  if (tmap_im2col_nhwc(&ma, x, g->n, g->h, g->w, g->c, -g->pad_l, -g->pad_t, upper_w, upper_h, g->stride_w, g->stride_h,
                       kBlockK, kBlockM))
    return -1;
  GemmParams p{};
  p.M = g->n * g->p * g->q;
  p.N = g->k;
  p.block_n = pick_block_n(g->k);
  p.num_m_tiles = (p.M + kBlockM - 1) / kBlockM;
  p.ctas = plan_ctas(p.block_n, p.num_m_tiles, g->r * g->s * ((g->c + kBlockK - 1) / kBlockK), false);
  const int kdim = g->r * g->s * g->c;
  if (tmap_tiled_2d(&mb, w, 2, (uint64_t)kdim, (uint64_t)g->k, (uint64_t)kdim * 2, kBlockK, p.block_n / p.ctas, true))
    return -1;
  p.num_n_tiles = (p.N + p.block_n - 1) / p.block_n;
  p.cchunks = (g->c + kBlockK - 1) / kBlockK;
  p.num_k_blocks = g->r * g->s * p.cchunks;
  p.a_im2col = 1;
  p.b_row_len = g->c;
  p.PQ = g->p * g->q;
  p.Q = g->q;
  p.stride_h = g->stride_h;
  p.stride_w = g->stride_w;
  p.base_h = -g->pad_t;
  p.base_w = -g->pad_l;
  p.R = g->r;
  p.S = g->s;
  p.out = y;
  p.ldo = g->k;
  p.col_stats = col_stats;
  return launch(ma, mb, p, static_cast<cudaStream_t>(stream));
}

static int64_t grid_for_tiles(int64_t m, int n, int num_k_blocks) {
  const int bn = pick_block_n(n);
  const int mt = (int)((m + kBlockM - 1) / kBlockM), nt = (n + bn - 1) / bn;
  return plan_grid(mt, nt, plan_ctas(bn, mt, num_k_blocks, false));
}

static int conv_k_blocks(const b200_conv_geom* g) { return g->r * g->s * ((g->c + kBlockK - 1) / kBlockK); }

int64_t b200_conv_stats_rows(const b200_conv_geom* g) {
  // per CTA of the persistent grid: 8 scratch rows (epilogue group x 32-row group) + 1 folded row at the end
  if (!g) return -1;
  if (halo_shape(g, g->c, g->k)) return halo_grid(g, g->k) * 9;
  return grid_for_tiles((int64_t)g->n * g->p * g->q, g->k, conv_k_blocks(g)) * 9;
}

int64_t b200_conv_stats_first_result_row(const b200_conv_geom* g) {
  if (!g) return -1;
  if (halo_shape(g, g->c, g->k)) return halo_grid(g, g->k) * 8;
  return grid_for_tiles((int64_t)g->n * g->p * g->q, g->k, conv_k_blocks(g)) * 8;
}

}  // extern "C"

// One stride phase of a general strided dgrad, expressed as a stride-1 dgrad over a sub-filter whose output rows are
// scattered into the real dx image (see b200_conv_dgrad_strided).
struct DgradPhase {
  int sy, sx, oy, ox, H, W;
};

static int dgrad_impl(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, const DgradPhase* ph,
                      void* stream) {
  CUtensorMap ma, mb;
  if (!ph && halo_shape(g, g->k, g->c)) {
    GemmParams p{};
    p.halo = 1;
    p.Himg = g->h;
    p.Wimg = g->w;
    p.tiles_y = (g->h + 15) / 16;
    p.tiles_x = (g->w + 7) / 8;
    p.M = g->n * g->h * g->w;
    p.N = g->c;
    p.block_n = 64;
    p.num_m_tiles = g->n * p.tiles_y * p.tiles_x;
    p.num_n_tiles = g->c / 64;
    p.num_k_blocks = 1;
    p.cchunks = 1;
    p.R = p.S = 3;
    p.b_mn_major = 1;
    p.b_row_len = g->c;
    p.out = dx;
    p.ldo = g->c;
    if (tmap_nhwc_tile(&ma, dy, g->n, g->p, g->q, g->k, 64, kHaloCols, kHaloRows)) return -1;
    const int row = 9 * g->c;
    if (tmap_tiled_2d(&mb, w, 2, (uint64_t)row, (uint64_t)g->k, (uint64_t)row * 2, 64, kBlockK, true)) return -1;
    return launch(ma, mb, p, static_cast<cudaStream_t>(stream));
  }
  GemmParams p{};
  const bool strided = g->stride_h != 1 || g->stride_w != 1;
  if (!strided) {
    // dx[n,h,w,:] = sum_{r,s} dy[n, h + pad_t - r, w + pad_l - s, :] . w[:, r, s, :]
    const int lower_h = g->pad_t - (g->r - 1), lower_w = g->pad_l - (g->s - 1);
    const int upper_h = lower_h + g->h - g->p, upper_w = lower_w + g->w - g->q;
    if (tmap_im2col_nhwc(&ma, dy, g->n, g->p, g->q, g->k, lower_w, lower_h, upper_w, upper_h, 1, 1, kBlockK, kBlockM))
      return -1;
    p.M = g->n * g->h * g->w;
    p.PQ = g->h * g->w;
    p.Q = g->w;
    p.base_h = lower_h;
    p.base_w = lower_w;
  } else {
    if (tmap_im2col_nhwc(&ma, dy, g->n, g->p, g->q, g->k, 0, 0, 0, 0, 1, 1, kBlockK, kBlockM)) return -1;
    p.M = g->n * g->p * g->q;
    p.PQ = g->p * g->q;
    p.Q = g->q;
    p.scatter = 1;
    p.scatter_fill = 1;
    p.sy = g->stride_h;
    p.sx = g->stride_w;
    p.H = g->h;
    p.W = g->w;
  }
  if (ph) {
    p.scatter = 1;
    p.scatter_fill = 0;
    p.sy = ph->sy;
    p.sx = ph->sx;
    p.oy = ph->oy;
    p.ox = ph->ox;
    p.H = ph->H;
    p.W = ph->W;
  }
  p.stride_h = p.stride_w = 1;
  p.N = g->c;
  // tile width: a multiple of 64 (the MN-major weight boxes are 64 columns wide); columns past Cin are zero-filled on
  // load and clipped on store
  p.block_n = g->c % 256 == 0 ? 256 : (g->c % 128 == 0 ? 128 : (g->c % 64 == 0 ? 64 : (g->c > 192 ? 256 : (g->c > 128 ? 192 : (g->c > 64 ? 128 : 64)))));
  const int row = g->r * g->s * g->c;
  if (tmap_tiled_2d(&mb, w, 2, (uint64_t)row, (uint64_t)g->k, (uint64_t)row * 2, 64, kBlockK, true)) return -1;
  p.num_m_tiles = (p.M + kBlockM - 1) / kBlockM;
  p.ctas = plan_ctas(p.block_n, p.num_m_tiles, g->r * g->s * ((g->k + kBlockK - 1) / kBlockK), true);
  p.num_n_tiles = (p.N + p.block_n - 1) / p.block_n;
  p.cchunks = (g->k + kBlockK - 1) / kBlockK;
  p.num_k_blocks = g->r * g->s * p.cchunks;
  p.a_im2col = 1;
  p.R = g->r;
  p.S = g->s;
  p.b_mn_major = 1;
  p.b_row_len = g->c;
  p.out = dx;
  p.ldo = g->c;
  return launch(ma, mb, p, static_cast<cudaStream_t>(stream));
}

namespace b200 {
namespace {
// w_ab[k][u][v][c] = w[k][a + sy*u][b + sx*v][c]: the taps of one stride phase, packed as a KRSC filter of its own
__global__ void __launch_bounds__(256) phase_filter_kernel(const uint4* __restrict__ w, uint4* __restrict__ out, int K, int R,
                                                           int S, int cvec, int a, int b, int sy, int sx, int U, int V) {
  const int64_t total = (int64_t)K * U * V * cvec;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % cvec);
    int64_t t = i / cvec;
    const int v = (int)(t % V);
    t /= V;
    const int u = (int)(t % U);
    const int k = (int)(t / U);
    out[i] = w[(((int64_t)k * R + a + sy * u) * S + b + sx * v) * cvec + cv];
  }
}
}  // namespace
}  // namespace b200

static inline int ceil_div_i(int a, int b) { return a >= 0 ? (a + b - 1) / b : -((-a) / b); }
static inline int floor_div_i(int a, int b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }

extern "C" {

int b200_conv_dgrad(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* stream) {
  B200_CHECK_ARG(dy && w && dx && g, "conv_dgrad: null pointer");
  B200_CHECK_ARG(b200_conv_supported(g, B200_CONV_DGRAD) == 1,
                 "conv_dgrad: unsupported geometry (general strided filters go through b200_conv_dgrad_strided)");
  return dgrad_impl(dy, w, dx, g, nullptr, stream);
}

int64_t b200_conv_dgrad_strided_workspace_bytes(const b200_conv_geom* g) {
  if (!g) return -1;
  // the largest phase filter: ceil(R/sy) x ceil(S/sx) taps
  const int U = (g->r + g->stride_h - 1) / g->stride_h, V = (g->s + g->stride_w - 1) / g->stride_w;
  return ((int64_t)g->k * U * V * g->c * 2 + 255) / 256 * 256;
}

int b200_conv_dgrad_strided(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* workspace,
                            void* stream) {
  B200_CHECK_ARG(dy && w && dx && g && workspace, "conv_dgrad_strided: null pointer");
  B200_CHECK_ARG(b200_conv_supported(g, B200_CONV_DGRAD) == 2, "conv_dgrad_strided: geometry is not a general strided dgrad");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int sy = g->stride_h, sx = g->stride_w;
  // Stride phases: with h' = h + pad_t = sy*i + a, only the taps r = a + sy*u reach dx[h], from dy[i - u]. Each phase
  // is a stride-1 dgrad of the phase's sub-filter whose output rows go to every sy-th / sx-th pixel of dx.
  struct Ph { int a, b, U, V, i0, j0, I, J; };
  Ph phs[16];
  int n_ph = 0;
  bool empty = false;
  for (int a = 0; a < sy; ++a)
    for (int b = 0; b < sx; ++b) {
      Ph q;
      q.a = a;
      q.b = b;
      q.U = a < g->r ? (g->r - a + sy - 1) / sy : 0;
      q.V = b < g->s ? (g->s - b + sx - 1) / sx : 0;
      q.i0 = ceil_div_i(g->pad_t - a, sy);
      q.j0 = ceil_div_i(g->pad_l - b, sx);
      q.I = floor_div_i(g->h - 1 + g->pad_t - a, sy) - q.i0 + 1;
      q.J = floor_div_i(g->w - 1 + g->pad_l - b, sx) - q.j0 + 1;
      if (q.I <= 0 || q.J <= 0) continue;             // this phase has no pixel in the image
      if (q.U == 0 || q.V == 0) {
        empty = true;                                // pixels no tap reaches: zero gradient
        continue;
      }
      phs[n_ph++] = q;
    }
  if (empty) B200_CUDA(cudaMemsetAsync(dx, 0, (size_t)g->n * g->h * g->w * g->c * 2, st));
  const int cvec = g->c / 8;
  for (int i = 0; i < n_ph; ++i) {
    const Ph& q = phs[i];
    const int64_t total = (int64_t)g->k * q.U * q.V * cvec;
    const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
    phase_filter_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const uint4*>(w), reinterpret_cast<uint4*>(workspace), g->k,
                                                g->r, g->s, cvec, q.a, q.b, sy, sx, q.U, q.V);
    B200_LAUNCH_CHECK();
    b200_conv_geom gp = *g;
    gp.h = q.I;
    gp.w = q.J;
    gp.r = q.U;
    gp.s = q.V;
    gp.stride_h = gp.stride_w = 1;
    gp.pad_t = q.i0;
    gp.pad_l = q.j0;
    gp.pad_b = g->p - q.I + q.U - 1 - q.i0;          // whatever makes the stride-1 geometry consistent; unused
    gp.pad_r = g->q - q.J + q.V - 1 - q.j0;
    DgradPhase ph{sy, sx, sy * q.i0 + q.a - g->pad_t, sx * q.j0 + q.b - g->pad_l, g->h, g->w};
    if (int e = dgrad_impl(dy, static_cast<const b200_bf16*>(workspace), dx, &gp, &ph, stream)) return e;
  }
  return 0;
}

int b200_gemm_nn(const b200_bf16* a, int32_t lda, const b200_bf16* b, b200_bf16* out, int32_t m, int32_t n, int32_t k,
                 void* stream) {
  // out[m,n] = a[m,k] . b[k,n]; a row-major with leading dimension lda (>= k, zero padded past k is fine),
  // b row-major [k][n] read as an MN-major operand; n % 64 == 0.
  B200_CHECK_ARG(a && b && out, "gemm_nn: null pointer");
  B200_CHECK_ARG(m > 0 && k > 0 && n > 0 && n % 64 == 0 && lda % 8 == 0 && lda >= k, "gemm_nn: need n%%64==0, lda%%8==0");
  CUtensorMap ma, mb;
  GemmParams p{};
  p.M = m;
  p.N = n;
  p.block_n = n % 256 == 0 ? 256 : (n % 128 == 0 ? 128 : 64);
  if (tmap_tiled_2d(&ma, a, 2, (uint64_t)lda, (uint64_t)m, (uint64_t)lda * 2, kBlockK, kBlockM, true)) return -1;
  if (tmap_tiled_2d(&mb, b, 2, (uint64_t)n, (uint64_t)k, (uint64_t)n * 2, 64, kBlockK, true)) return -1;
  p.num_m_tiles = (m + kBlockM - 1) / kBlockM;
  p.ctas = plan_ctas(p.block_n, p.num_m_tiles, (k + kBlockK - 1) / kBlockK, true);
  p.num_n_tiles = n / p.block_n;
  p.cchunks = (k + kBlockK - 1) / kBlockK;
  p.num_k_blocks = p.cchunks;
  p.R = p.S = 1;
  p.b_mn_major = 1;
  p.b_row_len = n;
  p.out = out;
  p.ldo = n;
  return launch(ma, mb, p, static_cast<cudaStream_t>(stream));
}

int b200_gemm_nt_stats(const b200_bf16* a, const b200_bf16* b, void* out, int32_t m, int32_t n, int32_t k,
                       int32_t ldo, int32_t out_fp32, float* col_stats, void* stream);

int b200_gemm_nt(const b200_bf16* a, const b200_bf16* b, void* out, int32_t m, int32_t n, int32_t k, int32_t ldo,
                 int32_t out_fp32, void* stream) {
  return b200_gemm_nt_stats(a, b, out, m, n, k, ldo, out_fp32, nullptr, stream);
}

// (the scratch rows of the GEMM-with-statistics call are sized from (m, n) alone, so that call keeps one CTA per tile)
int64_t b200_gemm_stats_rows(int32_t m, int32_t n) { return grid_for_tiles(m, n, 0) * 9; }
int64_t b200_gemm_stats_first_result_row(int32_t m, int32_t n) { return grid_for_tiles(m, n, 0) * 8; }

int b200_gemm_nt_stats(const b200_bf16* a, const b200_bf16* b, void* out, int32_t m, int32_t n, int32_t k,
                       int32_t ldo, int32_t out_fp32, float* col_stats, void* stream) {
  B200_CHECK_ARG(a && b && out, "gemm_nt: null pointer");
  // k need not be a multiple of the 64-wide k-block: the last block's columns past k are zero-filled by TMA in both
  // operands (rows must stay 16-byte aligned, hence k % 8)
  B200_CHECK_ARG(m > 0 && n > 0 && k > 0 && k % 8 == 0 && ldo >= n, "gemm_nt: need k%%8==0 and ldo>=n");
  CUtensorMap ma, mb;
  GemmParams p{};
  p.M = m;
  p.N = n;
  p.block_n = pick_block_n(n);
  p.num_m_tiles = (m + kBlockM - 1) / kBlockM;
  const int kblocks = (k + kBlockK - 1) / kBlockK;
  p.ctas = col_stats ? 1 : plan_ctas(p.block_n, p.num_m_tiles, kblocks, false);
  if (tmap_tiled_2d(&ma, a, 2, (uint64_t)k, (uint64_t)m, (uint64_t)k * 2, kBlockK, kBlockM, true)) return -1;
  if (tmap_tiled_2d(&mb, b, 2, (uint64_t)k, (uint64_t)n, (uint64_t)k * 2, kBlockK, p.block_n / p.ctas, true)) return -1;
  p.num_n_tiles = (n + p.block_n - 1) / p.block_n;
  p.cchunks = kblocks;
  p.num_k_blocks = p.cchunks;
  p.R = p.S = 1;
  p.out = out;
  p.ldo = ldo;
  p.out_fp32 = out_fp32;
  p.col_stats = col_stats;
  return launch(ma, mb, p, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
