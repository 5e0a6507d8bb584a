// allreduce.cu — allreduce (+ fused 1/N scale and momentum-SGD / LARS update), all-gather, broadcast, barrier and
// point-to-point probes over NVLink peer memory. See include/b200comm.h for the contract and the reference citations
// (Horovod's MPI_Allreduce over OpenMPI/UCX or Intel MPI/libfabric, followed by TF ApplyMomentum; the OSU
// micro-benchmark roles).
//
// Algorithm (two-shot, per 512-element block b of the message, owner(b) = b mod world):
//   barrier                      every rank's gradients for this bucket are complete and visible
//   owner: acc = sum_p peer_p[b]  in rank order (peer loads over NVLink), acc *= scale
//          peer_p[b] = acc        for every p (peer stores): the all-gather
//   barrier                      all shards have landed everywhere
//   (fused) every rank applies the optimizer update to its local master/momentum/bf16 weights from the reduced
//           gradient now sitting in its own buffer - identical inputs, identical arithmetic, identical weights.
// Cross-GPU barriers are per CTA: CTA c of every rank exchanges an epoch number with CTA c of every other rank
// through st.release.sys / ld.acquire.sys on flag words in the peers' signal pads. Epochs live in device memory and
// only ever grow, so nothing is reset between calls and a captured CUDA graph can be replayed.
//
// What a per-CTA barrier proves, and what the code after it may therefore touch: when CTA c passes the second
// barrier it knows that CTA c of every rank has finished its reduce + all-gather, i.e. that the blocks
// { j * world + p : j = c, c + G, c + 2G, ... ; p = 0 .. world-1 } (G = gridDim.x) hold their final value in this
// rank's buffer. It knows nothing about blocks written by OTHER CTA indices of other ranks. The fused momentum-SGD
// half therefore walks exactly that block set. The LARS variant needs whole-tensor norms, so it first joins all
// CTAs of this rank with a device-local grid barrier (every CTA has then passed its cross-rank barrier, so every
// block is final) and only then reads across blocks.
//
// All kernels of this file that use the signal pad must be ordered on ONE stream per device (they share the per-CTA
// epoch slots); parallel/symm.py serialises them on the context's communication stream.
#include "b200comm.h"
#include "common.cuh"

namespace b200 {
namespace {

// 128 threads x one 16-byte vector = one 512-element block. Small CTAs on purpose: a convolution CTA (576 threads x 96
// registers, <= 223 KiB shared memory) leaves 10,240 registers and a few KiB of shared memory free on its SM, and a
// 128-thread CTA of these kernels (<= 80 registers, a few bytes of shared memory) fits into that gap, so a collective
// launched on the side stream makes progress WHILE a persistent convolution kernel holds every SM.
constexpr int kThreads = 128;
constexpr int kBlock = kThreads * 4;          // fp32 elements per block
constexpr int kBlockBf16 = kThreads * 8;      // bf16 elements per block
constexpr int kPadFlagWords = B200_COMM_MAX_CTAS * B200_COMM_MAX_RANKS;   // written by peers
constexpr int kPadEpochWords = B200_COMM_MAX_CTAS;                        // per-CTA epoch counters (local)
// local grid barrier (count, generation), then the point-to-point probe's mailboxes: per partner rank one flag word
// (written by that partner) and one local message count, so exchanges with different partners do not interfere
constexpr int kPadGridCnt = kPadFlagWords + kPadEpochWords;
constexpr int kPadGridGen = kPadGridCnt + 1;
constexpr int kPadP2PFlag = kPadGridCnt + 2;                              // one word per sending rank
constexpr int kPadP2PSeq = kPadP2PFlag + B200_COMM_MAX_RANKS;             // one word per partner rank
constexpr int kPadWords = kPadP2PSeq + B200_COMM_MAX_RANKS + 2;

struct CommDev {
  uint64_t buf[B200_COMM_MAX_RANKS];
  uint64_t pad[B200_COMM_MAX_RANKS];
  uint64_t mc;      // multicast address of the data buffer (0 = none)
  int rank, world;
};

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

constexpr uint64_t kWatchdogNs = 30000000000ull;   // 30 s: a peer never arrived; fail loudly instead of hanging

// All threads of the CTA call this. s_epoch is a shared-memory word holding this CTA's last used epoch.
__device__ __forceinline__ void cta_barrier_all_ranks(const CommDev& c, uint32_t* s_epoch) {
  // Every thread's peer stores happen-before the bar.sync; the release stores below are cumulative over what their
  // threads observed through that barrier, so the data is visible to a peer that acquires the flag (the PTX release
  // pattern; one fence per signalling thread instead of one per thread of the CTA).
  __syncthreads();
  const uint32_t e = *s_epoch + 1;
  if ((int)threadIdx.x < c.world) {
    const int t = threadIdx.x;
    uint32_t* dst = reinterpret_cast<uint32_t*>(c.pad[t]) + blockIdx.x * B200_COMM_MAX_RANKS + c.rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(e) : "memory");
    const uint32_t* src = reinterpret_cast<const uint32_t*>(c.pad[c.rank]) + blockIdx.x * B200_COMM_MAX_RANKS + t;
    uint32_t v;
    const uint64_t t0 = globaltimer_ns();
    for (;;) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
      if ((int32_t)(v - e) >= 0) break;
      if (globaltimer_ns() - t0 > kWatchdogNs) {
        printf("b200 comm: rank %d CTA %d timed out waiting for rank %d (epoch %u, saw %u)\n", c.rank, blockIdx.x, t, e, v);
        __trap();
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *s_epoch = e;
  __syncthreads();
}

__device__ __forceinline__ uint32_t* epoch_slot(const CommDev& c) {
  return reinterpret_cast<uint32_t*>(c.pad[c.rank]) + kPadFlagWords + blockIdx.x;
}

// Join all CTAs of THIS rank's launch (they are co-resident: each of them spins on peer flags anyway). Sense
// reversal on a generation word: every CTA reads the generation before it arrives, the last arriver resets the count
// and bumps the generation. Works for any grid size per launch because the count returns to zero.
__device__ __forceinline__ void grid_barrier_local(const CommDev& c) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* cnt = reinterpret_cast<uint32_t*>(c.pad[c.rank]) + kPadGridCnt;
    uint32_t* gen = reinterpret_cast<uint32_t*>(c.pad[c.rank]) + kPadGridGen;
    __threadfence();
    uint32_t g;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(gen) : "memory");
    if (atomicAdd(cnt, 1u) == gridDim.x - 1) {
      *cnt = 0;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(gen), "r"(g + 1) : "memory");
    } else {
      uint32_t v;
      const uint64_t t0 = globaltimer_ns();
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(gen) : "memory");
        if (globaltimer_ns() - t0 > kWatchdogNs) {
          printf("b200 comm: rank %d CTA %d timed out in the local grid barrier\n", c.rank, blockIdx.x);
          __trap();
        }
      } while (v == g);
    }
  }
  __syncthreads();
}

struct SgdDev {
  float* master;
  float* mom;
  b200_bf16* shadow;
  const int32_t* block_tensor;
  const float* tdecay;
  const float* tlars;
  float* trust;
  const float* lr_dev;      // when non-null the learning rate is read from device memory (schedules under a CUDA graph)
  float lr, mu, wd;
  int enabled;
  // LARS (fused): per-block (kBlock elements) partial norms, per-tensor block ranges, the tensors that END in this bucket and the
  // block range [upd_lo, upd_hi) they span (it may start in an earlier bucket when a tensor was split)
  float* block_partials;
  const int32_t* tblk_lo;
  const int32_t* tblk_cnt;
  int t_lo, t_hi;
  int64_t upd_lo, upd_hi;
  float eta, eps;
};

// Peer (and post-barrier local) data must not come from a stale L1 line: a relaxed load at system scope is a strong
// operation (served from the point of coherence, never from this SM's non-coherent L1), and no_allocate keeps it from
// leaving a line behind.
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

__device__ __forceinline__ void st_peer_f4(float* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ void sgd_apply_block(const SgdDev& sgd, const float4 gv, int64_t gb, int tid, float lr) {
  float dec = sgd.wd, tr = 1.0f;
  if (sgd.block_tensor) {
    const int t = sgd.block_tensor[(gb * kBlock) >> 10];   // the table has one entry per 1024 elements
    if (t < 0) return;                         // padding between tensors stays untouched
    dec = sgd.wd * sgd.tdecay[t];
    if (sgd.trust && sgd.tlars[t] != 0.0f) tr = sgd.trust[t];
  }
  const int64_t i = gb * kBlock + tid * 4;
  const float4 wv = *reinterpret_cast<const float4*>(sgd.master + i);
  const float4 mv = *reinterpret_cast<const float4*>(sgd.mom + i);
  float ww[4] = {wv.x, wv.y, wv.z, wv.w}, mm[4] = {mv.x, mv.y, mv.z, mv.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float ge = (gg[k] + dec * ww[k]) * tr;
    mm[k] = sgd.mu * mm[k] + ge;
    ww[k] -= lr * mm[k];
  }
  *reinterpret_cast<float4*>(sgd.master + i) = make_float4(ww[0], ww[1], ww[2], ww[3]);
  *reinterpret_cast<float4*>(sgd.mom + i) = make_float4(mm[0], mm[1], mm[2], mm[3]);
  if (sgd.shadow) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(ww[0], ww[1]), hi = __floats2bfloat162_rn(ww[2], ww[3]);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(sgd.shadow + i) = pk;
  }
}

// Four blocks per trip: all twelve loads (gradient, master, momentum) are issued before the first use, so a thread has
// twelve 16-byte requests in flight instead of three. `gb[u] < 0` marks an empty slot.
__device__ __forceinline__ void sgd_apply_blocks4(const SgdDev& sgd, const float* g, const int64_t gb[4], int tid, float lr) {
  float4 gv[4], wv[4], mv[4];
  float dec[4], tr[4];
  bool live[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    live[u] = gb[u] >= 0;
    dec[u] = sgd.wd;
    tr[u] = 1.0f;
    if (live[u] && sgd.block_tensor) {
      const int t = sgd.block_tensor[(gb[u] * kBlock) >> 10];
      if (t < 0) {
        live[u] = false;
      } else {
        dec[u] = sgd.wd * sgd.tdecay[t];
        if (sgd.trust && sgd.tlars[t] != 0.0f) tr[u] = sgd.trust[t];
      }
    }
    if (live[u]) {
      const int64_t i = gb[u] * kBlock + tid * 4;
      gv[u] = ld_peer_f4(g + i);
      wv[u] = *reinterpret_cast<const float4*>(sgd.master + i);
      mv[u] = *reinterpret_cast<const float4*>(sgd.mom + i);
    }
  }
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (!live[u]) continue;
    const int64_t i = gb[u] * kBlock + tid * 4;
    float ww[4] = {wv[u].x, wv[u].y, wv[u].z, wv[u].w}, mm[4] = {mv[u].x, mv[u].y, mv[u].z, mv[u].w};
    const float gg[4] = {gv[u].x, gv[u].y, gv[u].z, gv[u].w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float ge = (gg[k] + dec[u] * ww[k]) * tr[u];
      mm[k] = sgd.mu * mm[k] + ge;
      ww[k] -= lr * mm[k];
    }
    *reinterpret_cast<float4*>(sgd.master + i) = make_float4(ww[0], ww[1], ww[2], ww[3]);
    *reinterpret_cast<float4*>(sgd.mom + i) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    if (sgd.shadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(ww[0], ww[1]), hi = __floats2bfloat162_rn(ww[2], ww[3]);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(sgd.shadow + i) = pk;
    }
  }
}

// Sum over the CTA of two per-thread values, in a fixed order (warp tree, then warps 0..7 in sequence).
__device__ __forceinline__ float2 cta_sum2(float a, float b, float (*red)[8]) {
  a = warp_sum(a);
  b = warp_sum(b);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) {
    red[0][warp] = a;
    red[1][warp] = b;
  }
  __syncthreads();
  float sa = 0.0f, sb = 0.0f;
#pragma unroll
  for (int k = 0; k < kThreads / 32; ++k) {
    sa += red[0][k];
    sb += red[1][k];
  }
  return make_float2(sa, sb);
}

// kW: compile-time bound on the rank count (2, 4 or 8) so the two-shot loop can keep 8 / kW owned blocks x kW peers =
// eight peer loads in flight per thread in registers. <= 80 registers x 128 threads: see the note at kThreads.
template <bool kNvls, bool kLars, int kW>
__global__ void __launch_bounds__(kThreads, 6) allreduce_f32_kernel(CommDev c, int64_t offset, int64_t nblk, float scale,
                                                                    SgdDev sgd) {
  __shared__ uint32_t s_epoch;
  __shared__ float s_red[2][8];
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);

  const int tid = threadIdx.x;
  // ---- reduce-scatter + all-gather of the blocks this rank owns: b = j*world + rank, j strided over CTAs
  for (int64_t j = blockIdx.x;; j += gridDim.x) {
    const int64_t b = j * c.world + c.rank;
    if (b >= nblk) break;
    if (kNvls) {
      // NVLS: one load returns the sum over all ranks' copies, computed inside the NVSwitch; one store lands in
      // every rank's copy. Link traffic per GPU: S/N reduced in + S/N out, instead of (N-1)/N*S each way.
      // Four owned blocks per trip keep four switch round trips in flight per thread.
      float4 acc[4];
      float* mp[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t bu = (j + (int64_t)u * gridDim.x) * c.world + c.rank;
        mp[u] = bu < nblk ? reinterpret_cast<float*>(c.mc) + offset + bu * kBlock + tid * 4 : nullptr;
        if (mp[u])
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(acc[u].x), "=f"(acc[u].y), "=f"(acc[u].z), "=f"(acc[u].w)
                       : "l"(mp[u])
                       : "memory");
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (!mp[u]) continue;
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mp[u]), "f"(acc[u].x * scale),
                     "f"(acc[u].y * scale), "f"(acc[u].z * scale), "f"(acc[u].w * scale)
                     : "memory");
      }
      j += 3 * (int64_t)gridDim.x;      // the loop header advances by one more gridDim.x
      continue;
    }
    // Peer loads are latency-bound (about 2 us each over NVLink), so a thread keeps eight of them in flight: with
    // two ranks it takes four owned blocks per trip, with up to four ranks two, with more one.
    constexpr int kU = 8 / kW;
    float4 v[kU][kW];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t bu = (j + (int64_t)u * gridDim.x) * c.world + c.rank;
      if (bu < nblk) {
#pragma unroll
        for (int p = 0; p < kW; ++p)
          if (p < c.world) v[u][p] = ld_peer_f4(reinterpret_cast<const float*>(c.buf[p]) + offset + bu * kBlock + tid * 4);
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t bu = (j + (int64_t)u * gridDim.x) * c.world + c.rank;
      if (bu < nblk) {
        float4 acc = v[u][0];
#pragma unroll
        for (int p = 1; p < kW; ++p)
          if (p < c.world) {
            acc.x += v[u][p].x;
            acc.y += v[u][p].y;
            acc.z += v[u][p].z;
            acc.w += v[u][p].w;
          }
        acc.x *= scale;
        acc.y *= scale;
        acc.z *= scale;
        acc.w *= scale;
#pragma unroll
        for (int p = 0; p < kW; ++p)
          if (p < c.world) st_peer_f4(reinterpret_cast<float*>(c.buf[p]) + offset + bu * kBlock + tid * 4, acc);
      }
    }
    j += (int64_t)(kU - 1) * gridDim.x;      // the loop header advances by one more gridDim.x
  }
  cta_barrier_all_ranks(c, &s_epoch);

  // ---- fused optimizer half: every rank updates its replica from the (now identical) reduced gradient
  if (sgd.enabled) {
    const float* g = reinterpret_cast<const float*>(c.buf[c.rank]);
    const float lr = sgd.lr_dev ? __ldg(sgd.lr_dev) : sgd.lr;
    if (!kLars) {
      // exactly the blocks this CTA's barrier covers: j in {blockIdx.x + k*gridDim.x}, all owners p of that j
      // item q of this CTA = block (blockIdx.x + (q / world) * gridDim.x) * world + q % world; four items per trip
      const int64_t base_blk = offset / kBlock;
      for (int64_t q = 0;; q += 4) {
        if ((blockIdx.x + (q / c.world) * (int64_t)gridDim.x) * c.world >= nblk) break;
        int64_t gb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t qq = q + u;
          const int64_t b = (blockIdx.x + (qq / c.world) * (int64_t)gridDim.x) * c.world + qq % c.world;
          gb[u] = b < nblk ? base_blk + b : -1;
        }
        sgd_apply_blocks4(sgd, g, gb, tid, lr);
      }
    } else {
      // (1) per-block partial norms of this bucket's blocks (any CTA may take any block after the grid barrier...
      //     but the partials are cheap, so they are taken right here for the covered set, then everyone joins)
      for (int64_t j = blockIdx.x; j * c.world < nblk; j += gridDim.x) {
        for (int p = 0; p < c.world; ++p) {
          const int64_t b = j * c.world + p;
          if (b >= nblk) break;
          const int64_t gb = offset / kBlock + b;
          const int64_t i = gb * kBlock + tid * 4;
          const float4 gv = ld_peer_f4(g + i);
          const float4 wv = *reinterpret_cast<const float4*>(sgd.master + i);
          const float2 s = cta_sum2(wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w,
                                    gv.x * gv.x + gv.y * gv.y + gv.z * gv.z + gv.w * gv.w, s_red);
          if (tid == 0) {
            sgd.block_partials[2 * gb] = s.x;
            sgd.block_partials[2 * gb + 1] = s.y;
          }
        }
      }
      grid_barrier_local(c);
      // (2) trust ratio of every tensor that ends in this bucket: its blocks' partials folded in a fixed order
      for (int t = sgd.t_lo + blockIdx.x; t < sgd.t_hi; t += gridDim.x) {
        const int lo = sgd.tblk_lo[t], cnt = sgd.tblk_cnt[t];
        float sw = 0.0f, sg = 0.0f;
        for (int b = tid; b < cnt; b += kThreads) {
          sw += __ldcg(sgd.block_partials + 2 * (lo + b));
          sg += __ldcg(sgd.block_partials + 2 * (lo + b) + 1);
        }
        const float2 s = cta_sum2(sw, sg, s_red);
        if (tid == 0) {
          float r = 1.0f;
          if (sgd.tlars[t] != 0.0f) {
            const float wn = sqrtf(s.x), gn = sqrtf(s.y);
            if (wn > 0.0f && gn > 0.0f) r = sgd.eta * wn / (gn + sgd.wd * sgd.tdecay[t] * wn + sgd.eps);
          }
          sgd.trust[t] = r;
        }
      }
      grid_barrier_local(c);
      // (3) update the blocks of those tensors (the range may reach back into an earlier bucket)
      for (int64_t gb0 = sgd.upd_lo + blockIdx.x; gb0 < sgd.upd_hi; gb0 += 4 * (int64_t)gridDim.x) {
        int64_t gb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t x = gb0 + (int64_t)u * gridDim.x;
          gb[u] = x < sgd.upd_hi ? x : -1;
        }
        sgd_apply_blocks4(sgd, g, gb, tid, lr);
      }
    }
  }
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// One-shot variant for small messages (one block per CTA): every rank reads the block from all peers and
// reduces it itself in rank order - a single NVLink traversal instead of two. The second barrier keeps a fast rank
// from overwriting its buffer while a slower peer is still reading it. n only has to be a multiple of 4 elements
// here (the tail block is masked), which is what lets the sweep go below 4 KiB without padding in the tool.
__global__ void __launch_bounds__(kThreads) allreduce_f32_oneshot_kernel(CommDev c, int64_t offset, int64_t n,
                                                                         float scale, SgdDev sgd) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);
  const int tid = threadIdx.x;
  const int64_t b = blockIdx.x;              // grid == number of blocks
  const int64_t off = offset + b * kBlock + tid * 4;
  const bool live = b * kBlock + tid * 4 < n;
  float4 v[B200_COMM_MAX_RANKS];
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (live) {
#pragma unroll
    for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) v[p] = ld_peer_f4(reinterpret_cast<const float*>(c.buf[p]) + off);
    acc = v[0];
#pragma unroll
    for (int p = 1; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) {
        acc.x += v[p].x;
        acc.y += v[p].y;
        acc.z += v[p].z;
        acc.w += v[p].w;
      }
    acc.x *= scale;
    acc.y *= scale;
    acc.z *= scale;
    acc.w *= scale;
  }
  cta_barrier_all_ranks(c, &s_epoch);
  if (live) *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.buf[c.rank]) + off) = acc;
  if (sgd.enabled && live) {
    const float lr = sgd.lr_dev ? __ldg(sgd.lr_dev) : sgd.lr;
    sgd_apply_block(sgd, acc, offset / kBlock + b, tid, lr);
  }
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// bf16 payload: 1024 elements per block (128 threads x 8), fp32 accumulation, one rounding. Same eight-loads-in-flight
// structure as the fp32 two-shot loop (kW = compile-time bound on the rank count).
template <int kW>
__global__ void __launch_bounds__(kThreads, 6) allreduce_bf16_kernel(CommDev c, int64_t offset, int64_t nblk, float scale) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);
  const int tid = threadIdx.x;
  constexpr int kU = 8 / kW;
  for (int64_t j = blockIdx.x;; j += (int64_t)kU * gridDim.x) {
    if (j * c.world + c.rank >= nblk) break;
    float4 raw[kU][kW];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t bu = (j + (int64_t)u * gridDim.x) * c.world + c.rank;
      if (bu < nblk) {
#pragma unroll
        for (int p = 0; p < kW; ++p)
          if (p < c.world)
            raw[u][p] = ld_peer_f4(reinterpret_cast<const float*>(reinterpret_cast<const b200_bf16*>(c.buf[p]) + offset +
                                                                  bu * kBlockBf16 + tid * 8));
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t bu = (j + (int64_t)u * gridDim.x) * c.world + c.rank;
      if (bu < nblk) {
        float acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
#pragma unroll
        for (int p = 0; p < kW; ++p)
          if (p < c.world) {
            float f[8];
            unpack8(*reinterpret_cast<const bf16x8*>(&raw[u][p]), f);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += f[k];
          }
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] *= scale;
        const bf16x8 out = pack8(acc);
        const float4 o4 = *reinterpret_cast<const float4*>(&out);
#pragma unroll
        for (int p = 0; p < kW; ++p)
          if (p < c.world)
            st_peer_f4(reinterpret_cast<float*>(reinterpret_cast<b200_bf16*>(c.buf[p]) + offset + bu * kBlockBf16 + tid * 8), o4);
      }
    }
  }
  cta_barrier_all_ranks(c, &s_epoch);
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

__global__ void __launch_bounds__(kThreads) barrier_kernel(CommDev c) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// Copy 16-byte vectors [0, nvec) of `src` to the same offset of every peer's buffer (or of the multicast mapping, one
// store for all of them). `skip` = a rank that needs no copy (the source's own buffer), -1 = none.
__device__ __forceinline__ void push_to_peers(const CommDev& c, const float* src, int64_t offset, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  const bool mc = c.mc != 0 && c.world > 2;      // with two ranks a plain peer store is the same traffic, minus the echo
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += 8 * stride) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (i + u * stride < nvec) v[u] = *reinterpret_cast<const float4*>(src + (i + u * stride) * 4);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int64_t e = i + u * stride;
      if (e >= nvec) continue;
      if (mc) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(
                         reinterpret_cast<float*>(c.mc) + offset + e * 4),
                     "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w)
                     : "memory");
      } else {
#pragma unroll
        for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
          if (p < c.world && p != c.rank) st_peer_f4(reinterpret_cast<float*>(c.buf[p]) + offset + e * 4, v[u]);
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads) broadcast_f32_kernel(CommDev c, int64_t offset, int64_t nvec, int root) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);     // peers are done with whatever they were doing to these buffers
  if (c.rank == root) push_to_peers(c, reinterpret_cast<const float*>(c.buf[root]) + offset, offset, nvec);
  cta_barrier_all_ranks(c, &s_epoch);
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// All-gather: rank r's segment [offset + r*n, +n) of its own buffer is written to the same place in every peer.
__global__ void __launch_bounds__(kThreads) allgather_f32_kernel(CommDev c, int64_t offset, int64_t nvec_per_rank) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);
  const int64_t seg = offset + (int64_t)c.rank * nvec_per_rank * 4;
  push_to_peers(c, reinterpret_cast<const float*>(c.buf[c.rank]) + seg, seg, nvec_per_rank);
  cta_barrier_all_ranks(c, &s_epoch);
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// Point-to-point bandwidth probe (osu_bw / osu_bibw roles): this rank stores (or loads) nvec 16-byte vectors to
// (from) ONE peer. No barrier inside: the caller brackets it with events; run it on one rank for one direction, on
// both ranks of a pair for the bidirectional figure.
__global__ void __launch_bounds__(kThreads) p2p_copy_kernel(CommDev c, int peer, int64_t offset, int64_t nvec, int pull) {
  const float* src = reinterpret_cast<const float*>(c.buf[pull ? peer : c.rank]) + offset;
  float* dst = reinterpret_cast<float*>(c.buf[pull ? c.rank : peer]) + offset;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  // eight 16-byte requests in flight per thread: a pull has to cover ~2 us of NVLink round trip at ~770 GB/s
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += 8 * stride) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (i + u * stride < nvec) v[u] = ld_peer_f4(src + (i + u * stride) * 4);
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (i + u * stride < nvec) st_peer_f4(dst + (i + u * stride) * 4, v[u]);
  }
}

// Point-to-point latency probe (osu_latency role): ranks a and b bounce a message of nvec vectors `iters` times;
// one CTA each, every other rank returns at once. Per bounce: payload stores, fence, flag store (release.sys); the
// receiver spins on the flag (acquire.sys), reads the payload, and answers. Rank a writes the mean ONE-WAY time in
// microseconds to out_us[0], measured with %globaltimer on the device.
__global__ void __launch_bounds__(kThreads) p2p_pingpong_kernel(CommDev c, int a, int b, int64_t offset, int nvec,
                                                                int iters, float* out_us) {
  if (c.rank != a && c.rank != b) return;
  const int peer = c.rank == a ? b : a;
  uint32_t* my_flag = reinterpret_cast<uint32_t*>(c.pad[c.rank]) + kPadP2PFlag + peer;
  uint32_t* peer_flag = reinterpret_cast<uint32_t*>(c.pad[peer]) + kPadP2PFlag + c.rank;
  uint32_t* seq_slot = reinterpret_cast<uint32_t*>(c.pad[c.rank]) + kPadP2PSeq + peer;
  __shared__ uint32_t s_seq;
  if (threadIdx.x == 0) s_seq = *seq_slot;
  __syncthreads();
  uint32_t seq = s_seq;
  float* mine = reinterpret_cast<float*>(c.buf[c.rank]) + offset;
  float* theirs = reinterpret_cast<float*>(c.buf[peer]) + offset;
  float acc = 0.0f;
  auto send = [&]() {
    for (int i = threadIdx.x; i < nvec; i += kThreads) st_peer_f4(theirs + i * 4, make_float4(acc, 1.f, 2.f, 3.f));
    __syncthreads();                 // the release store of thread 0 below is cumulative over the CTA's payload stores
    ++seq;
    if (threadIdx.x == 0) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer_flag), "r"(seq) : "memory");
  };
  auto recv = [&]() {
    if (threadIdx.x == 0) {
      uint32_t v;
      const uint64_t t0 = globaltimer_ns();
      do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(my_flag) : "memory");
        if (globaltimer_ns() - t0 > kWatchdogNs) {
          printf("b200 comm: rank %d timed out in the ping-pong probe\n", c.rank);
          __trap();
        }
      } while ((int32_t)(v - (seq + 1)) < 0);
    }
    __syncthreads();
    ++seq;
    for (int i = threadIdx.x; i < nvec; i += kThreads) acc += ld_peer_f4(mine + i * 4).x;
  };
  // both sides count every message (sent or received) in `seq`, so a flag value names one message of the exchange
  const uint64_t t0 = globaltimer_ns();
  for (int it = 0; it < iters; ++it) {
    if (c.rank == a) {
      send();
      recv();
    } else {
      recv();
      send();
    }
  }
  const uint64_t t1 = globaltimer_ns();
  if (threadIdx.x == 0) {
    *seq_slot = seq;
    if (acc == -1.2345f) out_us[1] = acc;      // keeps the payload reads alive
    if (c.rank == a && out_us) out_us[0] = (float)((double)(t1 - t0) / (2.0 * iters) * 1e-3);
  }
}

int to_dev(const b200_comm* comm, CommDev& d, int& ctas) {
  B200_CHECK_ARG(comm != nullptr, "comm: null descriptor");
  B200_CHECK_ARG(comm->world >= 1 && comm->world <= B200_COMM_MAX_RANKS && comm->rank >= 0 && comm->rank < comm->world,
                 "comm: bad rank/world (%d/%d)", comm->rank, comm->world);
  for (int i = 0; i < B200_COMM_MAX_RANKS; ++i) {
    d.buf[i] = comm->buf[i];
    d.pad[i] = comm->pad[i];
    if (i < comm->world) B200_CHECK_ARG(d.buf[i] && d.pad[i], "comm: missing peer pointer for rank %d", i);
  }
  d.mc = comm->mc;
  d.rank = comm->rank;
  d.world = comm->world;
  ctas = comm->ctas > 0 ? comm->ctas : 64;
  if (ctas > B200_COMM_MAX_CTAS) ctas = B200_COMM_MAX_CTAS;
  return 0;
}

// CTAs for a streaming kernel over `blocks` 2 KiB units: no more than the descriptor allows, no more than one per four
// units (small messages are latency-bound; more CTAs only add flag traffic), at least one.
int ctas_for(int64_t blocks, int cap) {
  int64_t want = (blocks + 3) / 4;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int64_t b200_comm_pad_bytes(void) { return (int64_t)kPadWords * 4; }

int32_t b200_comm_block_elems(void) { return kBlock; }

int b200_symm_alloc(int64_t bytes, void** dptr) {
  B200_CHECK_ARG(bytes > 0 && dptr, "symm_alloc: bad args");
  B200_CUDA(cudaMalloc(dptr, (size_t)bytes));
  B200_CUDA(cudaMemset(*dptr, 0, (size_t)bytes));
  return 0;
}

int b200_symm_free(void* dptr) {
  if (dptr) B200_CUDA(cudaFree(dptr));
  return 0;
}

int b200_symm_get_handle(void* dptr, void* handle_out) {
  B200_CHECK_ARG(dptr && handle_out, "symm_get_handle: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == B200_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  B200_CUDA(cudaIpcGetMemHandle(&h, dptr));
  memcpy(handle_out, &h, sizeof h);
  return 0;
}

int b200_symm_open_handle(const void* handle, void** peer_ptr) {
  B200_CHECK_ARG(handle && peer_ptr, "symm_open_handle: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  B200_CUDA(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int b200_symm_close_handle(void* peer_ptr) {
  if (peer_ptr) B200_CUDA(cudaIpcCloseMemHandle(peer_ptr));
  return 0;
}

int b200_allreduce_f32(const b200_comm* comm, int64_t offset, int64_t n, float scale, const b200_sgd_args* sgd,
                       void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int algo = comm->algo;
  B200_CHECK_ARG(algo >= B200_COMM_ALGO_AUTO && algo <= B200_COMM_ALGO_NVLS, "allreduce_f32: unknown algorithm %d", algo);
  B200_CHECK_ARG(n > 0 && n % 4 == 0 && offset >= 0 && offset % 4 == 0, "allreduce_f32: offset and n must be multiples of 4 elements");
  SgdDev s{};
  bool lars = false;
  if (sgd) {
    B200_CHECK_ARG(sgd->master && sgd->mom, "allreduce_f32: fused update needs master and momentum buffers");
    B200_CHECK_ARG(!sgd->block_tensor || sgd->tensor_decay, "allreduce_f32: block table needs tensor_decay");
    B200_CHECK_ARG(!sgd->trust || sgd->tensor_lars, "allreduce_f32: trust needs tensor_lars");
    B200_CHECK_ARG(n % 1024 == 0 && offset % 1024 == 0, "allreduce_f32: the fused update works on whole 1024-element blocks");
    s.master = sgd->master;
    s.mom = sgd->mom;
    s.shadow = sgd->shadow;
    s.block_tensor = sgd->block_tensor;
    s.tdecay = sgd->tensor_decay;
    s.tlars = sgd->tensor_lars;
    s.trust = sgd->trust;
    s.lr_dev = sgd->lr_dev;
    s.lr = sgd->lr;
    s.mu = sgd->mu;
    s.wd = sgd->wd;
    s.enabled = 1;
    if (sgd->lars_block_partials) {
      B200_CHECK_ARG(sgd->trust && sgd->block_tensor && sgd->tensor_blk_lo && sgd->tensor_blk_cnt,
                     "allreduce_f32: fused LARS needs trust, block_tensor and the per-tensor block ranges");
      B200_CHECK_ARG(sgd->lars_t_lo >= 0 && sgd->lars_t_hi >= sgd->lars_t_lo && sgd->lars_upd_lo >= 0 &&
                         sgd->lars_upd_hi >= sgd->lars_upd_lo && sgd->lars_upd_hi * kBlock <= offset + n,
                     "allreduce_f32: fused LARS tensor / block range is inconsistent");
      lars = true;
      s.block_partials = sgd->lars_block_partials;
      s.tblk_lo = sgd->tensor_blk_lo;
      s.tblk_cnt = sgd->tensor_blk_cnt;
      s.t_lo = sgd->lars_t_lo;
      s.t_hi = sgd->lars_t_hi;
      s.upd_lo = sgd->lars_upd_lo;
      s.upd_hi = sgd->lars_upd_hi;
      s.eta = sgd->lars_eta;
      s.eps = sgd->lars_eps;
    }
  }
  const int64_t nblk = (n + kBlock - 1) / kBlock;
  const int64_t one_shot_max = comm->one_shot_max_bytes > 0 ? comm->one_shot_max_bytes : 0;
  const bool one_shot = !lars && nblk <= B200_COMM_MAX_CTAS &&
                        (algo == B200_COMM_ALGO_ONE_SHOT || (algo == B200_COMM_ALGO_AUTO && n * 4 <= one_shot_max));
  if (one_shot) {
    allreduce_f32_oneshot_kernel<<<(int)nblk, kThreads, 0, st>>>(d, offset, n, scale, s);
    B200_LAUNCH_CHECK();
    return 0;
  }
  B200_CHECK_ARG(n % 1024 == 0 && offset % 1024 == 0,
                 "allreduce_f32: beyond the one-shot range offset and n must be multiples of 1024 elements");
  // each rank owns nblk/world blocks; the NVLS loop takes four per trip
  ctas = ctas_for((nblk + d.world - 1) / d.world, ctas);
  bool nvls = comm->mc != 0 && (algo == B200_COMM_ALGO_NVLS ||
                                (algo == B200_COMM_ALGO_AUTO && comm->nvls_min_bytes >= 0 && n * 4 >= comm->nvls_min_bytes));
  B200_CHECK_ARG(algo != B200_COMM_ALGO_NVLS || nvls, "allreduce_f32: NVLS requested but the buffer has no multicast mapping");
#define B200_AR_LAUNCH(NVLS, LARS, W) allreduce_f32_kernel<NVLS, LARS, W><<<ctas, kThreads, 0, st>>>(d, offset, nblk, scale, s)
  if (nvls) {
    if (lars) B200_AR_LAUNCH(true, true, 8);
    else B200_AR_LAUNCH(true, false, 8);
  } else if (d.world <= 2) {
    if (lars) B200_AR_LAUNCH(false, true, 2);
    else B200_AR_LAUNCH(false, false, 2);
  } else if (d.world <= 4) {
    if (lars) B200_AR_LAUNCH(false, true, 4);
    else B200_AR_LAUNCH(false, false, 4);
  } else {
    if (lars) B200_AR_LAUNCH(false, true, 8);
    else B200_AR_LAUNCH(false, false, 8);
  }
#undef B200_AR_LAUNCH
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_allreduce_bf16(const b200_comm* comm, int64_t offset, int64_t n, float scale, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  B200_CHECK_ARG(n > 0 && n % 2048 == 0 && offset >= 0 && offset % 2048 == 0,
                 "allreduce_bf16: offset and n must be multiples of 2048 elements");
  const int64_t nblk = n / kBlockBf16;
  ctas = ctas_for((nblk + d.world - 1) / d.world, ctas);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d.world <= 2) allreduce_bf16_kernel<2><<<ctas, kThreads, 0, st>>>(d, offset, nblk, scale);
  else if (d.world <= 4) allreduce_bf16_kernel<4><<<ctas, kThreads, 0, st>>>(d, offset, nblk, scale);
  else allreduce_bf16_kernel<8><<<ctas, kThreads, 0, st>>>(d, offset, nblk, scale);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_comm_barrier(const b200_comm* comm, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  barrier_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_broadcast_f32(const b200_comm* comm, int64_t offset, int64_t n, int32_t root, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  B200_CHECK_ARG(n > 0 && n % 4 == 0 && offset >= 0 && offset % 4 == 0 && root >= 0 && root < d.world,
                 "broadcast_f32: n and offset must be multiples of 4, root in range");
  ctas = ctas_for(n / kBlock, B200_COMM_MAX_CTAS);
  broadcast_f32_kernel<<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, n / 4, root);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_allgather_f32(const b200_comm* comm, int64_t offset, int64_t n_per_rank, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  B200_CHECK_ARG(n_per_rank > 0 && n_per_rank % 4 == 0 && offset >= 0 && offset % 4 == 0,
                 "allgather_f32: n_per_rank and offset must be multiples of 4 elements");
  ctas = ctas_for(n_per_rank / kBlock, B200_COMM_MAX_CTAS);
  allgather_f32_kernel<<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, n_per_rank / 4);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_p2p_copy(const b200_comm* comm, int32_t peer, int64_t offset, int64_t n, int32_t pull, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  B200_CHECK_ARG(peer >= 0 && peer < d.world && peer != d.rank && n > 0 && n % 4 == 0 && offset >= 0 && offset % 4 == 0,
                 "p2p_copy: peer must be another rank; n and offset multiples of 4 elements");
  ctas = ctas_for(n / kBlock, B200_COMM_MAX_CTAS);
  p2p_copy_kernel<<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, peer, offset, n / 4, pull);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_p2p_pingpong(const b200_comm* comm, int32_t rank_a, int32_t rank_b, int64_t offset, int64_t n, int32_t iters,
                      float* out_us, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  B200_CHECK_ARG(rank_a >= 0 && rank_a < d.world && rank_b >= 0 && rank_b < d.world && rank_a != rank_b && iters > 0 &&
                     n >= 0 && n % 4 == 0 && n <= (1 << 22) && offset >= 0 && offset % 4 == 0,
                 "p2p_pingpong: two distinct ranks, n a multiple of 4 elements (<= 4 Mi), iters > 0");
  p2p_pingpong_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, rank_a, rank_b, offset, (int)(n / 4), iters,
                                                                           out_us);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
