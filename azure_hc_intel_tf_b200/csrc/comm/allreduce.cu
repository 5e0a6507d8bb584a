// allreduce.cu — allreduce (+ fused 1/N scale and momentum-SGD update), broadcast and barrier over NVLink peer
// memory. See include/b200comm.h for the contract and the reference citations (Horovod's MPI_Allreduce over
// OpenMPI/UCX or Intel MPI/libfabric, followed by TF ApplyMomentum).
//
// Algorithm (two-shot, per 1024-element block b of the message, owner(b) = b mod world):
//   barrier                      every rank's gradients for this bucket are complete and visible
//   owner: acc = sum_p peer_p[b]  in rank order (peer loads over NVLink), acc *= scale
//          peer_p[b] = acc        for every p (peer stores): the all-gather
//   barrier                      all shards have landed everywhere
//   (fused) every rank applies the optimizer update to its local master/momentum/bf16 weights from the reduced
//           gradient now sitting in its own buffer - identical inputs, identical arithmetic, identical weights.
// Cross-GPU barriers are per CTA: CTA c of every rank exchanges an epoch number with CTA c of every other rank
// through st.release.sys / ld.acquire.sys on flag words in the peers' signal pads. Epochs live in device memory and
// only ever grow, so nothing is reset between calls and a captured CUDA graph can be replayed.
#include "b200comm.h"
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kThreads = 256;
constexpr int kPadFlagWords = B200_COMM_MAX_CTAS * B200_COMM_MAX_RANKS;   // written by peers
constexpr int kPadWords = kPadFlagWords + B200_COMM_MAX_CTAS;             // + per-CTA epoch counters (local)

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

// All threads of the CTA call this. s_epoch is a shared-memory word holding this CTA's last used epoch.
__device__ __forceinline__ void cta_barrier_all_ranks(const CommDev& c, uint32_t* s_epoch) {
  __threadfence_system();          // this thread's peer stores are performed before the CTA signals
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
      if (globaltimer_ns() - t0 > 30000000000ull) {   // 30 s: a peer never arrived; fail loudly instead of hanging
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

struct SgdDev {
  float* master;
  float* mom;
  b200_bf16* shadow;
  const int32_t* block_tensor;
  const float* tdecay;
  const float* tlars;
  const float* trust;
  float lr, mu, wd;
  int enabled;
};

__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  // peer (and post-barrier local) data must not come from a stale L1 line
  float4 v;
  asm volatile("ld.global.cv.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

template <bool kNvls>
__global__ void __launch_bounds__(kThreads) allreduce_f32_kernel(CommDev c, int64_t offset, int64_t nblk, float scale,
                                                                 SgdDev sgd) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);

  const int tid = threadIdx.x;
  // ---- reduce-scatter + all-gather of the blocks this rank owns: b = j*world + rank, j strided over CTAs
  for (int64_t j = blockIdx.x;; j += gridDim.x) {
    const int64_t b = j * c.world + c.rank;
    if (b >= nblk) break;
    const int64_t off = offset + b * 1024 + tid * 4;
    if (kNvls) {
      // NVLS: one load returns the sum over all ranks' copies, computed inside the NVSwitch; one store lands in
      // every rank's copy. Link traffic per GPU: S/N reduced in + S/N out, instead of (N-1)/N*S each way.
      // Four owned blocks per trip keep four switch round trips in flight per thread.
      float4 acc[4];
      float* mp[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t bu = (j + (int64_t)u * gridDim.x) * c.world + c.rank;
        mp[u] = bu < nblk ? reinterpret_cast<float*>(c.mc) + offset + bu * 1024 + tid * 4 : nullptr;
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
    float4 v[B200_COMM_MAX_RANKS];
#pragma unroll
    for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) v[p] = ld_peer_f4(reinterpret_cast<const float*>(c.buf[p]) + off);
    float4 acc = v[0];
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
#pragma unroll
    for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.buf[p]) + off) = acc;
  }
  cta_barrier_all_ranks(c, &s_epoch);

  // ---- fused optimizer half: every rank updates its replica from the (now identical) reduced gradient
  if (sgd.enabled) {
    const float* g = reinterpret_cast<const float*>(c.buf[c.rank]);
    for (int64_t b = blockIdx.x; b < nblk; b += gridDim.x) {
      const int64_t gb = offset / 1024 + b;          // global 1024-block index
      float dec = sgd.wd, tr = 1.0f;
      if (sgd.block_tensor) {
        const int t = sgd.block_tensor[gb];
        if (t < 0) continue;
        dec = sgd.wd * sgd.tdecay[t];
        if (sgd.trust && sgd.tlars[t] != 0.0f) tr = sgd.trust[t];
      }
      const int64_t i = gb * 1024 + tid * 4;
      const float4 gv = ld_peer_f4(g + i);
      float4 wv = *reinterpret_cast<const float4*>(sgd.master + i);
      float4 mv = *reinterpret_cast<const float4*>(sgd.mom + i);
      float ww[4] = {wv.x, wv.y, wv.z, wv.w}, mm[4] = {mv.x, mv.y, mv.z, mv.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float ge = (gg[k] + dec * ww[k]) * tr;
        mm[k] = sgd.mu * mm[k] + ge;
        ww[k] -= sgd.lr * mm[k];
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
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// One-shot variant for small messages (one 1024-element block per CTA): every rank reads the block from all peers and
// reduces it itself in rank order - a single NVLink traversal instead of two. The second barrier keeps a fast rank
// from overwriting its buffer while a slower peer is still reading it.
__global__ void __launch_bounds__(kThreads) allreduce_f32_oneshot_kernel(CommDev c, int64_t offset, int64_t nblk,
                                                                         float scale, SgdDev sgd) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);
  const int tid = threadIdx.x;
  const int64_t b = blockIdx.x;              // grid == nblk
  const int64_t off = offset + b * 1024 + tid * 4;
  float4 v[B200_COMM_MAX_RANKS];
#pragma unroll
  for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
    if (p < c.world) v[p] = ld_peer_f4(reinterpret_cast<const float*>(c.buf[p]) + off);
  float4 acc = v[0];
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
  cta_barrier_all_ranks(c, &s_epoch);
  *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.buf[c.rank]) + off) = acc;
  if (sgd.enabled) {
    const int64_t gb = offset / 1024 + b;
    float dec = sgd.wd, tr = 1.0f;
    bool skip = false;
    if (sgd.block_tensor) {
      const int t = sgd.block_tensor[gb];
      if (t < 0) {
        skip = true;
      } else {
        dec = sgd.wd * sgd.tdecay[t];
        if (sgd.trust && sgd.tlars[t] != 0.0f) tr = sgd.trust[t];
      }
    }
    if (!skip) {
      const int64_t i = gb * 1024 + tid * 4;
      float4 wv = *reinterpret_cast<const float4*>(sgd.master + i);
      float4 mv = *reinterpret_cast<const float4*>(sgd.mom + i);
      float ww[4] = {wv.x, wv.y, wv.z, wv.w}, mm[4] = {mv.x, mv.y, mv.z, mv.w}, gg[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float ge = (gg[k] + dec * ww[k]) * tr;
        mm[k] = sgd.mu * mm[k] + ge;
        ww[k] -= sgd.lr * mm[k];
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
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
}

// bf16 payload: 2048 elements per block (256 threads x 8), fp32 accumulation, one rounding.
__global__ void __launch_bounds__(kThreads) allreduce_bf16_kernel(CommDev c, int64_t offset, int64_t nblk, float scale) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);
  const int tid = threadIdx.x;
  for (int64_t j = blockIdx.x;; j += gridDim.x) {
    const int64_t b = j * c.world + c.rank;
    if (b >= nblk) break;
    const int64_t off = offset + b * 2048 + tid * 8;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    bf16x8 v[B200_COMM_MAX_RANKS];
#pragma unroll
    for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) {
        float4 raw = ld_peer_f4(reinterpret_cast<const float*>(reinterpret_cast<const b200_bf16*>(c.buf[p]) + off));
        v[p] = *reinterpret_cast<bf16x8*>(&raw);
      }
#pragma unroll
    for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) {
        float f[8];
        unpack8(v[p], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
      }
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] *= scale;
    const bf16x8 out = pack8(acc);
#pragma unroll
    for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
      if (p < c.world) st8(reinterpret_cast<b200_bf16*>(c.buf[p]) + off, out);
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

__global__ void __launch_bounds__(kThreads) broadcast_f32_kernel(CommDev c, int64_t offset, int64_t nvec, int root) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = *epoch_slot(c);
  __syncthreads();
  cta_barrier_all_ranks(c, &s_epoch);     // peers are done with whatever they were doing to these buffers
  if (c.rank == root) {
    const float* src = reinterpret_cast<const float*>(c.buf[root]) + offset;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * kThreads) {
      const float4 v = *reinterpret_cast<const float4*>(src + i * 4);
#pragma unroll
      for (int p = 0; p < B200_COMM_MAX_RANKS; ++p)
        if (p < c.world && p != root) *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.buf[p]) + offset + i * 4) = v;
    }
  }
  cta_barrier_all_ranks(c, &s_epoch);
  if (threadIdx.x == 0) *epoch_slot(c) = s_epoch;
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

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int64_t b200_comm_pad_bytes(void) { return (int64_t)kPadWords * 4; }

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
  B200_CHECK_ARG(n > 0 && n % 1024 == 0 && offset >= 0 && offset % 1024 == 0,
                 "allreduce_f32: offset and n must be multiples of 1024 elements");
  SgdDev s{};
  if (sgd) {
    B200_CHECK_ARG(sgd->master && sgd->mom, "allreduce_f32: fused update needs master and momentum buffers");
    B200_CHECK_ARG(!sgd->block_tensor || sgd->tensor_decay, "allreduce_f32: block table needs tensor_decay");
    B200_CHECK_ARG(!sgd->trust || sgd->tensor_lars, "allreduce_f32: trust needs tensor_lars");
    s = SgdDev{sgd->master, sgd->mom, sgd->shadow, sgd->block_tensor, sgd->tensor_decay, sgd->tensor_lars, sgd->trust,
               sgd->lr, sgd->mu, sgd->wd, 1};
  }
  const int64_t nblk = n / 1024;
  const int64_t one_shot_max = comm->one_shot_max_bytes > 0 ? comm->one_shot_max_bytes : 0;
  if (n * 4 <= one_shot_max && nblk <= B200_COMM_MAX_CTAS) {
    allreduce_f32_oneshot_kernel<<<(int)nblk, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, nblk, scale, s);
    B200_LAUNCH_CHECK();
    return 0;
  }
  if (nblk < ctas) ctas = (int)nblk;
  if (comm->mc && comm->nvls_min_bytes >= 0 && n * 4 >= comm->nvls_min_bytes && comm->world > 2)
    allreduce_f32_kernel<true><<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, nblk, scale, s);
  else
    allreduce_f32_kernel<false><<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, nblk, scale, s);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_allreduce_bf16(const b200_comm* comm, int64_t offset, int64_t n, float scale, void* stream) {
  CommDev d;
  int ctas;
  if (int e = to_dev(comm, d, ctas)) return e;
  B200_CHECK_ARG(n > 0 && n % 2048 == 0 && offset >= 0 && offset % 2048 == 0,
                 "allreduce_bf16: offset and n must be multiples of 2048 elements");
  const int64_t nblk = n / 2048;
  if (nblk < ctas) ctas = (int)nblk;
  allreduce_bf16_kernel<<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, nblk, scale);
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
  broadcast_f32_kernel<<<ctas, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d, offset, n / 4, root);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
