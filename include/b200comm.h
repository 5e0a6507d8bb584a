/*
 * b200comm.h — C ABI of the NVLink/NVSwitch communication layer in libb200cnn.so.
 *
 * Role in the reference stack: the MPI transport under Horovod — OpenMPI 4.0.0 + UCX 1.5.0 (`rc_x,sm,self`, hcoll)
 * or Intel MPI 2019.4 + libfabric (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:85-95,
 * run-tf-sing-libfabric-intelmpi.sh:86-101) — plus Horovod's own MPI_Allreduce call and TF's ApplyMomentum that
 * follows it. Here every rank maps every peer's buffers (CUDA IPC) and one sm_100a kernel
 * reads and writes them directly over NVLink: reduce-scatter by peer loads, all-gather by peer stores, the 1/N
 * scale and the momentum-SGD update, with flag-based cross-GPU barriers. No NCCL call is involved. Three algorithms,
 * chosen per message size: one-shot (every rank reads all peers; small messages), two-shot (owner reduces its shard
 * with peer loads and all-gathers it with peer stores) and NVLS (multimem.ld_reduce pulls the shard already reduced
 * inside the NVSwitch, multimem.st broadcasts it; needs a multicast mapping of the buffer).
 *
 * Symmetric buffers: each rank allocates the same size with b200_symm_alloc, exports an IPC handle, and opens the
 * handles of its peers (exchange them with any out-of-band channel, e.g. torch.distributed all_gather).
 */
#ifndef B200COMM_H_
#define B200COMM_H_

#include <stdint.h>

#include "b200cnn.h"

#ifdef __cplusplus
extern "C" {
#endif

#define B200_COMM_MAX_RANKS 8
#define B200_COMM_MAX_CTAS 128
#define B200_IPC_HANDLE_BYTES 64

/* Device allocation suitable for IPC export (zero-filled). These are the only allocating calls of this header. */
int b200_symm_alloc(int64_t bytes, void** dptr);
int b200_symm_free(void* dptr);
int b200_symm_get_handle(void* dptr, void* handle_out /* 64 bytes */);
int b200_symm_open_handle(const void* handle /* 64 bytes */, void** peer_ptr);
int b200_symm_close_handle(void* peer_ptr);
/* Symmetric memory on the virtual-memory-management API (csrc/comm/symm_alloc.cpp): the allocation that can carry an
 * NVLS multicast mapping. Collective protocol, driven by the caller (parallel/symm.py does it with these calls and a
 * barrier of its choice): every rank b200_vmm_create()s the same size and gets a POSIX file descriptor of its physical
 * allocation; descriptors travel between processes over abstract Unix-domain sockets (b200_fd_listen / _send / _recv,
 * SCM_RIGHTS); every rank b200_vmm_map_peer()s each peer's descriptor; for multicast rank 0 b200_vmm_mc_create()s and
 * sends that descriptor, the others b200_vmm_mc_import() it, ALL b200_vmm_mc_add_device(), barrier, ALL
 * b200_vmm_mc_bind(). b200_vmm_ptr(h, r) is rank r's buffer as mapped here, b200_vmm_mc_ptr(h) the multicast address.
 * The reference's counterpart is MPI/UCX endpoint set-up (run-tf-sing-ucx-openmpi.sh:85-95). */
int b200_vmm_supported(int32_t device, int32_t* multicast_ok);
int b200_vmm_create(int64_t bytes, int32_t world, int32_t rank, int32_t want_multicast, void** handle_out, int32_t* fd_out,
                    int64_t* padded_bytes);
int b200_vmm_map_peer(void* handle, int32_t peer, int32_t fd);
uint64_t b200_vmm_ptr(void* handle, int32_t rank);
int b200_vmm_mc_create(void* handle, int32_t* fd_out);
int b200_vmm_mc_import(void* handle, int32_t fd);
int b200_vmm_mc_add_device(void* handle);
int b200_vmm_mc_bind(void* handle);
uint64_t b200_vmm_mc_ptr(void* handle);
int b200_vmm_free(void* handle);
int b200_fd_listen(const char* name);
int b200_fd_send(int sock, const char* peer_name, int fd, int32_t tag);
int b200_fd_recv(int sock, int32_t* tag_out);
int b200_fd_close(int fd);

/* Bytes of the per-rank signal pad (flags + epoch counters) the kernels need. */
int64_t b200_comm_pad_bytes(void);
/* Elements of one fp32 work block of the collectives (512): ownership in the two-shot algorithms alternates between
 * ranks at this granularity, and the fused-LARS block tables below are in these units. */
int32_t b200_comm_block_elems(void);

typedef struct b200_comm {
  uint64_t buf[B200_COMM_MAX_RANKS];   /* device addresses of every rank's data buffer (own rank included) */
  uint64_t pad[B200_COMM_MAX_RANKS];   /* device addresses of every rank's signal pad */
  int32_t rank, world;
  int32_t ctas;                        /* CTAs per collective (<= B200_COMM_MAX_CTAS); the same on all ranks */
  int32_t one_shot_max_bytes;          /* messages up to this size use the one-shot algorithm */
  uint64_t mc;                         /* multicast (NVLS) address of the data buffer, or 0 when there is none */
  int64_t nvls_min_bytes;              /* messages of at least this size use the in-switch reduction when mc != 0;
                                          negative = never */
  int32_t algo;                        /* B200_COMM_ALGO_*: AUTO picks by the two thresholds above; the others force
                                          one algorithm (the sweep tool measures each to derive the thresholds) */
  int32_t reserved;
} b200_comm;

enum { B200_COMM_ALGO_AUTO = 0, B200_COMM_ALGO_ONE_SHOT = 1, B200_COMM_ALGO_TWO_SHOT = 2, B200_COMM_ALGO_NVLS = 3 };

/* Optimizer half of the fused kernel; all pointers are LOCAL flat buffers indexed like the gradient buffer. */
typedef struct b200_sgd_args {
  float* master;
  float* mom;
  b200_bf16* shadow;                   /* may be NULL */
  const int32_t* block_tensor;         /* per 1024 elements: owning tensor or -1; may be NULL */
  const float* tensor_decay;
  const float* tensor_lars;
  float* trust;                        /* per-tensor LARS trust ratios; may be NULL (plain momentum SGD). Read only,
                                          unless the fused LARS fields below are set: then the kernel writes it */
  float lr, mu, wd;
  const float* lr_dev;                 /* when not NULL the learning rate is read from this device float at run time
                                          instead of `lr`: a schedule can then change it under a captured CUDA graph */
  /* Fused LARS (all NULL/0 = off). "Block" = b200_comm_block_elems() elements. The kernel takes per-block partial
   * norms (|w|^2, |g|^2; lars_block_partials holds 2 floats per block of the WHOLE flat buffer) of the reduced bucket, folds them per tensor in a fixed order for the tensors
   * [lars_t_lo, lars_t_hi) that END in this bucket (tensor t owns blocks [tensor_blk_lo[t], +tensor_blk_cnt[t])),
   * writes trust[t] = eta*|w| / (|g| + wd*decay*|w| + eps), and updates the blocks [lars_upd_lo, lars_upd_hi) those
   * tensors span - a range that reaches back into the previous bucket when a tensor was split across buckets, whose
   * update was deferred for that reason. Buckets must be reduced in plan order on one stream. */
  float* lars_block_partials;
  const int32_t* tensor_blk_lo;
  const int32_t* tensor_blk_cnt;
  int32_t lars_t_lo, lars_t_hi;
  int64_t lars_upd_lo, lars_upd_hi;    /* in blocks of the flat buffer */
  float lars_eta, lars_eps;
} b200_sgd_args;

/* In-place sum over ranks of fp32 elements [offset, offset+n) of the symmetric buffer, times `scale`
 * (n and offset multiples of 1024 elements; within the one-shot range multiples of 4 suffice). Every rank ends with
 * bitwise identical data: each element is reduced by exactly one owner in rank order (two-shot), or by every rank in the same order (one-shot). With sgd != NULL the
 * same launch then applies the momentum-SGD update (b200_sgd_momentum_flat semantics, grad_scale = 1) to
 * [offset, offset+n) of the local optimizer buffers. All ranks must make the same call in the same order. */
int b200_allreduce_f32(const b200_comm* comm, int64_t offset, int64_t n, float scale, const b200_sgd_args* sgd,
                       void* stream);

/* The same for bf16 payloads (n % 2048 == 0): accumulation in fp32, one rounding at the end. No optimizer half. */
int b200_allreduce_bf16(const b200_comm* comm, int64_t offset, int64_t n, float scale, void* stream);

/* Cross-GPU barrier on `stream` (all ranks). */
int b200_comm_barrier(const b200_comm* comm, void* stream);

/* Root writes elements [offset, offset+n) of its buffer into every peer's buffer (n % 4 == 0 floats): one
 * multimem.st per vector when the buffer has a multicast mapping, else one store per peer; ends with a barrier.
 * Horovod's broadcast_global_variables role (K14; osu_bcast). */
int b200_broadcast_f32(const b200_comm* comm, int64_t offset, int64_t n, int32_t root, void* stream);

/* All-gather in place: rank r's segment [offset + r*n_per_rank, +n_per_rank) of its own buffer is written to the same
 * position of every peer's buffer (multicast store when available); barriers on both sides. n_per_rank % 4 == 0.
 * MPI_Allgather role (osu_allgather; Horovod's allgather op). */
int b200_allgather_f32(const b200_comm* comm, int64_t offset, int64_t n_per_rank, void* stream);

/* Point-to-point probes (osu_bw / osu_bibw / osu_latency roles; install-scripts/install_osu_bench.sh:13-17).
 * b200_p2p_copy: this rank moves n floats at `offset` to (pull == 0: peer stores) or from (pull != 0: peer loads) ONE
 * peer's buffer; no synchronisation inside - bracket it with events; call it on one rank for one direction or on both
 * ranks of a pair for the bidirectional figure.
 * b200_p2p_pingpong: ranks a and b bounce an n-float message `iters` times (flag + payload over NVLink, one CTA
 * each); every rank calls it, the others return at once; rank a writes the mean one-way time in microseconds,
 * measured on the device with %globaltimer, to out_us[0]. n <= 4 Mi floats. */
int b200_p2p_copy(const b200_comm* comm, int32_t peer, int64_t offset, int64_t n, int32_t pull, void* stream);
int b200_p2p_pingpong(const b200_comm* comm, int32_t rank_a, int32_t rank_b, int64_t offset, int64_t n, int32_t iters,
                      float* out_us, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200COMM_H_ */
