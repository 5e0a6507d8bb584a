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
/* Bytes of the per-rank signal pad (flags + epoch counters) the kernels need. */
int64_t b200_comm_pad_bytes(void);

typedef struct b200_comm {
  uint64_t buf[B200_COMM_MAX_RANKS];   /* device addresses of every rank's data buffer (own rank included) */
  uint64_t pad[B200_COMM_MAX_RANKS];   /* device addresses of every rank's signal pad */
  int32_t rank, world;
  int32_t ctas;                        /* CTAs per collective (<= B200_COMM_MAX_CTAS); the same on all ranks */
  int32_t one_shot_max_bytes;          /* messages up to this size use the one-shot algorithm */
  uint64_t mc;                         /* multicast (NVLS) address of the data buffer, or 0 when there is none */
  int64_t nvls_min_bytes;              /* messages of at least this size use the in-switch reduction when mc != 0
                                          and world > 2 (with two ranks a direct peer read is the shorter path) */
} b200_comm;

/* Optimizer half of the fused kernel; all pointers are LOCAL flat buffers indexed like the gradient buffer. */
typedef struct b200_sgd_args {
  float* master;
  float* mom;
  b200_bf16* shadow;                   /* may be NULL */
  const int32_t* block_tensor;         /* per 1024 elements: owning tensor or -1; may be NULL */
  const float* tensor_decay;
  const float* tensor_lars;
  const float* trust;                  /* may be NULL */
  float lr, mu, wd;
} b200_sgd_args;

/* In-place sum over ranks of fp32 elements [offset, offset+n) of the symmetric buffer, times `scale`
 * (n % 1024 == 0, offset % 1024 == 0). Every rank ends with bitwise identical data: each element is reduced by
 * exactly one owner in rank order (two-shot), or by every rank in the same order (one-shot). With sgd != NULL the
 * same launch then applies the momentum-SGD update (b200_sgd_momentum_flat semantics, grad_scale = 1) to
 * [offset, offset+n) of the local optimizer buffers. All ranks must make the same call in the same order. */
int b200_allreduce_f32(const b200_comm* comm, int64_t offset, int64_t n, float scale, const b200_sgd_args* sgd,
                       void* stream);

/* The same for bf16 payloads (n % 2048 == 0): accumulation in fp32, one rounding at the end. No optimizer half. */
int b200_allreduce_bf16(const b200_comm* comm, int64_t offset, int64_t n, float scale, void* stream);

/* Cross-GPU barrier on `stream` (all ranks). */
int b200_comm_barrier(const b200_comm* comm, void* stream);

/* Root writes elements [offset, offset+n) of its buffer into every peer's buffer (n % 4 == 0 floats); ends with a
 * barrier. Horovod's broadcast_global_variables role (K14). */
int b200_broadcast_f32(const b200_comm* comm, int64_t offset, int64_t n, int32_t root, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200COMM_H_ */
