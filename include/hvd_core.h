/*
 * hvd_core.h — host-side runtime of the B200-native Horovod-style data-parallel layer.
 *
 * Plays the role Horovod's C++ core plays for the reference harness (tensor fusion with
 * HOROVOD_FUSION_THRESHOLD=134217728, reference benchmark-scripts/run-tf-sing-ucx-openmpi.sh:105;
 * background coordinator + timeline + stall check are external to the reference tree, see
 * SURVEY.md §2.3/§2.4). The design is different on purpose: a CNN's gradient production order is
 * static, so there is no per-cycle negotiation. The plan is computed once, hashed, and the hash is
 * compared across ranks at the first step.
 *
 * Pure C ABI (ctypes-friendly). No CUDA dependency: loads and runs on a CPU-only box.
 */
#ifndef B200_HVD_CORE_H_
#define B200_HVD_CORE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- fusion planner */

/* One tensor handed to the planner, in gradient-readiness order (reverse of forward order). */
typedef struct hvd_tensor_desc {
  int64_t numel;      /* element count */
  int32_t elem_size;  /* bytes per element in the fusion buffer (4 = fp32, 2 = bf16) */
  int32_t dtype_id;   /* tensors of different dtype_id never share a bucket */
  int32_t flags;      /* bit0: apply weight decay, bit1: LARS-adapt this tensor */
  int32_t reserved;
} hvd_tensor_desc;

/* One contiguous piece of a tensor placed in a bucket. Oversize tensors become several. */
typedef struct hvd_segment {
  int32_t tensor;        /* index into the input array */
  int32_t bucket;        /* bucket index */
  int64_t tensor_offset; /* first element of the tensor covered by this segment */
  int64_t bucket_offset; /* element offset inside the bucket */
  int64_t numel;         /* elements in this segment */
} hvd_segment;

typedef struct hvd_bucket {
  int64_t numel;        /* padded element count (multiple of align_elems) */
  int32_t elem_size;
  int32_t dtype_id;
  int32_t first_segment;
  int32_t num_segments;
} hvd_bucket;

typedef struct hvd_plan hvd_plan;

/* Build a static plan. threshold_bytes<=0 selects the reference's 128 MiB. align_elems is the
 * padding granularity (elements) of every segment start and bucket size; it must be a power of
 * two. A tensor larger than the threshold is split into threshold-sized chunks (VGG-16 fc6). */
hvd_plan* hvd_plan_create(const hvd_tensor_desc* tensors, int32_t n, int64_t threshold_bytes,
                          int32_t align_elems);
/* The same, with the tensors that make up the last `tail_bytes` of the gradient stream cut off into a bucket of their
 * own: the final bucket's collective is the only one that cannot overlap with backward, so it is kept small. */
hvd_plan* hvd_plan_create_tail(const hvd_tensor_desc* tensors, int32_t n, int64_t threshold_bytes,
                               int32_t align_elems, int64_t tail_bytes);
void hvd_plan_destroy(hvd_plan* p);
int32_t hvd_plan_num_buckets(const hvd_plan* p);
int32_t hvd_plan_num_segments(const hvd_plan* p);
int32_t hvd_plan_get_bucket(const hvd_plan* p, int32_t i, hvd_bucket* out);
int32_t hvd_plan_get_segment(const hvd_plan* p, int32_t i, hvd_segment* out);
/* FNV-1a hash over the whole plan; ranks compare it once instead of negotiating every cycle. */
uint64_t hvd_plan_hash(const hvd_plan* p);

/* ---------------------------------------------------------------- readiness tracker */

typedef struct hvd_tracker hvd_tracker;

/* Counts down the segments of each bucket as their gradients materialise during backward. */
hvd_tracker* hvd_tracker_create(const hvd_plan* p);
void hvd_tracker_destroy(hvd_tracker* t);
/* Start a new step: all counters are re-armed. */
void hvd_tracker_begin_step(hvd_tracker* t);
/* Mark tensor ready. Writes the indices of buckets that just became complete into out_buckets
 * (capacity cap) and returns how many. Returns -1 on a double-ready or out-of-range tensor. */
int32_t hvd_tracker_tensor_ready(hvd_tracker* t, int32_t tensor, int32_t* out_buckets, int32_t cap);
/* Buckets not yet complete this step. */
int32_t hvd_tracker_pending(const hvd_tracker* t);
/* Mark a bucket's collective as finished (for the stall inspector). */
void hvd_tracker_bucket_done(hvd_tracker* t, int32_t bucket);
/* Seconds since the oldest launched-but-unfinished bucket was launched, or 0. */
double hvd_tracker_oldest_inflight_seconds(const hvd_tracker* t);

/* ---------------------------------------------------------------- timeline (Chrome trace) */

typedef struct hvd_timeline hvd_timeline;

hvd_timeline* hvd_timeline_open(const char* path, int32_t rank);
/* Complete event ("ph":"X") with explicit start and duration in microseconds. */
void hvd_timeline_event(hvd_timeline* tl, const char* name, const char* phase, int32_t lane,
                        double start_us, double dur_us);
/* Instant marker. */
void hvd_timeline_mark(hvd_timeline* tl, const char* name, int32_t lane, double ts_us);
double hvd_timeline_now_us(const hvd_timeline* tl);
void hvd_timeline_close(hvd_timeline* tl);

/* ---------------------------------------------------------------- stall inspector */

typedef struct hvd_stall hvd_stall;
/* Background thread: every period_s it checks the tracker; when a bucket has been in flight for
 * more than warn_s it prints a warning on stderr, and after abort_s (if > 0) it calls abort(). */
hvd_stall* hvd_stall_start(hvd_tracker* t, double period_s, double warn_s, double abort_s);
int32_t hvd_stall_warnings(const hvd_stall* s);
void hvd_stall_stop(hvd_stall* s);

/* ---------------------------------------------------------------- launcher arithmetic */

/* The worker arithmetic of the reference launcher (run-tf-sing-ucx-openmpi.sh:40-50).
 * out[0]=workers_per_node out[1]=cores_per_worker out[2]=intra_threads out[3]=total_workers.
 * Returns 0, or -1 for invalid input. */
int32_t hvd_launch_arith(int32_t num_nodes, int32_t workers_per_socket, int32_t num_sockets,
                         int32_t cores_per_socket, int32_t inter_threads, int32_t out[4]);

const char* hvd_core_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_HVD_CORE_H_ */
