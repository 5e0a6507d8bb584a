// hvd_core.cpp — host runtime: static fusion planner, readiness tracker, timeline, stall inspector.
// See include/hvd_core.h for the contract and the reference citations.
#include "hvd_core.h"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

using Clock = std::chrono::steady_clock;

inline int64_t round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

inline double seconds_since(Clock::time_point t0) {
  return std::chrono::duration<double>(Clock::now() - t0).count();
}

}  // namespace

struct hvd_plan {
  std::vector<hvd_bucket> buckets;
  std::vector<hvd_segment> segments;
  int32_t num_tensors = 0;
  std::vector<int32_t> segments_per_tensor;
};

extern "C" {

hvd_plan* hvd_plan_create(const hvd_tensor_desc* tensors, int32_t n, int64_t threshold_bytes,
                          int32_t align_elems) {
  return hvd_plan_create_tail(tensors, n, threshold_bytes, align_elems, 0);
}

hvd_plan* hvd_plan_create_tail(const hvd_tensor_desc* tensors, int32_t n, int64_t threshold_bytes,
                               int32_t align_elems, int64_t tail_bytes) {
  if (n < 0 || (n > 0 && tensors == nullptr)) return nullptr;
  if (align_elems <= 0 || (align_elems & (align_elems - 1)) != 0) return nullptr;
  if (threshold_bytes <= 0) threshold_bytes = 134217728;  // reference default, 128 MiB
  auto* p = new hvd_plan();
  p->num_tensors = n;
  p->segments_per_tensor.assign(n, 0);

  // The last bucket's collective cannot overlap with anything (backward is over when its final gradient lands), so
  // its size is pure exposed time: with tail_bytes > 0 the tensors that make up the last tail_bytes of the gradient
  // stream are cut off into a bucket of their own, however much room the open bucket still has.
  int64_t total_bytes = 0, seen_bytes = 0;
  for (int32_t t = 0; t < n; ++t) total_bytes += tensors[t].numel * tensors[t].elem_size;
  bool tail_cut = tail_bytes <= 0;

  int32_t open = -1;  // index of the bucket currently being filled
  auto open_bucket = [&](int32_t elem_size, int32_t dtype_id) {
    hvd_bucket b{};
    b.numel = 0;
    b.elem_size = elem_size;
    b.dtype_id = dtype_id;
    b.first_segment = static_cast<int32_t>(p->segments.size());
    b.num_segments = 0;
    p->buckets.push_back(b);
    open = static_cast<int32_t>(p->buckets.size()) - 1;
  };

  for (int32_t t = 0; t < n; ++t) {
    const hvd_tensor_desc& d = tensors[t];
    if (d.numel <= 0 || d.elem_size <= 0) {
      delete p;
      return nullptr;
    }
    // Largest element count per bucket for this dtype, a multiple of the alignment.
    int64_t cap = threshold_bytes / d.elem_size / align_elems * align_elems;
    if (cap < align_elems) cap = align_elems;
    int64_t done = 0;
    bool cut_here = false;
    if (!tail_cut && open >= 0 && p->buckets[open].numel > 0 && total_bytes - seen_bytes <= tail_bytes) {
      cut_here = true;
      tail_cut = true;
    }
    seen_bytes += d.numel * d.elem_size;
    while (done < d.numel) {
      int64_t remaining = d.numel - done;
      bool need_new = open < 0 || cut_here || p->buckets[open].dtype_id != d.dtype_id ||
                      p->buckets[open].elem_size != d.elem_size;
      cut_here = false;
      if (!need_new) {
        int64_t room = cap - p->buckets[open].numel;
        // A tensor that fits a fresh bucket is never split just to top up the open one
        // (keeps per-tensor segments to one in the common case; Horovod closes the buffer too).
        if (room <= 0 || (remaining > room && remaining <= cap)) need_new = true;
      }
      if (need_new) open_bucket(d.elem_size, d.dtype_id);
      hvd_bucket& b = p->buckets[open];
      int64_t room = cap - b.numel;
      int64_t take = remaining < room ? remaining : room;
      hvd_segment s{};
      s.tensor = t;
      s.bucket = open;
      s.tensor_offset = done;
      s.bucket_offset = b.numel;
      s.numel = take;
      p->segments.push_back(s);
      p->segments_per_tensor[t] += 1;
      b.num_segments += 1;
      b.numel = round_up(b.numel + take, align_elems);
      done += take;
    }
  }
  return p;
}

void hvd_plan_destroy(hvd_plan* p) { delete p; }

int32_t hvd_plan_num_buckets(const hvd_plan* p) {
  return p ? static_cast<int32_t>(p->buckets.size()) : -1;
}

int32_t hvd_plan_num_segments(const hvd_plan* p) {
  return p ? static_cast<int32_t>(p->segments.size()) : -1;
}

int32_t hvd_plan_get_bucket(const hvd_plan* p, int32_t i, hvd_bucket* out) {
  if (!p || !out || i < 0 || i >= static_cast<int32_t>(p->buckets.size())) return -1;
  *out = p->buckets[i];
  return 0;
}

int32_t hvd_plan_get_segment(const hvd_plan* p, int32_t i, hvd_segment* out) {
  if (!p || !out || i < 0 || i >= static_cast<int32_t>(p->segments.size())) return -1;
  *out = p->segments[i];
  return 0;
}

uint64_t hvd_plan_hash(const hvd_plan* p) {
  if (!p) return 0;
  uint64_t h = 1469598103934665603ull;
  auto mix = [&h](const void* data, size_t len) {
    const unsigned char* c = static_cast<const unsigned char*>(data);
    for (size_t i = 0; i < len; ++i) {
      h ^= c[i];
      h *= 1099511628211ull;
    }
  };
  for (const hvd_bucket& b : p->buckets) {
    mix(&b.numel, sizeof b.numel);
    mix(&b.elem_size, sizeof b.elem_size);
    mix(&b.dtype_id, sizeof b.dtype_id);
    mix(&b.num_segments, sizeof b.num_segments);
  }
  for (const hvd_segment& s : p->segments) {
    mix(&s.tensor, sizeof s.tensor);
    mix(&s.bucket, sizeof s.bucket);
    mix(&s.tensor_offset, sizeof s.tensor_offset);
    mix(&s.bucket_offset, sizeof s.bucket_offset);
    mix(&s.numel, sizeof s.numel);
  }
  return h;
}

}  // extern "C"

// ------------------------------------------------------------------------------- tracker

struct hvd_tracker {
  const hvd_plan* plan;
  mutable std::mutex mu;
  std::vector<int32_t> remaining;       // per bucket: segments whose tensor is not ready yet
  std::vector<uint8_t> tensor_ready;    // per tensor
  std::vector<uint8_t> inflight;        // per bucket: launched, not yet done
  std::vector<Clock::time_point> launched_at;
  int32_t pending = 0;
};

extern "C" {

hvd_tracker* hvd_tracker_create(const hvd_plan* p) {
  if (!p) return nullptr;
  auto* t = new hvd_tracker();
  t->plan = p;
  t->remaining.resize(p->buckets.size());
  t->tensor_ready.resize(p->num_tensors);
  t->inflight.assign(p->buckets.size(), 0);
  t->launched_at.resize(p->buckets.size());
  hvd_tracker_begin_step(t);
  return t;
}

void hvd_tracker_destroy(hvd_tracker* t) { delete t; }

void hvd_tracker_begin_step(hvd_tracker* t) {
  if (!t) return;
  std::lock_guard<std::mutex> lk(t->mu);
  for (size_t b = 0; b < t->remaining.size(); ++b) t->remaining[b] = t->plan->buckets[b].num_segments;
  std::fill(t->tensor_ready.begin(), t->tensor_ready.end(), 0);
  t->pending = static_cast<int32_t>(t->remaining.size());
}

int32_t hvd_tracker_tensor_ready(hvd_tracker* t, int32_t tensor, int32_t* out_buckets, int32_t cap) {
  if (!t) return -1;
  std::lock_guard<std::mutex> lk(t->mu);
  if (tensor < 0 || tensor >= t->plan->num_tensors) return -1;
  if (t->tensor_ready[tensor]) return -1;
  t->tensor_ready[tensor] = 1;
  int32_t n = 0;
  for (const hvd_segment& s : t->plan->segments) {
    if (s.tensor != tensor) continue;
    if (--t->remaining[s.bucket] == 0) {
      t->pending -= 1;
      t->inflight[s.bucket] = 1;
      t->launched_at[s.bucket] = Clock::now();
      if (out_buckets && n < cap) out_buckets[n] = s.bucket;
      n += 1;
    }
  }
  return n;
}

int32_t hvd_tracker_pending(const hvd_tracker* t) {
  if (!t) return -1;
  std::lock_guard<std::mutex> lk(t->mu);
  return t->pending;
}

void hvd_tracker_bucket_done(hvd_tracker* t, int32_t bucket) {
  if (!t) return;
  std::lock_guard<std::mutex> lk(t->mu);
  if (bucket >= 0 && bucket < static_cast<int32_t>(t->inflight.size())) t->inflight[bucket] = 0;
}

double hvd_tracker_oldest_inflight_seconds(const hvd_tracker* t) {
  if (!t) return 0.0;
  std::lock_guard<std::mutex> lk(t->mu);
  double worst = 0.0;
  for (size_t b = 0; b < t->inflight.size(); ++b) {
    if (!t->inflight[b]) continue;
    double s = seconds_since(t->launched_at[b]);
    if (s > worst) worst = s;
  }
  return worst;
}

}  // extern "C"

// ------------------------------------------------------------------------------- timeline

struct hvd_timeline {
  FILE* f = nullptr;
  int32_t rank = 0;
  bool first = true;
  Clock::time_point t0;
  std::mutex mu;
};

namespace {
void json_escape(const char* s, std::string* out) {
  for (; s && *s; ++s) {
    if (*s == '"' || *s == '\\') out->push_back('\\');
    if (static_cast<unsigned char>(*s) >= 0x20) out->push_back(*s);
  }
}
}  // namespace

extern "C" {

hvd_timeline* hvd_timeline_open(const char* path, int32_t rank) {
  if (!path) return nullptr;
  FILE* f = std::fopen(path, "w");
  if (!f) return nullptr;
  auto* tl = new hvd_timeline();
  tl->f = f;
  tl->rank = rank;
  tl->t0 = Clock::now();
  std::fputs("[\n", f);
  return tl;
}

double hvd_timeline_now_us(const hvd_timeline* tl) {
  if (!tl) return 0.0;
  return seconds_since(tl->t0) * 1e6;
}

void hvd_timeline_event(hvd_timeline* tl, const char* name, const char* phase, int32_t lane,
                        double start_us, double dur_us) {
  if (!tl) return;
  std::string n, c;
  json_escape(name, &n);
  json_escape(phase, &c);
  std::lock_guard<std::mutex> lk(tl->mu);
  std::fprintf(tl->f, "%s{\"name\":\"%s\",\"cat\":\"%s\",\"ph\":\"X\",\"pid\":%d,\"tid\":%d,\"ts\":%.3f,\"dur\":%.3f}",
               tl->first ? "" : ",\n", n.c_str(), c.c_str(), tl->rank, lane, start_us, dur_us);
  tl->first = false;
}

void hvd_timeline_mark(hvd_timeline* tl, const char* name, int32_t lane, double ts_us) {
  if (!tl) return;
  std::string n;
  json_escape(name, &n);
  std::lock_guard<std::mutex> lk(tl->mu);
  std::fprintf(tl->f, "%s{\"name\":\"%s\",\"ph\":\"i\",\"s\":\"t\",\"pid\":%d,\"tid\":%d,\"ts\":%.3f}",
               tl->first ? "" : ",\n", n.c_str(), tl->rank, lane, ts_us);
  tl->first = false;
}

void hvd_timeline_close(hvd_timeline* tl) {
  if (!tl) return;
  std::fputs("\n]\n", tl->f);
  std::fclose(tl->f);
  delete tl;
}

}  // extern "C"

// ------------------------------------------------------------------------------- stall inspector

struct hvd_stall {
  hvd_tracker* tracker;
  double period_s, warn_s, abort_s;
  std::atomic<int32_t> warnings{0};
  std::atomic<bool> stop{false};
  std::mutex mu;
  std::condition_variable cv;
  std::thread th;
};

extern "C" {

hvd_stall* hvd_stall_start(hvd_tracker* t, double period_s, double warn_s, double abort_s) {
  if (!t || period_s <= 0) return nullptr;
  auto* s = new hvd_stall();
  s->tracker = t;
  s->period_s = period_s;
  s->warn_s = warn_s;
  s->abort_s = abort_s;
  s->th = std::thread([s] {
    std::unique_lock<std::mutex> lk(s->mu);
    while (!s->stop.load()) {
      s->cv.wait_for(lk, std::chrono::duration<double>(s->period_s));
      if (s->stop.load()) break;
      double age = hvd_tracker_oldest_inflight_seconds(s->tracker);
      if (s->warn_s > 0 && age > s->warn_s) {
        s->warnings.fetch_add(1);
        std::fprintf(stderr,
                     "[hvd stall] a fused gradient bucket has been in flight for %.1f s; one or more "
                     "ranks may not have reached the collective\n", age);
      }
      if (s->abort_s > 0 && age > s->abort_s) {
        std::fprintf(stderr, "[hvd stall] aborting after %.1f s\n", age);
        std::abort();
      }
    }
  });
  return s;
}

int32_t hvd_stall_warnings(const hvd_stall* s) { return s ? s->warnings.load() : 0; }

void hvd_stall_stop(hvd_stall* s) {
  if (!s) return;
  {
    std::lock_guard<std::mutex> lk(s->mu);
    s->stop.store(true);
  }
  s->cv.notify_all();
  if (s->th.joinable()) s->th.join();
  delete s;
}

int32_t hvd_launch_arith(int32_t num_nodes, int32_t workers_per_socket, int32_t num_sockets,
                         int32_t cores_per_socket, int32_t inter_threads, int32_t out[4]) {
  if (!out || num_nodes <= 0 || workers_per_socket < 0 || num_sockets <= 0 || cores_per_socket <= 0 ||
      inter_threads <= 0)
    return -1;
  int32_t cores_per_worker, workers_per_node;
  if (workers_per_socket == 0) {
    cores_per_worker = cores_per_socket * num_sockets;
    workers_per_node = 1;
  } else {
    cores_per_worker = cores_per_socket / workers_per_socket;
    workers_per_node = workers_per_socket * num_sockets;
  }
  out[0] = workers_per_node;
  out[1] = cores_per_worker;
  out[2] = cores_per_worker / inter_threads;
  out[3] = num_nodes * workers_per_node;
  return 0;
}

const char* hvd_core_version(void) { return "b200-hvd-core 0.1"; }

}  // extern "C"
