// common.cuh — shared device helpers for the sm_100a kernel library.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>

#include "b200cnn.h"

namespace b200 {

void set_error(const char* fmt, ...);

#define B200_CHECK_ARG(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      ::b200::set_error(__VA_ARGS__);    \
      return -1;                         \
    }                                    \
  } while (0)

#define B200_CUDA(expr)                                                              \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return static_cast<int>(_e);                                                   \
    }                                                                                \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA(cudaGetLastError())

constexpr int kNumSMs = 148;

// Kernel function attributes (the opt-in to more than 48 KiB of dynamic shared memory) are per DEVICE, so a "set it
// once" flag has to be kept per device too: true the first time this is called for `seen` with the current device.
inline bool first_use_on_device(std::atomic<uint64_t>& seen) {
  int d = 0;
  cudaGetDevice(&d);
  const uint64_t bit = 1ull << (d & 63);
  return (seen.fetch_or(bit) & bit) == 0;
}

// Dynamic shared memory the persistent tensor-core kernels may take: 223 of the SM's 228 KiB. The rest (and the
// registers a 576-thread CTA leaves free) is the room in which a 128-thread communication CTA (csrc/comm) becomes
// resident NEXT TO a convolution CTA, so gradient collectives progress under a running convolution.
constexpr int kMaxConvSmem = 223 * 1024;

// 8 bf16 values = one 16-byte vector.
struct alignas(16) bf16x8 {
  __nv_bfloat162 v[4];
};

__device__ __forceinline__ bf16x8 ld8(const b200_bf16* p) {
  bf16x8 r;
  *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(p);
  return r;
}

// Streaming load: data that is read once (do not keep in L1).
__device__ __forceinline__ bf16x8 ld8_stream(const b200_bf16* p) {
  bf16x8 r;
  uint4 u;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
               : "l"(p));
  *reinterpret_cast<uint4*>(&r) = u;
  return r;
}

__device__ __forceinline__ void st8(b200_bf16* p, const bf16x8& v) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&v);
}

__device__ __forceinline__ void unpack8(const bf16x8& v, float f[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(v.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

__device__ __forceinline__ bf16x8 pack8(const float f[8]) {
  bf16x8 r;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

inline int grid_for(int64_t work_items, int threads, int max_blocks_per_sm = 8) {
  int64_t blocks = (work_items + threads - 1) / threads;
  int64_t cap = static_cast<int64_t>(kNumSMs) * max_blocks_per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace b200
