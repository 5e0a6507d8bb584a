// mma_rate.cu - cycles per tcgen05.mma (bf16, M=128, K=16, SS operands, 128B swizzle) as a function of N and operand
// majorness, with nothing else touching shared memory. Measured on B200 (148 CTAs, one per SM):
//   N=256: 128.0 cycles = the tensor-pipe floor, in every mode.
//   N<=128: 96 cycles with descriptors rebuilt per MMA, 78 with stepped descriptor words, 99.5 when a tcgen05.commit
//   follows every fourth MMA - i.e. each tcgen05 instruction costs the one issuing thread ~80 cycles, more than the
//   32 / 64 cycles of math of a 64 / 128-column MMA. Hence the second issuing warp of the halo mode in conv_gemm.cu. Build: nvcc -gencode arch=compute_100a,code=sm_100a -I../../azure_hc_intel_tf_b200/csrc -I../../include
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "conv/umma.cuh"
using namespace umma;

__global__ void __launch_bounds__(128, 1) rate_kernel(int n, int a_mn, int b_mn, int iters, int kblocks, int fast, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint64_t bar2;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&bar2, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0 && lane == 0) {
    const uint32_t idesc = make_idesc_bf16(128, (uint32_t)n, (uint32_t)a_mn, (uint32_t)b_mn);
    const int stage_bytes = 16384 + n * 128;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t sa = smem_u32(smem) + (uint32_t)((it % kblocks) * stage_bytes);
      const uint32_t sb = sa + 16384;
      if (fast) {
        // descriptor low words stepped with adds (what the production issue loops do)
        const uint32_t a_lo = a_mn ? sdesc_lo(sa, 8192) : sdesc_lo(sa, 16);
        const uint32_t b_lo = b_mn ? sdesc_lo(sb, 8192) : sdesc_lo(sb, 16);
        const uint32_t a_step = a_mn ? 128u : 2u, b_step = b_mn ? 128u : 2u, hi = sdesc_hi_sw128(1024);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_bf16_lohi(tmem, a_lo + k * a_step, hi, b_lo + k * b_step, hi, idesc, (it | k) ? 1u : 0u);
        if (fast == 2) mma_commit(&bar2);      // a tcgen05.commit per k-block, as the pipelined kernels do
        continue;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = a_mn ? make_sdesc_sw128(sa + k * 2048, 8192, 1024) : make_sdesc_sw128(sa + k * 32, 16, 1024);
        const uint64_t db = b_mn ? make_sdesc_sw128(sb + k * 2048, 8192, 1024) : make_sdesc_sw128(sb + k * 32, 16, 1024);
        mma_bf16(tmem, da, db, idesc, (it | k) ? 1u : 0u);
      }
    }
    mma_commit(&bar);
    mbar_wait(&bar, 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  const int iters = 4000;
  for (int fast : {0, 1, 2})
    for (int n : {64, 128, 256})
      for (int mode = 0; mode < 4; mode += 3) {
        const int grid = 148;
        const int a_mn = mode & 1, b_mn = mode >> 1;
        const int kblocks = (200 * 1024) / (16384 + n * 128);
        rate_kernel<<<grid, 128, 210 * 1024>>>(n, a_mn, b_mn, iters, kblocks, fast, d);
        long long h = 0;
        cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("%s grid %3d N=%3d A %s B %s: %.1f cycles per MMA (nominal %d)%s\n", fast == 2 ? "stepped+commit" : (fast ? "stepped-desc  " : "rebuilt-desc  "), grid, n, a_mn ? "MN" : "K ", b_mn ? "MN" : "K ",
               (double)h / (iters * 4.0), n / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}
