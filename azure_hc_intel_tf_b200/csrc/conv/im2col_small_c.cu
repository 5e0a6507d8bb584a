// im2col_small_c.cu — explicit im2col for first-layer convolutions (Cin = 3): the TMA im2col path needs 64-channel
// (128-byte) rows, which an RGB image does not have. The patch matrix A[N*P*Q][Kpad] (K = R*S*Cin, zero padded to a
// multiple of 64) is written once and then feeds the tcgen05 GEMM kernels: forward is gemm_nt(A, W), the weight
// gradient is dense_wgrad(A, dY). The input gradient of a first layer is never needed.
// Role in the reference stack: MKL-DNN's convolution on the 3-channel input layer (K1/K3 for conv0).
#include "b200conv.h"
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kQT = 32;          // output pixels (along Q) per block
constexpr int kMaxSmemElems = 20 * 1024;

// A block owns kQT consecutive output pixels of one output row. The input rows they touch are staged in shared
// memory with coalesced loads; a k -> (row, offset) table removes all divisions from the gather; every output
// vector (8 consecutive k of one pixel) leaves as one coalesced 16-byte store.
__global__ void __launch_bounds__(256) im2col_kernel(const __nv_bfloat16* __restrict__ x, b200_bf16* __restrict__ a,
                                                     b200_conv_geom g, int kdim, int kpad, int span, int qblocks) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int* ktab = reinterpret_cast<int*>(smem_raw);                                  // [kpad]
  __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(smem_raw + kpad * 4);   // [R][span]
  const int sc = g.s * g.c;
  int b = blockIdx.x;
  const int qb = b % qblocks;
  b /= qblocks;
  const int p = b % g.p;
  const int n = b / g.p;
  const int q0 = qb * kQT;
  const int h0 = p * g.stride_h - g.pad_t;
  const int wc0 = (q0 * g.stride_w - g.pad_l) * g.c;       // first staged element's offset within an image row
  for (int k = threadIdx.x; k < kpad; k += blockDim.x) ktab[k] = k < kdim ? (k / sc) * span + (k % sc) : -1;
  const int rowlen = g.w * g.c;
  for (int idx = threadIdx.x; idx < g.r * span; idx += blockDim.x) {
    const int r = idx / span, e = idx - r * span;
    const int h = h0 + r, wc = wc0 + e;
    __nv_bfloat16 v = __float2bfloat16(0.0f);
    if (h >= 0 && h < g.h && wc >= 0 && wc < rowlen) v = x[((int64_t)n * g.h + h) * rowlen + wc];
    tile[idx] = v;
  }
  __syncthreads();
  const int kvecs = kpad / 8;
  const int step = g.stride_w * g.c;
  const int64_t row0 = ((int64_t)n * g.p + p) * g.q + q0;
  for (int v = threadIdx.x; v < kQT * kvecs; v += blockDim.x) {
    const int pix = v / kvecs, kv = v - pix * kvecs;
    if (q0 + pix >= g.q) continue;
    __nv_bfloat16 o[8];
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      const int t = ktab[kv * 8 + jj];
      o[jj] = t >= 0 ? tile[t + pix * step] : __float2bfloat16(0.0f);
    }
    *reinterpret_cast<uint4*>(a + ((row0 + pix) * kpad + kv * 8)) = *reinterpret_cast<const uint4*>(o);
  }
}

}  // namespace
}  // namespace b200

extern "C" int b200_im2col_small_c(const b200_bf16* x, b200_bf16* a, const b200_conv_geom* g, int32_t kpad, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(x && a && g, "im2col_small_c: null pointer");
  const int kdim = g->r * g->s * g->c;
  B200_CHECK_ARG(kpad >= kdim && kpad % 64 == 0, "im2col_small_c: kpad must be a multiple of 64 and >= R*S*C");
  const int span = ((kQT - 1) * g->stride_w + g->s) * g->c;
  B200_CHECK_ARG(g->r * span <= kMaxSmemElems, "im2col_small_c: filter too large for the staging tile");
  const int qblocks = (g->q + kQT - 1) / kQT;
  const size_t smem = (size_t)kpad * 4 + (size_t)g->r * span * 2;
  const int64_t blocks = (int64_t)g->n * g->p * qblocks;
  im2col_kernel<<<(unsigned)blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), a, *g, kdim, kpad, span, qblocks);
  B200_LAUNCH_CHECK();
  return 0;
}
