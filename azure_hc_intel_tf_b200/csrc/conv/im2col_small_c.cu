// im2col_small_c.cu — explicit im2col for first-layer convolutions (Cin = 3): the TMA im2col path needs 64-channel
// (128-byte) rows, which an RGB image does not have. The patch matrix A[N*P*Q][Kpad] (K = R*S*Cin, zero padded to a
// multiple of 64) is written once and then feeds the tcgen05 GEMM kernels: forward is gemm_nt(A, W), the weight
// gradient is dense_wgrad(A, dY). The input gradient of a first layer is never needed.
// Role in the reference stack: MKL-DNN's convolution on the 3-channel input layer (K1/K3 for conv0).
#include "b200conv.h"
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kMaxSmemElems = 24 * 1024;
constexpr int kRowsPerBlock = 4;          // output rows handled by one block (the k table is built once per block)

// A block owns `qt` consecutive output pixels (a whole output row when it fits) of kRowsPerBlock output rows. The
// input rows a strip touches are staged in shared memory with 32-bit loads; thread t always produces k-vector
// (t % kvecs) of its pixels, so the eight (row, offset) table entries of that vector live in registers and the gather
// is eight 16-bit shared loads + one coalesced 16-byte store per vector.
__global__ void __launch_bounds__(256) im2col_kernel(const __nv_bfloat16* __restrict__ x, b200_bf16* __restrict__ a,
                                                     b200_conv_geom g, int kdim, int kpad, int span, int qblocks, int qt) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(smem_raw);   // [R][span], span even
  const int sc = g.s * g.c;
  int b = blockIdx.x;
  const int qb = b % qblocks;
  b /= qblocks;
  const int pblocks = (g.p + kRowsPerBlock - 1) / kRowsPerBlock;
  const int pb = b % pblocks;
  const int n = b / pblocks;
  const int q0 = qb * qt;
  const int kvecs = kpad / 8;
  const int jv = threadIdx.x % kvecs;                 // this thread's k-vector
  const int pix_lane = threadIdx.x / kvecs, pix_step = blockDim.x / kvecs;
  int tab[8];
#pragma unroll
  for (int jj = 0; jj < 8; ++jj) {
    const int k = jv * 8 + jj;
    tab[jj] = k < kdim ? (k / sc) * span + (k % sc) : -1;
  }
  // the staged window starts one element early when needed so that it begins on an even element of the image row
  // (32-bit loads); wc0 may be negative (left padding)
  const int wc0 = (q0 * g.stride_w - g.pad_l) * g.c;
  const int lead = wc0 & 1;
  const int rowlen = g.w * g.c;
  const int step = g.stride_w * g.c;
  const bool rows_word_aligned = (rowlen & 1) == 0;
  for (int pr = 0; pr < kRowsPerBlock; ++pr) {
    const int p = pb * kRowsPerBlock + pr;
    if (p >= g.p) break;
    const int h0 = p * g.stride_h - g.pad_t;
    __syncthreads();                                  // the previous row's gather is done with the tile
    const int words = span / 2;
    for (int idx = threadIdx.x; idx < g.r * words; idx += blockDim.x) {
      const int r = idx / words, wd = idx - r * words;
      const int h = h0 + r, wc = wc0 - lead + 2 * wd;          // even element offset within the image row
      uint32_t v = 0;
      if (h >= 0 && h < g.h) {
        const __nv_bfloat16* row = x + ((int64_t)n * g.h + h) * rowlen;
        if (rows_word_aligned && wc >= 0 && wc + 1 < rowlen) {
          v = *reinterpret_cast<const uint32_t*>(row + wc);
        } else {
          const uint32_t lo = (wc >= 0 && wc < rowlen) ? (uint32_t)__bfloat16_as_ushort(row[wc]) : 0u;
          const uint32_t hi = (wc + 1 >= 0 && wc + 1 < rowlen) ? (uint32_t)__bfloat16_as_ushort(row[wc + 1]) : 0u;
          v = lo | (hi << 16);
        }
      }
      reinterpret_cast<uint32_t*>(tile)[idx] = v;
    }
    __syncthreads();
    const int64_t row0 = ((int64_t)n * g.p + p) * g.q + q0;
    const unsigned short* t16 = reinterpret_cast<const unsigned short*>(tile) + lead;
    for (int pix = pix_lane; pix < qt && q0 + pix < g.q; pix += pix_step) {
      const unsigned short* src = t16 + pix * step;
      uint32_t o[4];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const uint32_t lo = tab[2 * jj] >= 0 ? (uint32_t)src[tab[2 * jj]] : 0u;
        const uint32_t hi = tab[2 * jj + 1] >= 0 ? (uint32_t)src[tab[2 * jj + 1]] : 0u;
        o[jj] = lo | (hi << 16);
      }
      *reinterpret_cast<uint4*>(a + ((row0 + pix) * kpad + jv * 8)) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

}  // namespace
}  // namespace b200

extern "C" int b200_im2col_small_c(const b200_bf16* x, b200_bf16* a, const b200_conv_geom* g, int32_t kpad, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(x && a && g, "im2col_small_c: null pointer");
  const int kdim = g->r * g->s * g->c;
  B200_CHECK_ARG(kpad >= kdim && kpad % 8 == 0, "im2col_small_c: kpad must be a multiple of 8 and >= R*S*C");
  // pixels per strip: a whole output row when its staged window fits, else the largest strip that does
  const int kvecs = kpad / 8;
  int qt = g->q;
  auto span_of = [&](int t) { return (((t - 1) * g->stride_w + g->s) * g->c + 2) / 2 * 2; };   // +1 lead element, even
  while (qt > 8 && g->r * span_of(qt) > kMaxSmemElems) qt = (qt + 1) / 2;
  const int span = span_of(qt);
  B200_CHECK_ARG(g->r * span <= kMaxSmemElems, "im2col_small_c: filter too large for the staging tile");
  B200_CHECK_ARG(kvecs <= 256, "im2col_small_c: kpad too large");
  const int qblocks = (g->q + qt - 1) / qt;
  const int pblocks = (g->p + kRowsPerBlock - 1) / kRowsPerBlock;
  const size_t smem = (size_t)g->r * span * 2;
  const int64_t blocks = (int64_t)g->n * pblocks * qblocks;
  const int threads = 256 / kvecs * kvecs;             // a whole number of pixels per pass
  static std::atomic<uint64_t> attr_set{0};
  if (first_use_on_device(attr_set))
    B200_CUDA(cudaFuncSetAttribute(im2col_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmemElems * 2));
  im2col_kernel<<<(unsigned)blocks, threads, smem, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), a, *g, kdim, kpad, span, qblocks, qt);
  B200_LAUNCH_CHECK();
  return 0;
}
