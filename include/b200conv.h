/*
 * b200conv.h — C ABI of the tcgen05/TMEM/TMA implicit-GEMM convolution and GEMM kernels in libb200cnn.so.
 *
 * Layouts: activations NHWC bf16; filters KRSC bf16 (PyTorch [K,C,R,S] in channels_last memory format);
 * filter gradients KRSC fp32. All functions only enqueue work on `stream` (a cudaStream_t) and return 0 on
 * success. Reference counterparts: MKL-DNN convolution forward / backward-data / backward-weights and
 * inner-product under --mkl=TRUE (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76); K1-K3, K8 in SURVEY.md.
 */
#ifndef B200CONV_H_
#define B200CONV_H_

#include <stdint.h>

#include "b200cnn.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_conv_geom {
  int32_t n, h, w, c;          /* input  N,H,W,Cin  */
  int32_t k, r, s;             /* filter Cout,R,S   */
  int32_t p, q;                /* output spatial    */
  int32_t stride_h, stride_w;
  int32_t pad_t, pad_l, pad_b, pad_r;
} b200_conv_geom;

enum { B200_CONV_FPROP = 0, B200_CONV_DGRAD = 1, B200_CONV_WGRAD = 2 };

/* Non-zero if the tensor-core kernels cover this geometry for `op`, else 0 (callers fall back to a library). For
 * B200_CONV_DGRAD the value says which entry point: 1 = b200_conv_dgrad, 2 = b200_conv_dgrad_strided. */
int b200_conv_supported(const b200_conv_geom* g, int op);

/* Smallest input-channel count b200_conv_supported accepts (default 8: thinner-than-64 layers run on the same kernels
 * through TMA zero-fill; B200_CONV_MIN_C=64 hands them to the library, which measured faster on them; read once). */
int b200_conv_min_channels(void);

/* y[n,p,q,k] = sum_{r,s,c} x[n, p*sh - pad_t + r, q*sw - pad_l + s, c] * w[k,r,s,c].
 * col_stats (optional): [b200_conv_stats_rows(g)][2][k] fp32 buffer for the batch-norm statistics of the stored
 * (bf16-rounded) outputs, produced by the epilogue so that y is not read again. The first
 * b200_conv_stats_first_result_row(g) rows are kernel scratch; the rows from there on (one per CTA of the persistent
 * grid) hold partial column sums [0][k] and sums of squares [1][k] — fold those with b200_bn_finalize_partials.
 * The kernel initialises everything it reads; no memset is needed. */
int b200_conv_fprop(const b200_bf16* x, const b200_bf16* w, b200_bf16* y, const b200_conv_geom* g, float* col_stats,
                    void* stream);
int64_t b200_conv_stats_rows(const b200_conv_geom* g);
int64_t b200_conv_stats_first_result_row(const b200_conv_geom* g);

/* dx = gradient of fprop wrt x. Stride 1 (any filter), or strided 1x1 without padding. */
int b200_conv_dgrad(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* stream);

/* dx for any other strided filter (3x3/2 of ResNet v1.5 and the Inception reductions, 7x7/2, ...), strides up to 4:
 * one stride-1 tensor-core launch per stride phase (a, b) over the sub-filter w[:, a::sh, b::sw, :], whose rows the
 * epilogue scatters to every sh-th / sw-th pixel of dx - no zero-upsampled copy of dy and no wasted MMAs.
 * `workspace` (b200_conv_dgrad_strided_workspace_bytes) holds the packed sub-filter of the phase in flight. */
int b200_conv_dgrad_strided(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* workspace,
                            void* stream);
int64_t b200_conv_dgrad_strided_workspace_bytes(const b200_conv_geom* g);

/* dw[k,r,s,c] (fp32) = sum over pixels of dy[.,k] * x[window]. dw is overwritten. The pixel range is split over
 * CTAs; with `workspace` of b200_conv_wgrad_workspace_bytes(g) bytes the per-slice partials are folded in a fixed
 * order (bitwise reproducible); with workspace == NULL they are combined with float atomics in dw instead. */
int b200_conv_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, void* workspace,
                    void* stream);
int64_t b200_conv_wgrad_workspace_bytes(const b200_conv_geom* g);

/* Depthwise convolution (groups == channels; g->k == g->c, c % 8 == 0, filters up to 7x7, any stride; filters [C][R][S]
 * = the KRSC layout of a [C,1,R,S] weight, filter gradient fp32 in the same layout). Bandwidth-bound vector kernels
 * (csrc/kernels/depthwise.cu) on the mixed-precision FMA (bf16 x bf16 + fp32): a register-blocked fast path for 3x3 at
 * stride 1 / 2 (MobileNet), shared-memory-staged taps for everything else (NASNet's 5x5 / 7x7). wgrad folds its per-CTA
 * partials in a fixed order (bitwise reproducible). */
int b200_dwconv_supported(const b200_conv_geom* g);
int b200_dwconv_fprop(const b200_bf16* x, const b200_bf16* w, b200_bf16* y, const b200_conv_geom* g, void* stream);
int b200_dwconv_dgrad(const b200_bf16* dy, const b200_bf16* w, b200_bf16* dx, const b200_conv_geom* g, void* stream);
int b200_dwconv_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, const b200_conv_geom* g, void* workspace,
                      void* stream);
int64_t b200_dwconv_wgrad_workspace_bytes(const b200_conv_geom* g);

/* out[m, n] = a[m, k] . b[n, k]^T ; a, b row-major bf16, k % 8 == 0 (a partial last 64-wide k-block is zero-filled by
 * TMA); out bf16 or fp32 with leading dimension ldo. */
int b200_gemm_nt(const b200_bf16* a, const b200_bf16* b, void* out, int32_t m, int32_t n, int32_t k, int32_t ldo,
                 int32_t out_fp32, void* stream);

/* b200_gemm_nt that also emits per-column statistics of the stored outputs, like b200_conv_fprop
 * (col_stats: [b200_gemm_stats_rows(m, n)][2][n] fp32; results start at row b200_gemm_stats_first_result_row). */
int b200_gemm_nt_stats(const b200_bf16* a, const b200_bf16* b, void* out, int32_t m, int32_t n, int32_t k,
                       int32_t ldo, int32_t out_fp32, float* col_stats, void* stream);
int64_t b200_gemm_stats_rows(int32_t m, int32_t n);
int64_t b200_gemm_stats_first_result_row(int32_t m, int32_t n);

/* Patch matrix for small-Cin (first-layer) convolutions: a[n*p*q][kpad] bf16, column (r*S + s)*Cin + c, zero padded
 * from R*S*Cin up to kpad (a multiple of 8). Forward: gemm_nt(a, w_padded); weight gradient: dense_wgrad(a, dy). */
int b200_im2col_small_c(const b200_bf16* x, b200_bf16* a, const b200_conv_geom* g, int32_t kpad, void* stream);

/* out[m, n] (bf16) = a[m, k] . b[k, n]; a has leading dimension lda >= k (multiple of 8; elements past k must be
 * zero or lda == k), b is row-major [k][n] and is read in place as an MN-major operand; n % 64 == 0. */
int b200_gemm_nn(const b200_bf16* a, int32_t lda, const b200_bf16* b, b200_bf16* out, int32_t m, int32_t n, int32_t k,
                 void* stream);

/* Dense-layer weight gradient: dw[out][in] (fp32, overwritten) = dy[batch][out]^T . x[batch][in]; dy rows have
 * leading dimension dy_ld (multiple of 8, >= out; the padding columns are ignored). in % 64 == 0. */
int b200_dense_wgrad(const b200_bf16* x, const b200_bf16* dy, float* dw, int32_t batch, int32_t in_features,
                     int32_t out_features, int32_t dy_ld, void* workspace, void* stream);
int64_t b200_dense_wgrad_workspace_bytes(int32_t batch, int32_t in_features, int32_t out_features);

/* Reduce the per-group partials written by b200_conv_fprop into sums[0..k) = Σy, sums[k..2k) = Σy². */
int b200_conv_stats_reduce(const float* col_stats, int64_t rows, int32_t k, float* sums, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CONV_H_ */
