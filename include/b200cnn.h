/*
 * b200cnn.h — C ABI of libb200cnn.so: every hand-written sm_100a kernel of the framework.
 *
 * Conventions
 *   - activations are NHWC bf16 (uint16 storage), M = N*H*W rows of C channels; C % 8 == 0 unless noted;
 *   - statistics, gradients of parameters, master weights and optimizer state are fp32;
 *   - `stream` is a cudaStream_t passed as void*; every function only enqueues work on it;
 *   - return value: 0 on success, otherwise a cudaError_t (or -1 for an argument the kernel rejects);
 *   - nothing here allocates device memory unless the comment says so; scratch is supplied by the caller.
 *
 * Reference counterparts: the reference has no kernels of its own. Each entry names the
 * external primitive that plays its role when the reference launcher passes `--mkl=TRUE --data_format=NCHW
 * --optimizer=momentum --variable_update=horovod` (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:72-77).
 */
#ifndef B200CNN_H_
#define B200CNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint16_t b200_bf16;

const char* b200_version(void);
/* Last error string recorded by the library on this thread (argument checks, driver errors). */
const char* b200_last_error(void);

/* ------------------------------------------------------------------ optimizer (K10; TF ApplyMomentum) */

/* Momentum SGD over a flat range of n elements (n % 1024 == 0):
 *   g' = grad*grad_scale + wd*decay[t]*w ; (LARS: g' *= trust[t]) ; m = mu*m + g' ; w -= lr*m ; shadow = bf16(w)
 * block_tensor[i] names the tensor owning elements [1024 i, 1024 i + 1024) or -1 for padding; it may be NULL
 * (then decay applies everywhere and trust is ignored). shadow may be NULL. lr_dev (optional): device float that
 * overrides `lr` at run time, so a learning-rate schedule works under a captured CUDA graph
 * (run-tf-sing-ucx-openmpi.sh:73). */
int b200_sgd_momentum_flat(float* master, float* mom, const float* grad, b200_bf16* shadow, int64_t n,
                           const int32_t* block_tensor, const float* tensor_decay, const float* tensor_lars,
                           const float* trust, float lr, float mu, float wd, float grad_scale, const float* lr_dev,
                           void* stream);

/* RMSProp / Adam on the same flat buffers, TensorFlow update rules (tf_cnn_benchmarks --optimizer=rmsprop|adam):
 *   rmsprop: ms = rho*ms + (1-rho)*g^2; mom = mu*mom + lr*g/sqrt(ms+eps); w -= mom      (ms starts at 1, mom at 0)
 *   adam:    m = b1*m + (1-b1)*g; v = b2*v + (1-b2)*g^2; w -= lr_t*m/(sqrt(v)+eps), lr_t = lr*sqrt(1-b2^t)/(1-b1^t)
 * g = grad*grad_scale + wd*decay[tensor]*w. lr_dev (optional) overrides lr / lr_t from device memory (schedules and the
 * Adam bias correction under a captured CUDA graph). Same block table and bf16 shadow as b200_sgd_momentum_flat. */
int b200_rmsprop_flat(float* master, float* mom, float* ms, const float* grad, b200_bf16* shadow, int64_t n,
                      const int32_t* block_tensor, const float* tensor_decay, float lr, float rho, float mu, float eps,
                      float wd, float grad_scale, const float* lr_dev, void* stream);
int b200_adam_flat(float* master, float* m, float* v, const float* grad, b200_bf16* shadow, int64_t n,
                   const int32_t* block_tensor, const float* tensor_decay, float lr_t, float beta1, float beta2, float eps,
                   float wd, float grad_scale, const float* lr_dev, void* stream);

/* LARS: per-tensor squared norms of w and grad*grad_scale over a flat range → norms[2t], norms[2t+1]
 * (norms must be zeroed by the caller), then trust[t] = eta*|w| / (|g| + wd*decay[t]*|w| + eps) for tensors
 * with tensor_lars[t] != 0, else 1. No reference counterpart beyond the product definition (BASELINE.json). */
int b200_lars_norms(const float* master, const float* grad, int64_t n, const int32_t* block_tensor,
                    float* norms, float grad_scale, void* stream);
int b200_lars_trust(const float* norms, const float* tensor_decay, const float* tensor_lars, float* trust,
                    int32_t num_tensors, float eta, float wd, float eps, void* stream);

/* Deterministic variant used by the framework: per-1024-block partial norms (block_partials: 2 floats per block)
 * folded per tensor in a fixed order; tensor t owns blocks [tensor_blk_lo[t], +tensor_blk_cnt[t]). Identical inputs
 * give bitwise identical trust ratios, so data-parallel replicas stay bit-identical. */
int b200_lars_trust_det(const float* master, const float* grad, int64_t n, float* block_partials,
                        const int32_t* tensor_blk_lo, const int32_t* tensor_blk_cnt, const float* tensor_decay,
                        const float* tensor_lars, float* trust, int32_t num_tensors, float eta, float wd, float eps,
                        float grad_scale, void* stream);

/* ------------------------------------------------------------------ batch norm (K4/K5; MKL-DNN fused BN) */

/* Bytes of scratch b200_bn_stats needs for a tensor of `rows` x `c`. */
int64_t b200_bn_stats_scratch_bytes(int64_t rows, int32_t c);

/* Per-channel batch statistics of x[rows, c] (bf16), training mode:
 *   mean[c], rstd[c] = 1/sqrt(var+eps) (biased var), scale[c] = gamma*rstd, shift[c] = beta - mean*scale,
 *   running_mean = decay*running_mean + (1-decay)*mean, running_var likewise with the unbiased variance.
 * `scratch` holds per-block partial sums (b200_bn_stats_scratch_bytes); its first 4 bytes are a ticket
 * counter that must be zero on entry and is left zero on exit. */
int b200_bn_stats(const b200_bf16* x, int64_t rows, int32_t c, const float* gamma, const float* beta,
                  float* running_mean, float* running_var, float decay, float eps, float* mean, float* rstd,
                  float* scale, float* shift, void* scratch, void* stream);

/* b200_bn_stats for callers that cannot allocate device memory (memory-planned regions of a host framework,
 * plug-ins given fixed buffers only, call sites where the host allocator must not be re-entered): the scratch is a
 * fixed 2 MiB array inside the library image (b200_bn_stats_noalloc_max_bytes()), held by every GPU that loads the
 * library until the process exits. Same inputs, outputs and results as b200_bn_stats, minus the scratch argument.
 * This entry point has no counterpart in the reference.
 *  - Sizes whose scratch need (b200_bn_stats_scratch_bytes) exceeds the array are rejected with -1 and a message in
 *    b200_last_error(). The present kernel geometry needs at most ~0.6 MiB for any shape up to c = 4096, so every
 *    BatchNorm layer of the model zoo fits; the check guards larger channel counts and future geometries.
 *  - The first call on each device looks up the array's device address (with that device current); make that
 *    first call before starting any stream capture.
 *  - All calls on a device share the one array: two calls that may run on the GPU at the same time (different
 *    streams, or different threads without ordering) race. Order them on one stream or with events.
 *  - It performs no cudaMalloc/cudaFree and calls no allocating library routine. */
int b200_bn_stats_noalloc(const b200_bf16* x, int64_t rows, int32_t c, const float* gamma, const float* beta,
                          float* running_mean, float* running_var, float decay, float eps, float* mean, float* rstd,
                          float* scale, float* shift, void* stream);
int64_t b200_bn_stats_noalloc_max_bytes(void);

/* Same finalisation, but from per-channel sums already produced by a convolution epilogue:
 * sums[0..c) = Σx, sums[c..2c) = Σx². */
int b200_bn_finalize(const float* sums, int64_t rows, int32_t c, const float* gamma, const float* beta,
                     float* running_mean, float* running_var, float decay, float eps, float* mean, float* rstd,
                     float* scale, float* shift, void* stream);

/* The same, from [nrows][2][c] partial sums as written by b200_conv_fprop (folded here in row order). */
int b200_bn_finalize_partials(const float* partials, int64_t nrows, int64_t rows, int32_t c, const float* gamma,
                              const float* beta, float* running_mean, float* running_var, float decay, float eps,
                              float* mean, float* rstd, float* scale, float* shift, void* stream);

/* z = act(scale*x + shift (+ residual)); relu == 1 applies max(0, .), relu == 2 ReLU6 = min(max(0, .), 6) (MobileNet;
 * pass relu_mask then: the backward kernels take the pass-through set 0 < f < 6 from it). residual may be NULL. In
 * place allowed.
 * relu_mask (optional, rows*c/8 bytes): bit k of byte v = (element 8v+k passed the ReLU); the backward kernels read
 * it instead of z, 16x fewer bytes — used when the mask cannot be recomputed from x alone (residual add). */
int b200_bn_apply(const b200_bf16* x, const b200_bf16* residual, b200_bf16* z, uint8_t* relu_mask, int64_t rows,
                  int32_t c, const float* scale, const float* shift, int32_t relu, void* stream);

/* Backward, pass 1: dbeta[c] = Σ dzm, dgamma[c] = Σ dzm*xhat with dzm = dz masked by the ReLU
 * (mask from relu_mask when given, else from z when z != NULL, else recomputed from scale*x+shift; relu == 0 → no
 * mask). Needs the same scratch as b200_bn_stats. */
int b200_bn_bwd_reduce(const b200_bf16* dz, const b200_bf16* x, const b200_bf16* z, const uint8_t* relu_mask,
                       int64_t rows, int32_t c, const float* mean, const float* rstd, const float* scale, const float* shift, int32_t relu,
                       float* dgamma, float* dbeta, void* scratch, void* stream);

/* Backward, pass 2: dx = scale*(dzm - dbeta/rows - xhat*dgamma/rows); dres = dzm when dres != NULL. */
int b200_bn_bwd_apply(const b200_bf16* dz, const b200_bf16* x, const b200_bf16* z, const uint8_t* relu_mask,
                      b200_bf16* dx, b200_bf16* dres, int64_t rows, int32_t c, const float* mean, const float* rstd, const float* scale,
                      const float* shift, const float* dgamma, const float* dbeta, int32_t relu, void* stream);

/* ------------------------------------------------------------------ bias / ReLU (VGG-style conv, affine) */

/* z = act(x + bias[c]); in place allowed. */
int b200_bias_act(const b200_bf16* x, const float* bias, b200_bf16* z, int64_t rows, int32_t c, int32_t relu,
                  void* stream);
/* dx = dz masked by (z > 0) when relu; dbias[c] = Σ dx. Needs b200_bn_stats-sized scratch. dx may alias dz. */
int b200_bias_act_bwd(const b200_bf16* dz, const b200_bf16* z, b200_bf16* dx, float* dbias, int64_t rows,
                      int32_t c, int32_t relu, void* scratch, void* stream);

/* ------------------------------------------------------------------ pooling (K7; MKL-DNN pooling) */

typedef struct b200_pool_geom {
  int32_t n, h, w, c;        /* input */
  int32_t p, q;              /* output spatial */
  int32_t kh, kw, sh, sw;
  int32_t pad_t, pad_l;      /* leading pads; trailing pads are implied by p, q */
} b200_pool_geom;

/* Max pooling; idx[n,p,q,c] (uint8) records the winning tap (r*kw+s), first maximum in scan order. */
int b200_maxpool_fwd(const b200_bf16* x, b200_bf16* y, uint8_t* idx, const b200_pool_geom* g, void* stream);
int b200_maxpool_bwd(const b200_bf16* dy, const uint8_t* idx, b200_bf16* dx, const b200_pool_geom* g, void* stream);
/* Average pooling that divides by the number of taps inside the image (TF 'SAME' semantics). */
int b200_avgpool_fwd(const b200_bf16* x, b200_bf16* y, const b200_pool_geom* g, void* stream);
int b200_avgpool_bwd(const b200_bf16* dy, b200_bf16* dx, const b200_pool_geom* g, void* stream);
/* Global average over H*W: x[n,hw,c] → y[n,c] (spatial_mean), and its backward. */
int b200_spatial_mean_fwd(const b200_bf16* x, b200_bf16* y, int32_t n, int32_t hw, int32_t c, void* stream);
int b200_spatial_mean_bwd(const b200_bf16* dy, b200_bf16* dx, int32_t n, int32_t hw, int32_t c, void* stream);

/* ------------------------------------------------------------------ loss (K8; TF softmax_xent) */

/* logits[b, classes] fp32 row-major with leading dimension ld. loss_rows[b] = -log softmax[label];
 * *loss = mean over rows; dlogits = (softmax - onehot) * (grad_scale / b) as bf16 with leading dimension ld_d
 * (columns >= classes are zero-filled up to ld_d). */
int b200_softmax_xent(const float* logits, int32_t ld, const int64_t* labels, int32_t b, int32_t classes,
                      float* loss_rows, float* loss, b200_bf16* dlogits, int32_t ld_d, float grad_scale, void* stream);

/* ------------------------------------------------------------------ input (K11) */

/* uint8 NHWC (c_in channels) → bf16 NHWC with c_out >= c_in channels (extra channels zero), x/127.5 - 1. */
int b200_normalize_u8(const uint8_t* x, b200_bf16* y, int64_t pixels, int32_t c_in, int32_t c_out, void* stream);

/* ------------------------------------------------------------------ elementwise helpers */

/* out = a + b (bf16), n % 8 == 0. Used to sum the two gradient branches at a residual fork. */
int b200_add(const b200_bf16* a, const b200_bf16* b, b200_bf16* out, int64_t n, void* stream);
/* out = a + (mask bit ? b : 0), mask = one byte per 8 consecutive elements, bit k gates element k (the ReLU mask
 * b200_bn_apply writes). Sums the two gradient branches at a residual fork when one of them still has to pass the
 * block-output ReLU, without materialising the gated copy. n % 8 == 0. */
int b200_add_masked(const b200_bf16* a, const b200_bf16* b, const uint8_t* mask, b200_bf16* out, int64_t n,
                    void* stream);
/* out = relu ? max(a + b, 0) : a + b - the residual sum of blocks whose sum is not behind a batch-norm (CIFAR ResNets,
 * pre-activation ResNets); b200_relu_bwd: dx = z > 0 ? dz : 0 with z the output above. n % 8 == 0, in place allowed. */
int b200_add_act(const b200_bf16* a, const b200_bf16* b, b200_bf16* out, int64_t n, int32_t relu, void* stream);
int b200_relu_bwd(const b200_bf16* dz, const b200_bf16* z, b200_bf16* dx, int64_t n, void* stream);
/* fp32 → bf16 cast of n elements. */
int b200_cast_f32_bf16(const float* x, b200_bf16* y, int64_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CNN_H_ */
