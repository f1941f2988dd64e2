/* b200lmsfc.h -- C ABI of libb200lmsfc.so: the L-MSFC rate-distortion training step on B200 (sm_100a).
 *
 * The reference (herok97/L-MSFC-V2-Training = CompressAI-Vision 1.1.5.dev0 + CompressAI 1.2.4) has
 * no FFI for this path: every function below replaces a PyTorch-eager interface (or, for the CDF
 * builder, the pybind11 `_CXX` module) and cites it.  Paths are relative to
 * /root/reference/compressai/compressai/.  INTEGRATION.md shows the ctypes stubs a maintainer adds.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is a DEVICE pointer unless it says "host";
 *  - activations are NHWC bf16 ([B][H][W][C], C a multiple of 64); latents / likelihoods are NHWC fp32;
 *  - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it, no internal threads;
 *  - return 0 on success, non-zero on error (1 bad argument, 2 CUDA error, 3 domain error, the
 *    ValueError cases of the reference); lmsfc_last_error() returns the thread-local message;
 *  - caller owns all buffers; nothing is allocated or kept between calls.
 */
#ifndef B200LMSFC_H
#define B200LMSFC_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* lmsfc_last_error(void);
int lmsfc_version(void);
/* number of CUDA kernels this library has launched (or captured into a graph) in this process */
uint64_t lmsfc_launch_count(void);

/* ------------------------------------------------------------------ conv / deconv (tcgen05 implicit GEMM)
 * Replaces models/utils.py:128-146: conv = nn.Conv2d(k, stride, padding=k/2),
 * deconv = nn.ConvTranspose2d(k, stride, padding=k/2, output_padding=stride-1), and their autograd.
 * Weights live in "GEMM layout" [D0][T][D1] with T = k*k taps: conv D0=Cout, D1=Cin (torch
 * [Cout][Cin][kh][kw] permuted (0,2,3,1)); deconv D0=Cin, D1=Cout (torch [Cin][Cout][kh][kw] permuted).
 * `w` is always the ONE bf16 copy in that layout: forward and data-gradient kernels read it either as a
 * K-major or as an MN-major tensor-core operand, so no transposed copy exists. */
typedef struct {
  int kind;          /* 0 conv, 1 deconv */
  int k, stride;     /* 5/2 or 3/1 in L-MSFC */
  int B, Hin, Win;   /* input spatial size */
  int Hout, Wout;    /* output spatial size: conv ceil(Hin/stride); deconv <= stride*Hin (cropped) */
  int c0, c1;        /* input channels of source 0 and of the optional concatenated source 1 */
  int Cout;
} lmsfc_conv_desc;

/* Thread-local switch for the entropy-coded path (models/google.py:313-333): with on != 0 every conv / deconv / GDN
 * launch of the calling thread uses one fixed tiling (128-wide tiles, no split K), so a tensor's bits do not depend on
 * the batch it was computed in or on the device's SM count -- h_s(z_hat) picks CDF rows through hard thresholds
 * (entropy_models.py:684-689) and must be bit-identical in compress() and decompress(). */
int lmsfc_set_batch_invariant(int on);
/* out = act(conv(cat[src0,src1]) + bias).  relu: 0 none, 1 ReLU, 2 LeakyReLU(0.01) (the mean-scale hyperprior's
 * activation, models/google.py:377-391).  Any of out_bf16 / out_f32 / out_abs (|.| as bf16, the
 * abs(y) fed to h_a, models/google.py:294) may be null. */
int lmsfc_conv_forward(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* w,
                       const float* bias, int relu, void* out_bf16, float* out_f32, void* out_abs, void* stream);
/* grad wrt the first n_in input channels (n_in <= c0; source 1 is data and gets none). */
int lmsfc_conv_dgrad(const lmsfc_conv_desc* d, const void* gout, const void* w, int n_in, void* gin_bf16,
                     void* stream);
/* gw (fp32, GEMM layout, caller-zeroed) += weight gradient; gbias (fp32 [Cout], nullable) += column sums. */
int lmsfc_conv_wgrad(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* gout, float* gw,
                     float* gbias, void* stream);

/* conv / deconv with the GDN / IGDN that follows it fused into the same kernel (models/utils.py:128-146 followed by
 * layers/gdn.py:77-92): the CTA that owns all 256 channels of 128 pixels squares its own accumulator tile in shared
 * memory and runs the channel contraction as a second tcgen05 GEMM -- x^2 never makes a round trip through HBM.
 * Cout must be 256.  x_out (conv output, read by the backward pass), x2_out (x^2, read by gamma's gradient) and rn_out
 * (rsqrt(norm)) are nullable (inference); y_out = x * rsqrt(beta' + gamma' . x^2) (inverse: * sqrt).
 * lmsfc_conv_gdn_fusable: *fusable = 1 when the launch (forward, or with dgrad != 0 the data gradient w.r.t. the first
 * n_in input channels) would use whole-K 256-wide tiles anyway, i.e. fusing costs no occupancy; callers fall back to
 * the separate calls otherwise. */
int lmsfc_conv_gdn_fusable(const lmsfc_conv_desc* d, int dgrad, int n_in, int* fusable);
int lmsfc_conv_gdn_forward(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* w,
                           const float* bias, const void* gamma_eff, const float* beta_eff, int inverse, void* x_out,
                           void* x2_out, void* y_out, void* rn_out, void* stream);
/* The backward counterpart: the data gradient of conv `d` w.r.t. its first n_in (= 256) input channels IS the gradient
 * w.r.t. the output y of the GDN that produced them; the same kernel adds `addend` (the gradient from y's other
 * consumer, nullable), forms t = dL/dnorm (stored to t_out for gamma's / beta's gradient, lmsfc_gdn_wgrad), contracts
 * it with gamma and writes gx_out = the gradient w.r.t. the GDN's INPUT x.  x, rn: saved by the forward pass. */
int lmsfc_conv_dgrad_gdn_backward(const lmsfc_conv_desc* d, const void* gout, const void* w, int n_in, const void* addend,
                                  const void* x, const void* rn, const void* gamma_eff, int inverse, void* t_out,
                                  void* gx_out, void* stream);

/* ------------------------------------------------------------------ GDN / IGDN   (layers/gdn.py:77-92)
 * y = x * rsqrt(beta' + gamma' . x^2)  (inverse: * sqrt).  gamma_eff bf16 [C][C] ([out][in]) and
 * beta_eff fp32 [C] are the reparametrised values (lmsfc_reparam_forward).  rn = rsqrt(norm) is
 * saved (bf16) for backward.  x2_scratch (P*C bf16) receives x^2, which the backward pass's gamma gradient reads again;
 * NULL (inference): x^2 is formed inside the GEMM by its producer warps and nothing is kept. */
int lmsfc_gdn_forward(const void* x, const void* gamma_eff, const float* beta_eff, int inverse, int64_t P, int C,
                      void* x2_scratch, void* y, void* rn, void* stream);
/* gx written.  x2 = x^2 (bf16): x2_ready != 0 -> kept from the forward pass, else recomputed into it.  t_scratch receives
 * t = d loss / d norm (bf16), formed by the GEMM's producer warps.  The parameter gradients g_gamma_eff fp32 [C][C] and
 * g_beta_eff fp32 [C] (accumulated, caller-zeroed) are either both given -- computed here -- or both NULL and computed
 * later / on another stream by lmsfc_gdn_wgrad(t, x2): they are off the critical path. */
int lmsfc_gdn_backward(const void* gy, const void* x, const void* rn, const void* gamma_eff, int inverse, int64_t P,
                       int C, void* x2, int x2_ready, void* t_scratch, void* gx, float* g_gamma_eff, float* g_beta_eff,
                       void* stream);
int lmsfc_gdn_wgrad(const void* t, const void* x2, int64_t P, int C, float* g_gamma_eff, float* g_beta_eff, void* stream);
/* NonNegativeParametrizer (ops/parametrizers.py:38-64) with LowerBound's gradient (ops/bound_ops.py:40-42):
 * out = max(p, bound)^2 - pedestal ; g_p += g_out * 2 max(p,bound) * [(p >= bound) | (g < 0)]. */
int lmsfc_reparam_forward(const float* p, int64_t n, float bound, float pedestal, float* out_f32, void* out_bf16,
                          void* stream);
int lmsfc_reparam_backward(const float* g_out, const float* p, int64_t n, float bound, float* g_p, void* stream);
/* The same two maps over a device table of segments (every GDN layer of a model in one launch):
 * table[s] = {param offset, n, out_f32 offset or -1, out_bf16 offset or -1, g_eff offset} (int64, offsets in
 * elements from params / out_f32 / out_bf16 / g_eff); bounds[s] fp32; max_n = largest n. */
int lmsfc_reparam_forward_multi(const float* params, const int64_t* table, const float* bounds, int nseg, int64_t max_n,
                                float pedestal, float* out_f32, void* out_bf16, void* stream);
int lmsfc_reparam_backward_multi(const float* g_eff, const float* params, const int64_t* table, const float* bounds,
                                 int nseg, int64_t max_n, float* g_params, void* stream);

/* ------------------------------------------------------------------ entropy models (entropy_models/entropy_models.py)
 * mode: 0 "noise" (:161-165), 1 "dequantize" (:167-176); lmsfc_gc_* also 2 "given" (input already quantised).
 * lmsfc_gc_forward with scales == NULL only quantises (y_hat / y_hat_bf16), so the likelihood can run later / on
 * another stream from the quantised values.  noise: injected U(-.5,.5) draw (parity
 * tests) or null -> in-kernel Philox keyed by (seed, offset [+ rng_state[1]]).  sum_ln (nullable):
 * += sum(ln likelihood), the rate term of losses/rate_distortion.py:52-55.  y_hat_bf16 (nullable): bf16
 * copy of y_hat, the operand of the next conv.  Backward: g_yhat (fp32) or g_yhat_bf16 (a dgrad output),
 * either nullable. */
int lmsfc_gc_forward(const float* y, const float* scales, const float* means, const float* noise, int64_t n, int mode,
                     float scale_bound, float lik_bound, uint64_t seed, uint64_t offset, const uint64_t* rng_state,
                     float* y_hat, void* y_hat_bf16, float* lik, float* sum_ln, void* stream);   /* :650-682 */
/* g_lik null -> g = rate_coef / lik (fused rate gradient).  g_yhat nullable.  g_means nullable. */
int lmsfc_gc_backward(const float* y_hat, const float* scales, const float* means, const float* g_lik,
                      const float* g_yhat, const void* g_yhat_bf16, int64_t n, int mode, float scale_bound,
                      float lik_bound, float rate_coef, float* g_y, float* g_scales, float* g_means, void* stream);
/* EntropyBottleneck (:436-516).  x [P][C] fp32 NHWC; params [C][58] packed per channel as
 * M0(3) b0(3) f0(3) | (M(9) b(3) f(3)) x3 | M4(3) b4(1); medians [C] (quantiles[:,0,1]) for mode 1. */
int lmsfc_eb_forward(const float* x, const float* params, const float* medians, const float* noise, int64_t P, int C,
                     int mode, float lik_bound, uint64_t seed, uint64_t offset, const uint64_t* rng_state, float* y_hat,
                     void* y_hat_bf16, float* lik, float* sum_ln, void* stream);
int lmsfc_eb_backward(const float* y_hat, const float* params, const float* g_lik, const float* g_yhat,
                      const void* g_yhat_bf16, int64_t P, int C,
                      int mode, float lik_bound, float rate_coef, float* g_x, float* g_params, void* stream);
/* aux loss (:431-434): loss += sum |logits(quantiles) - target| ; g_quantiles [C][3] written. */
int lmsfc_eb_aux_loss(const float* params, const float* quantiles, const float* target, int C, float* loss,
                      float* g_quantiles, void* stream);
int lmsfc_rng_advance(uint64_t* rng_state, void* stream);

/* ------------------------------------------------------------------ RD loss (losses/rate_distortion.py:47-59) */
/* sse += sum (xhat-x)^2 ; grad (nullable) = grad_coef*(xhat-x), bf16. */
int lmsfc_mse_forward_backward(const void* xhat, const void* x, int64_t n, float grad_coef, float* sse, void* grad,
                               void* stream);
/* The same for up to 8 tensors (pyramid levels) in one launch: HOST arrays of device pointers / sizes / coefficients,
 * sse[l] += ..., grad (nullable array, nullable entries). */
int lmsfc_mse_forward_backward_multi(int nlev, const void* const* xhat, const void* const* x, const int64_t* n,
                                     const float* grad_coef, float* sse, void* const* grad, void* stream);
/* RateDistortionLoss.forward + the gradient w.r.t. every reconstructed level in ONE launch (losses/rate_distortion.py:
 * 47-59): lmsfc_mse_forward_backward_multi with the per-level sums reduced in a fixed order (block partials, last block),
 * and that last block forms out4 = {loss, bpp, D, *extra}: bpp = rate_coef * sum_k *rate[k] (device scalars sum(ln L_k)
 * left by the likelihood kernels, HOST array of n_rate <= 8 pointers), D = sum_l dist_coef[l] * sse_l (host array),
 * loss = lambda_scale * D + bpp; extra (nullable device scalar) rides along, e.g. the step's aux loss.  Nothing has to be
 * zeroed beforehand, and the loss value is bit-reproducible (no float atomics). */
int lmsfc_rd_loss(int nlev, const void* const* xhat, const void* const* x, const int64_t* n, const float* grad_coef,
                  void* const* grad, int n_rate, const float* const* rate, float rate_coef, const float* dist_coef,
                  float lambda_scale, const float* extra, float* out4, void* stream);
/* out[0]=bpp=sum_{i<n_rate} accum[i]*coef[i]; out[1]=D=sum over the next n_dist; out[2]=lambda_scale*D+bpp */
int lmsfc_loss_finalize(const float* accum, const float* coef, int n_rate, int n_dist, float lambda_scale, float* out,
                        void* stream);

/* ------------------------------------------------------------------ small gradients / layout */
int lmsfc_relu_backward(void* g_bf16_inplace, const void* out_bf16, int64_t n, void* stream);
/* nn.LeakyReLU(slope)'s gradient from its output (models/google.py:377-391): g *= out > 0 ? 1 : slope, 0 <= slope < 1 */
int lmsfc_leaky_relu_backward(void* g_bf16_inplace, const void* out_bf16, int64_t n, float slope, void* stream);
/* gradient of a latent head that returned (fp32 y, bf16 |y|): g = [relu & 1 ? (y>0) : 1]*g_direct + sign(y)*g_abs;
 * relu & 2: the head returned (fp32 y, bf16 y) (mean-scale hyperprior: h_a reads y itself, models/google.py:395): the
 * sign factor is 1 */
int lmsfc_abs_backward(const float* g_direct, const void* g_abs_bf16, const float* y, int64_t n, int relu,
                       void* g_out_bf16, void* stream);
int lmsfc_colsum(const void* g_bf16, int64_t P, int C, float* out, void* stream);
/* lmsfc_colsum with NO allocation inside.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and result as lmsfc_colsum (out[c] += sum_p g[p][c], g bf16 [P][C],
 * C a multiple of 8, out 16-byte aligned); the partial sums and arrival counters live in a file-scope device array of
 * 296 x LMSFC_COLSUM_STATIC_MAX_C floats (1.16 MiB) + 256 B, held by every GPU that loads the library until exit.
 * C > LMSFC_COLSUM_STATIC_MAX_C is rejected (return 1, message in lmsfc_last_error()); P is unlimited.  The call makes
 * no cudaMalloc / cudaFree / allocating library call: one cudaMemsetAsync of the counters and one kernel on `stream`.
 * All calls on a device share the one array: calls that overlap on the GPU (different streams without ordering) race;
 * calls ordered by a stream or events do not. */
#define LMSFC_COLSUM_STATIC_MAX_C 1024
int lmsfc_colsum_static(const void* g_bf16, int64_t P, int C, float* out, void* stream);
int lmsfc_cast_f32_bf16(const float* in, void* out, int64_t n, void* stream);
int lmsfc_cast_bf16_f32(const void* in, float* out, int64_t n, void* stream);
int lmsfc_nchw_to_nhwc_bf16(const float* in, void* out, int B, int C, int HW, void* stream);

/* ------------------------------------------------------------------ optimiser (examples/train.py:100-107) */
/* accum += sum(g^2), deterministic (fixed-order two-stage reduction); set_flag (nullable device int) = 1 afterwards */
int lmsfc_sumsq(const float* g, int64_t n, float* accum, int* set_flag, void* stream);
/* The same in slices, so that the norm overlaps the backward pass instead of trailing it: *out = sum(g[0:n]^2) of one
 * arena slice (overwritten), launched as soon as the slice's gradients are complete; then out = sum of the parts in index
 * order (set_flag as above). */
int lmsfc_sumsq_part(const float* g, int64_t n, float* out, void* stream);
int lmsfc_sum_parts(const float* parts, int n, float* out, int* set_flag, void* stream);
/* the closing slice: parts[index] = sum(g[0:n]^2) and, in the same launch, *out = parts[0] + ... + parts[nparts-1] in
 * index order (the other parts are complete on `stream`); saves the separate lmsfc_sum_parts launch on the step's tail */
int lmsfc_sumsq_part_final(const float* g, int64_t n, float* parts, int index, int nparts, float* out, int* set_flag,
                           void* stream);
/* clip_grad_norm_(max_norm) + torch.optim.Adam step over a flat fp32 arena; p_bf16 (nullable) gets
 * the bf16 shadow; step (device int) is incremented; scratch (>=8 floats) receives
 * {clip_coef, bias_corr1, bias_corr2, grad_norm, lr, active, -, -}.  sumsq null or max_norm<=0 -> no clipping. */
int lmsfc_adam_step(float* p, float* m, float* v, const float* g, void* p_bf16, int64_t n, float lr, float beta1,
                    float beta2, float eps, float max_norm, const float* sumsq, int* step, float* scratch,
                    void* stream);

/* The same step in two parts, so that the element-wise pass can run slice by slice (in the order the next forward
 * pass needs the weights) on a stream of its own: adam_prep bumps `step` and fills scratch from sumsq; adam_update
 * applies it to one 16-byte aligned slice and, with zero_g, clears the slice's gradient in the same pass.
 * lr_dev (nullable device float) overrides lr: a captured CUDA graph then follows torch.optim.lr_scheduler
 * (examples/train.py:279).  pending (nullable device int): the update is applied only when *pending != 0, and the flag is
 * cleared (set again by lmsfc_sumsq at the end of the next backward pass); force != 0: apply regardless and clear the
 * flag (an eager step); gate (nullable): take the decision from another optimiser's scratch[5] instead. */
int lmsfc_adam_prep(float lr, const float* lr_dev, float beta1, float beta2, float max_norm, const float* sumsq, int* step,
                    float* scratch, int* pending, int force, const float* gate, void* stream);
int lmsfc_adam_update(float* p, float* m, float* v, float* g, void* p_bf16, int64_t n, float beta1, float beta2, float eps,
                      const float* scratch, int zero_g, void* stream);

/* ------------------------------------------------------------------ the whole step as one replayable launch
 * Replaces the per-operation eager dispatch of one training iteration (examples/train.py:85-118: every module call is
 * a separate host-side launch) by ONE launch: every lmsfc_* call made between capture_begin and capture_end on `stream`
 * (and on streams forked from / joined to it through events) is recorded instead of executed -- all entry points above are
 * capturable: no allocation, no host synchronisation (lmsfc_xchg_init and the first call of the cudaMalloc'ing reductions
 * must have happened before) -- and lmsfc_step_launch replays the recorded step on the same buffers; the host only
 * rewrites the input buffers' contents between replays.  The graph keeps no reference to host memory; the caller keeps
 * every device buffer alive until lmsfc_step_destroy.  (train.TrainStep captures the same calls with torch.cuda.CUDAGraph
 * because torch's caching allocator has to follow the capture; a host without torch uses these.) */
typedef struct lmsfc_step_graph lmsfc_step_graph;
int lmsfc_step_capture_begin(void* stream);
/* ends the capture and instantiates it; *kernels (nullable) = number of kernel nodes recorded */
int lmsfc_step_capture_end(void* stream, lmsfc_step_graph** out, int* kernels);
int lmsfc_step_launch(lmsfc_step_graph* g, void* stream);
int lmsfc_step_destroy(lmsfc_step_graph* g);

/* ------------------------------------------------------------------ data-parallel exchange over NVLink peer memory
 * Replaces nn.DataParallel (examples/train.py:64-71, 275-276) + clip_grad_norm_ + Adam (:100-103) for N <= 8 ranks of one
 * node, one process per GPU, with this library's own kernels on peer-mapped buffers (csrc/exchange.cu): each rank OWNS
 * 1/N of every arena slice, pulls and averages that part of every rank's gradient (reduce-scatter by P2P loads in fixed
 * rank order), the N partial squared norms visit every rank (identical clip coefficient everywhere), Adam runs on the
 * owned part only and the new operands are stored into every rank's arena (all-gather by P2P stores): bf16 for the GEMM
 * weights, fp32 as well for [0, fp32_end) (GDN, entropy bottleneck, biases).
 * All pointer arrays hold the SAME buffer of every rank, peer-mapped into this process (e.g. torch symmetric memory):
 * grads / params fp32 arenas, shadow bf16 arena, flags uint64 [LMSFC_XCHG_SLOTS][LMSFC_MAX_RANKS] (zero-initialised),
 * norm_all float [LMSFC_MAX_RANKS]; tick: a local device uint64 (zero-initialised), the step epoch.
 * Call order per step, identical on all ranks (flag slots are the caller's choice, distinct per call site):
 *   xchg_reduce(slice) ... for every slice, after the rank's own gradients of the slice are complete on `stream`;
 *   xchg_norm_publish(all the slices' norm parts); xchg_norm_wait -> sumsq; lmsfc_adam_prep(sumsq);
 *   xchg_adam_push(slice) ... ; xchg_wait(slot of a push) before the first kernel that reads the slice's parameters. */
#define LMSFC_MAX_RANKS 8
#define LMSFC_XCHG_SLOTS 32
typedef struct {
  int world, rank;
  float* grads[LMSFC_MAX_RANKS];
  float* params[LMSFC_MAX_RANKS];
  void* shadow[LMSFC_MAX_RANKS];
  uint64_t* flags[LMSFC_MAX_RANKS];
  float* norm_all[LMSFC_MAX_RANKS];
  uint64_t* tick;
  /* optional NVSwitch multicast addresses of the same three arenas (NULL: plain peer loads / stores): the reduction
   * then happens inside the switch (multimem.ld_reduce: 1/N of the inbound traffic) and one store reaches every rank
   * (multimem.st) */
  float* mc_grads;
  float* mc_params;
  void* mc_shadow;
} lmsfc_xchg_peers;
/* loads the exchange kernels and their scratch (call once before the first exchange; not capturable) */
int lmsfc_xchg_init(void);
/* grads[rank][lo:hi) = mean over ranks (the caller owns [lo, hi), multiples of 4); *norm_part = its squared norm */
int lmsfc_xchg_reduce(const lmsfc_xchg_peers* peers, int slot, int64_t lo, int64_t hi, float* norm_part, int max_ctas,
                      void* stream);
/* set_flag (nullable device int) = 1 afterwards: the pending flag lmsfc_adam_prep consumes */
int lmsfc_xchg_norm_publish(const lmsfc_xchg_peers* peers, const float* norm_parts, int nparts, int slot, int* set_flag,
                            void* stream);
int lmsfc_xchg_norm_wait(const lmsfc_xchg_peers* peers, int slot, float* sumsq_out, void* stream);
/* Adam (scratch from lmsfc_adam_prep) on [own_lo, own_hi) inside the slice [unit_lo, unit_hi); clears the whole slice's
 * local gradient; signals `slot` on every rank when this rank's part has landed everywhere */
int lmsfc_xchg_adam_push(const lmsfc_xchg_peers* peers, int slot, int64_t unit_lo, int64_t unit_hi, int64_t own_lo,
                         int64_t own_hi, int64_t fp32_end, float* m, float* v, float beta1, float beta2, float eps,
                         const float* scratch, int max_ctas, void* stream);
int lmsfc_xchg_wait(const lmsfc_xchg_peers* peers, int slot, void* stream);

/* ------------------------------------------------------------------ CDF tables (cpp_exts/ops/ops.cpp:40-109) */
/* host pointers; cdf must hold n+1 entries.  Errors mirror the reference's ValueError cases. */
int lmsfc_pmf_to_quantized_cdf(const float* pmf_host, int n, int precision, uint32_t* cdf_host);
/* device: row r = pmf[r][0:pmf_length[r]] ++ tail[r] -> cdf[r][0:pmf_length[r]+2], zero padded
 * (entropy_models.py:204-212).  status[r]: 0 ok, 1 bad element, 2 all-zero, 3 nothing to steal. */
int lmsfc_pmf_rows_to_cdf(const float* pmf, int pmf_stride, const float* tail, const int* pmf_length, int rows,
                          int precision, int32_t* cdf, int cdf_stride, int* status, void* stream);

/* The float pmf grids in front of the integer stage, on the device (per-frame table refresh):
 * EntropyBottleneck.update's (entropy_models.py:412-424): pmf[c][j] for sample j + pmf_start[c], j < max_length, from
 * the packed parameters; tail[c] as :423.  GaussianConditional.update's (:632-642): pmf[s][j] for |j - pmf_center[s]|
 * under scale_table[s]; tail[s] = 2 * lower[0].  fp32 with precise libdevice functions: within a few ulp of the
 * reference's CPU evaluation, i.e. tables equal up to single counts (the reference tolerates +-2 on rebuilt tables). */
int lmsfc_eb_pmf(const float* params, const float* pmf_start, int C, int max_length, float* pmf, int pmf_stride,
                 float* tail, void* stream);
int lmsfc_gc_pmf(const float* scale_table, const int32_t* pmf_center, int S, int max_length, float* pmf, int pmf_stride,
                 float* tail, void* stream);

/* ------------------------------------------------------------------ range-ANS entropy coding
 * Replaces the pybind11 module `compressai.ans` (cpp_exts/rans/rans_interface.cpp:108-288, ryg rans64.h) behind
 * EntropyModel.compress / decompress (entropy_models/entropy_models.py:235-325).  symbols / indexes are int32
 * [items][n] (an item flattened in NCHW order, like `symbols[i].reshape(-1)`); cdfs int32 [rows][cdf_stride],
 * sizes = `_cdf_length`, offsets = `_offset`.  Lane k of `lanes` codes symbols[item][k::lanes] as one reference
 * string (lanes = 1: the reference's bytes).  16-bit probabilities, 4-bit bypass digits, as in the reference.
 *
 * encode: scratch holds items*lanes*lane_cap words (lane_cap >= lmsfc_rans_lane_cap(n, lanes)), lane_words
 * [items*lanes], lane_offsets [items*lanes + 1] (word offset of every lane in `packed`, last = total words),
 * packed [packed_cap] words; status (device int, caller-zeroed): 1 index outside the tables, 2 out of space. */
int64_t lmsfc_rans_lane_cap(int64_t n, int lanes);
int lmsfc_rans_encode(const int32_t* symbols, const int32_t* indexes, int items, int64_t n, int lanes,
                      const int32_t* cdfs, int rows, int cdf_stride, const int32_t* sizes, const int32_t* offsets,
                      uint32_t* scratch, int64_t lane_cap, uint32_t* lane_words, uint32_t* lane_offsets,
                      uint32_t* packed, int64_t packed_cap, int* status, void* stream);
/* decode: packed [packed_words] = the lanes' strings back to back, lane_offsets [items*lanes] their word offsets. */
int lmsfc_rans_decode(const uint32_t* packed, int64_t packed_words, const uint32_t* lane_offsets,
                      const int32_t* indexes, int items, int64_t n, int lanes, const int32_t* cdfs, int rows,
                      int cdf_stride, const int32_t* sizes, const int32_t* offsets, int32_t* symbols_out, int* status,
                      void* stream);
/* The same coder with DECODER CHECKPOINTS, so that ONE reference-format string decodes in parallel (a string is a serial
 * chain on the coder state: rans_interface.cpp:215-288 decodes it symbol by symbol).  While coding, the encoder notes, at
 * every `ckpt_interval`-th symbol of a lane (a power of two >= 128), the coder state and the number of words written so
 * far: ckpt[string][c - 1] = {state low, state high, words written} for symbol c * ckpt_interval, c = 1 .. ncp, ncp =
 * lmsfc_rans_checkpoints(n, lanes, ckpt_interval).  That state is exactly the decoder's state in front of that symbol,
 * so lmsfc_rans_decode_indexed starts one warp per (string, segment).  The strings themselves do not change -- they stay
 * the reference's bytes, decodable by the stock decoder; the table travels beside them (coder.py appends it behind the
 * string, where the stock decoder never looks).  For decode the caller rewrites the third word of every entry to the
 * absolute word position in `packed` where the segment starts: lane_offsets[string] + words of the string - words
 * written.  status 3: a segment did not end in the next checkpoint's state (corrupt table or string).
 * ckpt_interval = 0: no checkpoints, identical to lmsfc_rans_encode / lmsfc_rans_decode. */
int lmsfc_rans_checkpoints(int64_t n, int lanes, int ckpt_interval);
int lmsfc_rans_encode_indexed(const int32_t* symbols, const int32_t* indexes, int items, int64_t n, int lanes,
                              const int32_t* cdfs, int rows, int cdf_stride, const int32_t* sizes, const int32_t* offsets,
                              uint32_t* scratch, int64_t lane_cap, uint32_t* lane_words, uint32_t* lane_offsets,
                              uint32_t* packed, int64_t packed_cap, int* status, int ckpt_interval, int ncp, uint32_t* ckpt,
                              void* stream);
int lmsfc_rans_decode_indexed(const uint32_t* packed, int64_t packed_words, const uint32_t* lane_offsets,
                              const int32_t* indexes, int items, int64_t n, int lanes, const int32_t* cdfs, int rows,
                              int cdf_stride, const int32_t* sizes, const int32_t* offsets, int32_t* symbols_out, int* status,
                              int ckpt_interval, int ncp, const uint32_t* ckpt, void* stream);
/* GaussianConditional.build_indexes (entropy_models.py:684-689) */
int lmsfc_gc_build_indexes(const float* scales, int64_t n, const float* scale_table, int table_len, float scale_bound,
                           int32_t* indexes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200LMSFC_H */
