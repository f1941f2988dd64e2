// Internal launch interfaces of the memory-bound kernels (entropy.cu, elementwise.cu, cdf.cu).
#pragma once
#include "common.cuh"

namespace lmsfc {

enum : int { QUANT_NOISE = 0, QUANT_DEQUANTIZE = 1, QUANT_GIVEN = 2 };

struct GcArgs {
  const float *y, *scales, *means, *noise;   // means / noise nullable
  float *y_hat, *lik, *sum_ln;               // sum_ln nullable (atomicAdd of sum ln L)
  __nv_bfloat16* y_hat_bf16;                 // nullable: bf16 copy of y_hat (operand of the next conv)
  int64_t n;
  int mode;
  float scale_bound, lik_bound;
  uint64_t seed, offset;
  const uint64_t* rng_state;                 // device {seed, step}; nullable
};
struct GcBwdArgs {
  const float *y_hat, *scales, *means, *g_lik, *g_yhat;   // means / g_lik / g_yhat nullable
  const __nv_bfloat16* g_yhat_bf16;                       // alternative to g_yhat (dgrad output), nullable
  float *g_y, *g_scales, *g_means;                        // g_means nullable
  int64_t n;
  int mode;
  float scale_bound, lik_bound, rate_coef;                // g_lik == null -> g = rate_coef / lik
};
struct EbArgs {
  const float *x, *params, *medians, *noise;  // x: [P][C] (NHWC); params: [C][58]
  float *y_hat, *lik, *sum_ln;
  __nv_bfloat16* y_hat_bf16;                  // nullable
  int64_t P;
  int C, mode;
  float lik_bound;
  uint64_t seed, offset;
  const uint64_t* rng_state;
};
struct EbBwdArgs {
  const float *y_hat, *params, *g_lik, *g_yhat;
  const __nv_bfloat16* g_yhat_bf16;           // alternative to g_yhat, nullable
  float *g_x, *g_params;                      // g_params [C][58], accumulated with atomicAdd
  int64_t P;
  int C, mode;
  float lik_bound, rate_coef;
};

int gc_forward(const GcArgs& a, cudaStream_t s);
int gc_backward(const GcBwdArgs& a, cudaStream_t s);
int eb_forward(const EbArgs& a, cudaStream_t s);
int eb_backward(const EbBwdArgs& a, cudaStream_t s);
int eb_aux(const float* params, const float* quantiles, const float* target, int C, float* loss, float* g_quantiles,
           cudaStream_t s);
int rng_advance(uint64_t* state, cudaStream_t s);
int eb_pmf(const float* params, const float* pmf_start, int C, int max_length, float* pmf, int stride, float* tail,
           cudaStream_t s);
int gc_pmf(const float* table, const int* center, int S, int max_length, float* pmf, int stride, float* tail, cudaStream_t s);

// ---- elementwise.cu
int mse_fwd_bwd(const __nv_bfloat16* xhat, const __nv_bfloat16* x, int64_t n, float grad_coef, float* sse,
                __nv_bfloat16* grad, cudaStream_t s);
constexpr int kMseMaxLevels = 8;
int mse_fwd_bwd_multi(int nlev, const void* const* xhat, const void* const* x, const int64_t* n, const float* grad_coef,
                      float* sse, void* const* grad, cudaStream_t s);
int rd_loss_multi(int nlev, const void* const* xhat, const void* const* x, const int64_t* n, const float* grad_coef,
                  void* const* grad, int n_rate, const float* const* rate, float rate_coef, const float* dist_coef,
                  float lambda_scale, const float* extra, float* out4, cudaStream_t s);
int loss_finalize(const float* accum, const float* coef, int n_rate, int n_dist, float lambda_scale, float* out,
                  cudaStream_t s);
int square_bf16(const __nv_bfloat16* x, __nv_bfloat16* x2, int64_t n, cudaStream_t s);
int gdn_t(const __nv_bfloat16* gy, const __nv_bfloat16* x, const __nv_bfloat16* rn, int inverse, int64_t P, int C,
          __nv_bfloat16* t, float* g_beta, cudaStream_t s);
int reparam_fwd(const float* p, int64_t n, float bound, float pedestal, float* out_f32, __nv_bfloat16* out_bf16,
                cudaStream_t s);
int reparam_bwd(const float* g_out, const float* p, int64_t n, float bound, float* g_p, cudaStream_t s);
int reparam_fwd_multi(const float* params, const int64_t* table, const float* bounds, int nseg, int64_t max_n,
                      float pedestal, float* out_f32, __nv_bfloat16* out_bf16, cudaStream_t s);
int reparam_bwd_multi(const float* g_eff, const float* params, const int64_t* table, const float* bounds, int nseg,
                      int64_t max_n, float* g_params, cudaStream_t s);
int colsum_bf16(const __nv_bfloat16* g, int64_t P, int C, float* out, cudaStream_t s);
int colsum_bf16_static(const __nv_bfloat16* g, int64_t P, int C, float* out, cudaStream_t s);   // no allocation
int relu_bwd_bf16(__nv_bfloat16* g, const __nv_bfloat16* out, int64_t n, float slope, cudaStream_t s);
int abs_bwd(const float* g_direct, const __nv_bfloat16* g_abs, const float* y, int64_t n, int relu, __nv_bfloat16* g_out,
            cudaStream_t s);
int cast_f32_bf16(const float* in, __nv_bfloat16* out, int64_t n, cudaStream_t s);
int cast_bf16_f32(const __nv_bfloat16* in, float* out, int64_t n, cudaStream_t s);
int nchw_to_nhwc_bf16(const float* in, __nv_bfloat16* out, int B, int C, int HW, cudaStream_t s);
int sumsq_f32(const float* g, int64_t n, float* accum, int* set_flag, int overwrite, cudaStream_t s, const float* parts = nullptr,
              int nparts = 0, float* total = nullptr);
int sum_parts(const float* parts, int n, float* out, int* set_flag, cudaStream_t s);
struct AdamArgs {
  float *p, *m, *v;
  float* g;                  // read; zeroed in the same pass when zero_g
  __nv_bfloat16* p_bf16;     // nullable: bf16 shadow written in the same pass
  int64_t n;
  float lr, beta1, beta2, eps, max_norm;   // max_norm <= 0: no clipping
  const float* sumsq;        // device scalar with sum(g^2) (clip_grad_norm_); nullable
  int* step;                 // device step counter, incremented by the prep kernel
  float* scratch;            // >= 8 floats: {clip_coef, bias_corr1, bias_corr2, grad_norm, lr, active, -, -}
  int zero_g;
  const float* lr_dev;       // prep only, nullable: learning rate in device memory (overrides lr)
  int* pending;              // prep only, nullable: "update pending" flag, consumed
  int force;                 // prep only: apply regardless of the flag (and clear it)
  const float* gate;         // prep only, nullable: another optimiser's scratch[5]
};
// reduction scratch shared by the two-stage reductions (elementwise.cu): one slot per stream, 296 x 2048 floats and 64
// self-resetting counters each
int ensure_reduce_scratch();
float* red_scratch(cudaStream_t s);
unsigned int* red_count(cudaStream_t s);
int adam_step(const AdamArgs& a, cudaStream_t s);
int adam_prep(const AdamArgs& a, cudaStream_t s);     // step counter + clip / bias-correction scalars
int adam_update(const AdamArgs& a, cudaStream_t s);   // the element-wise pass, given a prepared scratch

// ---- cdf.cu
int pmf_to_quantized_cdf_host(const float* pmf, int n, int precision, uint32_t* cdf);
int pmf_rows_to_cdf_device(const float* pmf, int pmf_stride, const float* tail, const int* pmf_length, int rows,
                           int precision, int32_t* cdf, int cdf_stride, int* status, cudaStream_t s);

}  // namespace lmsfc
