// Fused, vectorised, warp-shuffle-reduced entropy-model kernels (fp32 arithmetic, HBM-bound).
//
//   GaussianConditional  compressai/entropy_models/entropy_models.py:650-682 (+ quantize :155-180)
//   EntropyBottleneck    entropy_models.py:436-516 ; aux loss :431-434
//   LowerBound gradient  compressai/ops/bound_ops.py:40-42
//
// Every forward kernel fuses quantise + likelihood + lower bound + the sum(ln L) reduction of
// RateDistortionLoss (losses/rate_distortion.py:52-55); every backward kernel recomputes the
// likelihood analytically instead of reading saved intermediates.
#include "kernels.cuh"

namespace lmsfc {

// ---------------------------------------------------------------- Philox4x32-10
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}
// uniform in [-0.5, 0.5) for element idx of draw `offset` under `seed`
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t offset, uint64_t idx) {
  uint32_t c[4] = {(uint32_t)(idx >> 2), (uint32_t)(idx >> 34), (uint32_t)offset, (uint32_t)(offset >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return (float)(c[idx & 3] >> 8) * (1.0f / 16777216.0f) - 0.5f;
}

__global__ void rng_advance_kernel(uint64_t* state) {
  pdl_sync(); state[1] += 1; }

int rng_advance(uint64_t* state, cudaStream_t s) {
  LMSFC_LAUNCH(rng_advance_kernel, dim3(1), dim3(1), 0, s, state);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- GaussianConditional
__device__ __forceinline__ float std_cumulative(float x) {  // entropy_models.py:604-608
  return 0.5f * erfcf(-0.70710678118654752440f * x);
}
__device__ __forceinline__ float normal_pdf(float t) { return 0.3989422804014327f * __expf(-0.5f * t * t); }

// all four uniforms of one Philox block (elements 4q .. 4q+3 of draw `offset`): the same values philox_uniform returns
__device__ __forceinline__ void philox_uniform4(uint64_t seed, uint64_t offset, uint64_t q, float (&u)[4]) {
  uint32_t c[4] = {(uint32_t)q, (uint32_t)(q >> 32), (uint32_t)offset, (uint32_t)(offset >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) u[j] = (float)(c[j] >> 8) * (1.0f / 16777216.0f) - 0.5f;
}

__device__ __forceinline__ float gc_one(const GcArgs& a, float y, float mu, float u, float scale, bool has_scale, float& yh) {
  if (a.mode == QUANT_NOISE) yh = y + u;                       // entropy_models.py:161-165
  else if (a.mode == QUANT_DEQUANTIZE) yh = rintf(y - mu) + mu;   // :167-176 (torch.round = half to even)
  else yh = y;                                                 // QUANT_GIVEN: the caller quantised already
  if (!has_scale) return 1.f;
  const float v = fabsf(yh - mu);
  const float s = fmaxf(scale, a.scale_bound);
  const float L = std_cumulative((0.5f - v) / s) - std_cumulative((-0.5f - v) / s);
  return fmaxf(L, a.lik_bound);
}

// Four consecutive elements per thread: one Philox block serves all four (each thread used to recompute the block for
// one value), 16-byte loads and stores.  Instruction-bound (two erfc per element), not HBM-bound.
__global__ void __launch_bounds__(256) gc_forward_kernel(GcArgs a) {
  pdl_sync();
  float local = 0.f;
  // key = seed perturbed per module (`offset`), counter = device step count
  const uint64_t seed = (a.rng_state ? a.rng_state[0] : a.seed) + a.offset * 0x9E3779B97F4A7C15ull;
  const uint64_t off = a.rng_state ? a.rng_state[1] : 0;
  const int64_t n4 = a.n >> 2;
  const bool draw = a.mode == QUANT_NOISE && !a.noise;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (int64_t)gridDim.x * blockDim.x) {
    const float4 y4 = reinterpret_cast<const float4*>(a.y)[q];
    const float4 m4 = a.means ? reinterpret_cast<const float4*>(a.means)[q] : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 s4 = a.scales ? reinterpret_cast<const float4*>(a.scales)[q] : make_float4(1.f, 1.f, 1.f, 1.f);
    float u[4] = {0.f, 0.f, 0.f, 0.f};
    if (draw) {
      philox_uniform4(seed, off, (uint64_t)q, u);
    } else if (a.noise) {
      const float4 t = reinterpret_cast<const float4*>(a.noise)[q];
      u[0] = t.x; u[1] = t.y; u[2] = t.z; u[3] = t.w;
    }
    const float yy[4] = {y4.x, y4.y, y4.z, y4.w}, mm[4] = {m4.x, m4.y, m4.z, m4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
    float yh[4], lk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) lk[j] = gc_one(a, yy[j], mm[j], u[j], ss[j], a.scales != nullptr, yh[j]);
    if (a.y_hat) reinterpret_cast<float4*>(a.y_hat)[q] = make_float4(yh[0], yh[1], yh[2], yh[3]);
    if (a.y_hat_bf16) reinterpret_cast<uint2*>(a.y_hat_bf16)[q] = make_uint2(pack_bf16(yh[0], yh[1]), pack_bf16(yh[2], yh[3]));
    if (a.scales) {
      reinterpret_cast<float4*>(a.lik)[q] = make_float4(lk[0], lk[1], lk[2], lk[3]);
      local += __logf(lk[0]) + __logf(lk[1]) + __logf(lk[2]) + __logf(lk[3]);
    }
  }
  if (blockIdx.x == 0 && (int64_t)threadIdx.x < a.n - n4 * 4) {          // scalar tail (n not a multiple of 4)
    const int64_t i = n4 * 4 + threadIdx.x;
    const float mu = a.means ? a.means[i] : 0.f;
    const float u = a.mode == QUANT_NOISE ? (a.noise ? a.noise[i] : philox_uniform(seed, off, (uint64_t)i)) : 0.f;
    float yh;
    const float lk = gc_one(a, a.y[i], mu, u, a.scales ? a.scales[i] : 1.f, a.scales != nullptr, yh);
    if (a.y_hat) a.y_hat[i] = yh;
    if (a.y_hat_bf16) a.y_hat_bf16[i] = __float2bfloat16(yh);
    if (a.scales) { a.lik[i] = lk; local += __logf(lk); }
  }
  if (a.sum_ln) {
    __shared__ float red[32];
    local = block_sum(local, red);
    if (threadIdx.x == 0) atomicAdd(a.sum_ln, local);
  }
}

__global__ void __launch_bounds__(256) gc_backward_kernel(GcBwdArgs a) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * blockDim.x) {
    const float yh = a.y_hat[i], mu = a.means ? a.means[i] : 0.f;
    const float d = yh - mu, v = fabsf(d);
    const float sraw = a.scales[i];
    const float s = fmaxf(sraw, a.scale_bound);
    const float ta = (0.5f - v) / s, tb = (-0.5f - v) / s;
    const float L = std_cumulative(ta) - std_cumulative(tb);
    float g = a.g_lik ? a.g_lik[i] : a.rate_coef / fmaxf(L, a.lik_bound);
    if (!(L >= a.lik_bound || g < 0.f)) g = 0.f;   // bound_ops.py:40-42
    const float pa = normal_pdf(ta), pb = normal_pdf(tb);
    const float gv = g * (pb - pa) / s;
    float gs = g * (tb * pb - ta * pa) / s;
    if (!(sraw >= a.scale_bound || gs < 0.f)) gs = 0.f;
    const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
    const float gyh = a.g_yhat ? a.g_yhat[i] : (a.g_yhat_bf16 ? __bfloat162float(a.g_yhat_bf16[i]) : 0.f);
    if (a.mode != QUANT_DEQUANTIZE) {              // noise, or an already quantised input: identity path
      a.g_y[i] = gyh + gv * sgn;
      if (a.g_means) a.g_means[i] = -gv * sgn;
    } else {
      a.g_y[i] = 0.f;
      if (a.g_means) a.g_means[i] = gyh;
    }
    a.g_scales[i] = gs;
  }
}

static int grid_for(int64_t n, int threads) {
  int64_t b = (n + threads - 1) / threads;
  const int64_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

int gc_forward(const GcArgs& a, cudaStream_t s) {
  if (a.n <= 0) return LMSFC_OK;
  LMSFC_CHECK_ARG((((uintptr_t)a.y | (uintptr_t)a.scales | (uintptr_t)a.means | (uintptr_t)a.noise | (uintptr_t)a.y_hat |
                    (uintptr_t)a.lik) & 15) == 0 && ((uintptr_t)a.y_hat_bf16 & 7) == 0, "gc_forward: operands must be 16-byte aligned");
  LMSFC_LAUNCH(gc_forward_kernel, dim3(grid_for((a.n + 3) / 4, 256)), dim3(256), 0, s, a);
  return LMSFC_OK;
}
int gc_backward(const GcBwdArgs& a, cudaStream_t s) {
  if (a.n <= 0) return LMSFC_OK;
  LMSFC_LAUNCH(gc_backward_kernel, dim3(grid_for(a.n, 256)), dim3(256), 0, s, a);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- EntropyBottleneck
// Packed per-channel parameters, 58 floats (filters (3,3,3,3), entropy_models.py:355-377):
//   [M0(3) b0(3) f0(3)] [M1(9) b1(3) f1(3)] [M2(9) b2(3) f2(3)] [M3(9) b3(3) f3(3)] [M4(3) b4(1)]
constexpr int EB_NP = 58;

__device__ __forceinline__ float softplus_t(float x) { return x > 20.f ? x : log1pf(__expf(x)); }  // torch threshold 20
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + __expf(-x)); }

// tanh of a pre-activation: 1 - 2 / (e^{2a} + 1) with ex2.approx / rcp.approx (absolute error ~1e-7, saturates to +-1).
// The precise tanhf is ~25 instructions with branches and there are 24 of them per element: the likelihood kernels
// were bound by them (0.08 of the HBM roofline at 8 M elements).  Per-channel constants (tanh(factor)) stay precise.
__device__ __forceinline__ float tanh_fast(float x) { return 1.f - __fdividef(2.f, __expf(2.f * x) + 1.f); }

struct EbChan {
  float sp0[3], b0[3], tf0[3];
  float sp[3][9], b[3][3], tf[3][3];
  float sp4[3], b4;
};

__device__ __forceinline__ void eb_load(const float* __restrict__ q, EbChan& c) {
#pragma unroll
  for (int j = 0; j < 3; ++j) { c.sp0[j] = softplus_t(q[j]); c.b0[j] = q[3 + j]; c.tf0[j] = tanhf(q[6 + j]); }
#pragma unroll
  for (int l = 0; l < 3; ++l) {
    const float* r = q + 9 + 15 * l;
#pragma unroll
    for (int j = 0; j < 9; ++j) c.sp[l][j] = softplus_t(r[j]);
#pragma unroll
    for (int j = 0; j < 3; ++j) { c.b[l][j] = r[9 + j]; c.tf[l][j] = tanhf(r[12 + j]); }
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) c.sp4[j] = softplus_t(q[54 + j]);
  c.b4 = q[57];
}

// logits_cumulative (entropy_models.py:436-455) for one scalar; optionally keeps the activations
struct EbTape { float ta[4][3], h[4][3]; };   // tanh(pre-activation), gated output per layer

template <bool TAPE>
__device__ __forceinline__ float eb_logits(const EbChan& c, float v, EbTape* tp) {
  float h[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float a = c.sp0[j] * v + c.b0[j];
    const float t = tanh_fast(a);
    h[j] = a + c.tf0[j] * t;
    if (TAPE) { tp->ta[0][j] = t; tp->h[0][j] = h[j]; }
  }
#pragma unroll
  for (int l = 0; l < 3; ++l) {
    float n[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float a = c.sp[l][3 * i] * h[0] + c.sp[l][3 * i + 1] * h[1] + c.sp[l][3 * i + 2] * h[2] + c.b[l][i];
      const float t = tanh_fast(a);
      n[i] = a + c.tf[l][i] * t;
      if (TAPE) { tp->ta[l + 1][i] = t; tp->h[l + 1][i] = n[i]; }
    }
    h[0] = n[0]; h[1] = n[1]; h[2] = n[2];
  }
  return c.sp4[0] * h[0] + c.sp4[1] * h[1] + c.sp4[2] * h[2] + c.b4;
}

__device__ __forceinline__ float eb_lik_raw(float lower, float upper) {  // entropy_models.py:458-469
  const float t = lower + upper;
  const float sg = (t > 0.f) ? -1.f : ((t < 0.f) ? 1.f : 0.f);
  return fabsf(sigmoid_f(sg * upper) - sigmoid_f(sg * lower));
}

// x is NHWC-flattened [P][C]; thread <-> channel, grid.y strides over pixels.
__global__ void __launch_bounds__(128) eb_forward_kernel(EbArgs a) {
  pdl_sync();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  float local = 0.f;
  if (c < a.C) {
    EbChan ch;
    eb_load(a.params + (size_t)c * EB_NP, ch);
    const float med = a.medians ? a.medians[c] : 0.f;
    // key = seed perturbed per module (`offset`), counter = device step count
    const uint64_t seed = (a.rng_state ? a.rng_state[0] : a.seed) + a.offset * 0x9E3779B97F4A7C15ull;
    const uint64_t off = a.rng_state ? a.rng_state[1] : 0;
    for (int64_t p = blockIdx.y; p < a.P; p += gridDim.y) {
      const int64_t i = p * a.C + c;
      const float x = a.x[i];
      float yh;
      if (a.mode == QUANT_NOISE) {
        const float u = a.noise ? a.noise[i] : philox_uniform(seed, off, (uint64_t)i);
        yh = x + u;
      } else {
        yh = rintf(x - med) + med;
      }
      const float lo = eb_logits<false>(ch, yh - 0.5f, nullptr);
      const float up = eb_logits<false>(ch, yh + 0.5f, nullptr);
      const float lik = fmaxf(eb_lik_raw(lo, up), a.lik_bound);
      a.y_hat[i] = yh;
      if (a.y_hat_bf16) a.y_hat_bf16[i] = __float2bfloat16(yh);
      a.lik[i] = lik;
      local += __logf(lik);
    }
  }
  if (a.sum_ln) {
    __shared__ float red[32];
    local = block_sum(local, red);
    if (threadIdx.x == 0) atomicAdd(a.sum_ln, local);
  }
}

// reverse pass through one logits evaluation; accumulates into the "effective" parameter
// gradients (w.r.t. softplus(M), b, tanh(f)); returns d logits / d v times g.
__device__ __forceinline__ float eb_logits_bwd(const EbChan& c, const EbTape& tp, float v, float g, float* G) {
  // G layout mirrors the packed parameters (58)
  float gh[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) { G[54 + j] += g * tp.h[3][j]; gh[j] = g * c.sp4[j]; }
  G[57] += g;
#pragma unroll
  for (int l = 2; l >= 0; --l) {
    float ga[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float t = tp.ta[l + 1][i];
      G[9 + 15 * l + 12 + i] += gh[i] * t;
      ga[i] = gh[i] * (1.f + c.tf[l][i] * (1.f - t * t));
      G[9 + 15 * l + 9 + i] += ga[i];
#pragma unroll
      for (int j = 0; j < 3; ++j) G[9 + 15 * l + 3 * i + j] += ga[i] * tp.h[l][j];
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) gh[j] = ga[0] * c.sp[l][j] + ga[1] * c.sp[l][3 + j] + ga[2] * c.sp[l][6 + j];
  }
  float gv = 0.f;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float t = tp.ta[0][j];
    G[6 + j] += gh[j] * t;
    const float ga = gh[j] * (1.f + c.tf0[j] * (1.f - t * t));
    G[3 + j] += ga;
    G[j] += ga * v;
    gv += ga * c.sp0[j];
  }
  return gv;
}

__global__ void __launch_bounds__(128) eb_backward_kernel(EbBwdArgs a) {
  pdl_sync();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= a.C) return;
  EbChan ch;
  const float* q = a.params + (size_t)c * EB_NP;
  eb_load(q, ch);
  float G[EB_NP];
#pragma unroll
  for (int j = 0; j < EB_NP; ++j) G[j] = 0.f;
  for (int64_t p = blockIdx.y; p < a.P; p += gridDim.y) {
    const int64_t i = p * a.C + c;
    const float yh = a.y_hat[i];
    EbTape tl, tu;
    const float lo = eb_logits<true>(ch, yh - 0.5f, &tl);
    const float up = eb_logits<true>(ch, yh + 0.5f, &tu);
    const float t = lo + up;
    const float sg = (t > 0.f) ? -1.f : ((t < 0.f) ? 1.f : 0.f);
    const float su = sigmoid_f(sg * up), sl = sigmoid_f(sg * lo);
    const float D = su - sl, L = fabsf(D);
    float g = a.g_lik ? a.g_lik[i] : a.rate_coef / fmaxf(L, a.lik_bound);
    if (!(L >= a.lik_bound || g < 0.f)) g = 0.f;
    const float sd = (D > 0.f) ? 1.f : ((D < 0.f) ? -1.f : 0.f);
    const float gu = g * sd * su * (1.f - su) * sg;
    const float gl = -g * sd * sl * (1.f - sl) * sg;
    float gx = eb_logits_bwd(ch, tl, yh - 0.5f, gl, G);
    gx += eb_logits_bwd(ch, tu, yh + 0.5f, gu, G);
    const float gyh = a.g_yhat ? a.g_yhat[i] : (a.g_yhat_bf16 ? __bfloat162float(a.g_yhat_bf16[i]) : 0.f);
    a.g_x[i] = (a.mode == QUANT_NOISE) ? gx + gyh : 0.f;
  }
  // chain through softplus / tanh of the raw parameters, then one atomic per (channel, param)
  float* out = a.g_params + (size_t)c * EB_NP;
#pragma unroll
  for (int j = 0; j < EB_NP; ++j) {
    const bool is_m = (j < 3) || (j >= 9 && j < 54 && ((j - 9) % 15) < 9) || (j >= 54 && j < 57);
    const bool is_f = (j >= 6 && j < 9) || (j >= 9 && j < 54 && ((j - 9) % 15) >= 12);
    float g = G[j];
    if (is_m) g *= (q[j] > 20.f) ? 1.f : sigmoid_f(q[j]);
    if (is_f) { const float t = tanhf(q[j]); g *= 1.f - t * t; }
    atomicAdd(out + j, g);
  }
}

// aux loss (entropy_models.py:431-434): value and d/d quantiles in one pass; parameters detached.
__global__ void __launch_bounds__(128) eb_aux_kernel(const float* params, const float* quantiles, const float* target,
                                                     int C, float* loss, float* g_quantiles) {
  pdl_sync();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  float local = 0.f;
  if (c < C) {
    EbChan ch;
    eb_load(params + (size_t)c * EB_NP, ch);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float v = quantiles[c * 3 + k];
      EbTape tp;
      const float lg = eb_logits<true>(ch, v, &tp);
      const float d = lg - target[k];
      local += fabsf(d);
      const float sd = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
      float G[EB_NP];
#pragma unroll
      for (int j = 0; j < EB_NP; ++j) G[j] = 0.f;
      g_quantiles[c * 3 + k] = eb_logits_bwd(ch, tp, v, sd, G);
    }
  }
  __shared__ float red[32];
  local = block_sum(local, red);
  if (threadIdx.x == 0) atomicAdd(loss, local);
}

// ---------------------------------------------------------------- float pmf grids for the CDF tables (update())
// EntropyBottleneck.update (entropy_models.py:389-429, pmf at :412-424) and GaussianConditional.update (:625-648, pmf at
// :632-642) evaluate the likelihood on an integer grid per channel / scale before the integer stage (cdf.cu).  Here the
// grid is evaluated on the device in fp32 with the PRECISE libdevice functions (expf, log1pf, tanhf, erfcf: no fast
// intrinsics).  The reference's CPU float ops differ from these by a few ulp, so the 16-bit tables can differ by one
// count where pmf * 2^16 lands next to .5 -- inside the reference's own tolerance for rebuilt tables (+-2 counts,
// tests/test_entropy_models.py:393-398); the default update() keeps the host evaluation that is bit-identical.
__device__ __forceinline__ float softplus_p(float x) { return x > 20.f ? x : log1pf(expf(x)); }
__device__ __forceinline__ float sigmoid_p(float x) { return 1.f / (1.f + expf(-x)); }
__device__ float eb_logits_precise(const float* __restrict__ q, float v) {
  float h[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float a = softplus_p(q[j]) * v + q[3 + j];
    h[j] = a + tanhf(q[6 + j]) * tanhf(a);
  }
#pragma unroll
  for (int l = 0; l < 3; ++l) {
    const float* r = q + 9 + 15 * l;
    float n[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float a = softplus_p(r[3 * i]) * h[0] + softplus_p(r[3 * i + 1]) * h[1] + softplus_p(r[3 * i + 2]) * h[2] + r[9 + i];
      n[i] = a + tanhf(r[12 + i]) * tanhf(a);
    }
    h[0] = n[0]; h[1] = n[1]; h[2] = n[2];
  }
  return softplus_p(q[54]) * h[0] + softplus_p(q[55]) * h[1] + softplus_p(q[56]) * h[2] + q[57];
}
__global__ void __launch_bounds__(128) eb_pmf_kernel(const float* __restrict__ params, const float* __restrict__ pmf_start,
                                                     int max_length, float* __restrict__ pmf, int stride,
                                                     float* __restrict__ tail) {
  pdl_sync();
  const int c = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= max_length) return;
  const float* q = params + (size_t)c * EB_NP;
  const float sample = (float)j + pmf_start[c];                      // arange(max_length) + pmf_start, :408-409
  const float lower = eb_logits_precise(q, sample - 0.5f), upper = eb_logits_precise(q, sample + 0.5f);
  const float t = lower + upper;
  const float sg = (t > 0.f) ? -1.f : ((t < 0.f) ? 1.f : 0.f);
  pmf[(size_t)c * stride + j] = fabsf(sigmoid_p(sg * upper) - sigmoid_p(sg * lower));
  if (j == 0) {                                                      // tail mass: first lower + last upper of the grid, :423
    const float last = (float)(max_length - 1) + pmf_start[c];
    tail[c] = sigmoid_p(lower) + sigmoid_p(-eb_logits_precise(q, last + 0.5f));
  }
}
__global__ void __launch_bounds__(128) gc_pmf_kernel(const float* __restrict__ table, const int* __restrict__ center,
                                                     int max_length, float* __restrict__ pmf, int stride,
                                                     float* __restrict__ tail) {
  pdl_sync();
  const int s = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= max_length) return;
  const float scale = table[s];
  const float v = fabsf((float)(j - center[s]));                     // |arange(max_length) - pmf_center|, :632-634
  const float c = -0.70710678118654752440f;                          // -(2 ** -0.5), :636
  const float upper = 0.5f * erfcf(c * ((0.5f - v) / scale)), lower = 0.5f * erfcf(c * ((-0.5f - v) / scale));
  pmf[(size_t)s * stride + j] = upper - lower;
  if (j == 0) tail[s] = 2.f * lower;                                 // :642
}
int eb_pmf(const float* params, const float* pmf_start, int C, int max_length, float* pmf, int stride, float* tail,
           cudaStream_t s) {
  if (C <= 0 || max_length <= 0) return LMSFC_OK;
  LMSFC_LAUNCH(eb_pmf_kernel, dim3((max_length + 127) / 128, C), dim3(128), 0, s, params, pmf_start, max_length, pmf, stride, tail);
  return LMSFC_OK;
}
int gc_pmf(const float* table, const int* center, int S, int max_length, float* pmf, int stride, float* tail, cudaStream_t s) {
  if (S <= 0 || max_length <= 0) return LMSFC_OK;
  LMSFC_LAUNCH(gc_pmf_kernel, dim3((max_length + 127) / 128, S), dim3(128), 0, s, table, center, max_length, pmf, stride, tail);
  return LMSFC_OK;
}

static dim3 eb_grid(int C, int64_t P) {
  const int bx = (C + 127) / 128;
  int64_t by = (148 * 4 + bx - 1) / bx;
  if (by > P) by = P;
  if (by < 1) by = 1;
  return dim3(bx, (unsigned)by);
}

int eb_forward(const EbArgs& a, cudaStream_t s) {
  if (a.P <= 0 || a.C <= 0) return LMSFC_OK;
  LMSFC_LAUNCH(eb_forward_kernel, dim3(eb_grid(a.C, a.P)), dim3(128), 0, s, a);
  return LMSFC_OK;
}
int eb_backward(const EbBwdArgs& a, cudaStream_t s) {
  if (a.P <= 0 || a.C <= 0) return LMSFC_OK;
  LMSFC_LAUNCH(eb_backward_kernel, dim3(eb_grid(a.C, a.P)), dim3(128), 0, s, a);
  return LMSFC_OK;
}
int eb_aux(const float* params, const float* quantiles, const float* target, int C, float* loss, float* g_quantiles,
           cudaStream_t s) {
  LMSFC_LAUNCH(eb_aux_kernel, dim3((C + 127) / 128), dim3(128), 0, s, params, quantiles, target, C, loss, g_quantiles);
  return LMSFC_OK;
}

}  // namespace lmsfc
