// extern "C" surface of libb200lmsfc.so (declared in include/b200lmsfc.h): argument checking and
// the mapping from conv/deconv/GDN semantics onto the two tcgen05 GEMM kernels.
#include "../../include/b200lmsfc.h"

#include <string.h>

#include "igemm.cuh"
#include "kernels.cuh"

namespace lmsfc {
const char* get_error();
}
using namespace lmsfc;
typedef __nv_bfloat16 bf16;
#define ST(s) ((cudaStream_t)(s))

extern "C" const char* lmsfc_last_error(void) { return lmsfc::get_error(); }
extern "C" int lmsfc_version(void) { return 100; }
extern "C" uint64_t lmsfc_launch_count(void) { return lmsfc::g_launch_count; }

extern "C" int lmsfc_set_batch_invariant(int on) {
  lmsfc::set_batch_invariant(on != 0);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- tap tables
static void taps_regular(int k, int pad, int* ntaps, int8_t* dy, int8_t* dx, int16_t* widx) {
  int n = 0;
  for (int r = 0; r < k; ++r)
    for (int t = 0; t < k; ++t) { dy[n] = (int8_t)(r - pad); dx[n] = (int8_t)(t - pad); widx[n] = (int16_t)(r * k + t); ++n; }
  *ntaps = n;
}
// sub-pixel phase (py,px) of a stride-s transposed conv: out[s*h+py] gathers in[h + dy], dy = (py+pad-r)/s
static void taps_phase(int k, int s, int pad, int py, int px, int* ntaps, int8_t* dy, int8_t* dx, int16_t* widx) {
  int n = 0;
  for (int r = 0; r < k; ++r) {
    if ((py + pad - r) % s != 0) continue;
    for (int t = 0; t < k; ++t) {
      if ((px + pad - t) % s != 0) continue;
      dy[n] = (int8_t)((py + pad - r) / s); dx[n] = (int8_t)((px + pad - t) / s); widx[n] = (int16_t)(r * k + t); ++n;
    }
  }
  *ntaps = n;
}

static int check_desc(const lmsfc_conv_desc* d) {
  LMSFC_CHECK_ARG(d != nullptr, "conv: null descriptor");
  LMSFC_CHECK_ARG(d->kind == 0 || d->kind == 1, "conv: kind=%d", d->kind);
  LMSFC_CHECK_ARG((d->k == 5 || d->k == 3 || d->k == 1) && d->stride >= 1 && d->stride <= 2, "conv: k=%d stride=%d", d->k,
                  d->stride);
  LMSFC_CHECK_ARG(d->B > 0 && d->Hin > 0 && d->Win > 0 && d->Hout > 0 && d->Wout > 0, "conv: empty shape");
  LMSFC_CHECK_ARG(d->c0 > 0 && d->c0 % 64 == 0 && d->c1 % 64 == 0 && d->Cout % 128 == 0, "conv: channels %d+%d -> %d",
                  d->c0, d->c1, d->Cout);
  if (d->kind == 0)
    LMSFC_CHECK_ARG(d->Hout == (d->Hin + d->stride - 1) / d->stride && d->Wout == (d->Win + d->stride - 1) / d->stride,
                    "conv: output %dx%d does not match input %dx%d / stride %d", d->Hout, d->Wout, d->Hin, d->Win,
                    d->stride);
  else
    LMSFC_CHECK_ARG(d->Hout <= d->stride * d->Hin && d->Wout <= d->stride * d->Win && d->c1 == 0,
                    "deconv: output %dx%d exceeds %d x input %dx%d", d->Hout, d->Wout, d->stride, d->Hin, d->Win);
  return LMSFC_OK;
}

// "regular" structure: out[h][w] = sum_taps src[h*s + r - pad][w*s + t - pad] . W[tap]
static int run_regular(const bf16* s0, const bf16* s1, int c0, int c1, int B, int Hs, int Ws, int Hg, int Wg, int k, int s,
                       const bf16* w, int w_rows, int N, IgemmParams& p, cudaStream_t st) {
  p.src0 = s0; p.src1 = s1; p.c0 = c0; p.c1 = c1;
  p.B = B; p.Hin = Hs; p.Win = Ws; p.sin = s;
  p.nphase = 1;
  p.ph[0].Hg = Hg; p.ph[0].Wg = Wg; p.ph[0].M = B * Hg * Wg; p.ph[0].oy = 0; p.ph[0].ox = 0; p.ph[0].tap_begin = 0;
  taps_regular(k, k / 2, &p.ph[0].ntaps, p.dy, p.dx, p.widx);
  p.Hout = Hg; p.Wout = Wg; p.sout = 1; p.N = N;
  p.w = w; p.w_rows = w_rows; p.w_cols = k * k * (c0 + c1); p.w_mn = 0;
  return launch_igemm(p, st);
}
// "transposed" structure: out[s*h+py][s*w+px] = sum_{taps of phase} src[h+dy][w+dx] . W[tap];
// the s*s sub-pixel phases partition the k*k taps and run as blockIdx.z slices of ONE launch
// The weight block is the layer's direct [D0 = c0][T][D1 = w_d1] copy, consumed as an MN-major B operand.
static int run_transposed(const bf16* s0, int c0, int B, int Hs, int Ws, int Ho, int Wo, int k, int s, const bf16* w,
                          int w_d1, int N, IgemmParams& p, cudaStream_t st) {
  p.src0 = s0; p.src1 = nullptr; p.c0 = c0; p.c1 = 0;
  p.B = B; p.Hin = Hs; p.Win = Ws; p.sin = 1;
  p.Hout = Ho; p.Wout = Wo; p.sout = s; p.N = N;
  p.w = w; p.w_rows = c0; p.w_cols = k * k * w_d1; p.w_mn = 1; p.w_ld_tap = w_d1;
  int np = 0, tb = 0;
  for (int py = 0; py < s; ++py)
    for (int px = 0; px < s; ++px) {
      const int Hg = (Ho - py + s - 1) / s, Wg = (Wo - px + s - 1) / s;
      if (Hg <= 0 || Wg <= 0) continue;
      int nt = 0;
      taps_phase(k, s, k / 2, py, px, &nt, p.dy + tb, p.dx + tb, p.widx + tb);
      if (nt == 0) continue;
      IgemmPhase& ph = p.ph[np++];
      ph.Hg = Hg; ph.Wg = Wg; ph.M = B * Hg * Wg; ph.oy = py; ph.ox = px; ph.tap_begin = tb; ph.ntaps = nt;
      tb += nt;
    }
  p.nphase = np;
  return launch_igemm(p, st);
}

extern "C" int lmsfc_conv_forward(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* w,
                                  const float* bias, int relu, void* out_bf16, float* out_f32, void* out_abs,
                                  void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  LMSFC_CHECK_ARG(src0 && w && (d->c1 == 0 || src1), "conv_forward: null operand");
  IgemmParams p;
  memset(&p, 0, sizeof(p));
  p.epi = EPI_BIAS; p.relu = relu; p.bias = bias;
  p.out_bf16 = (bf16*)out_bf16; p.out_f32 = out_f32; p.out_abs = (bf16*)out_abs;
  if (d->kind == 0)
    return run_regular((const bf16*)src0, (const bf16*)src1, d->c0, d->c1, d->B, d->Hin, d->Win, d->Hout, d->Wout, d->k,
                       d->stride, (const bf16*)w, d->Cout, d->Cout, p, ST(stream));
  if (d->stride == 1)
    LMSFC_CHECK_ARG(false, "deconv with stride 1 is not part of the L-MSFC path");
  return run_transposed((const bf16*)src0, d->c0, d->B, d->Hin, d->Win, d->Hout, d->Wout, d->k, d->stride, (const bf16*)w,
                        d->Cout, d->Cout, p, ST(stream));
}

extern "C" int lmsfc_conv_dgrad(const lmsfc_conv_desc* d, const void* gout, const void* w, int n_in, void* gin_bf16,
                                void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  LMSFC_CHECK_ARG(gout && w && gin_bf16, "conv_dgrad: null operand");
  LMSFC_CHECK_ARG(n_in > 0 && n_in <= d->c0 + d->c1 && n_in % 128 == 0, "conv_dgrad: n_in=%d", n_in);
  IgemmParams p;
  memset(&p, 0, sizeof(p));
  p.epi = EPI_BIAS; p.out_bf16 = (bf16*)gin_bf16;
  const int cin = d->c0 + d->c1;
  if (d->kind == 0) {
    // grad of a strided conv = transposed structure over gout; stride-1 convs degenerate to one phase
    return run_transposed((const bf16*)gout, d->Cout, d->B, d->Hout, d->Wout, d->Hin, d->Win, d->k, d->stride,
                          (const bf16*)w, cin, n_in, p, ST(stream));
  }
  // grad of a transposed conv = regular strided conv over (cropped) gout
  return run_regular((const bf16*)gout, nullptr, d->Cout, 0, d->B, d->Hout, d->Wout, d->Hin, d->Win, d->k, d->stride,
                     (const bf16*)w, cin, n_in, p, ST(stream));
}

// ---------------------------------------------------------------- conv / data gradient with the GDN fused behind it
static int conv_fwd_params(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* w, IgemmParams& p,
                           cudaStream_t st) {
  if (d->kind == 0)
    return run_regular((const bf16*)src0, (const bf16*)src1, d->c0, d->c1, d->B, d->Hin, d->Win, d->Hout, d->Wout, d->k,
                       d->stride, (const bf16*)w, d->Cout, d->Cout, p, st);
  LMSFC_CHECK_ARG(d->stride != 1, "deconv with stride 1 is not part of the L-MSFC path");
  return run_transposed((const bf16*)src0, d->c0, d->B, d->Hin, d->Win, d->Hout, d->Wout, d->k, d->stride, (const bf16*)w,
                        d->Cout, d->Cout, p, st);
}
static int conv_dgrad_params(const lmsfc_conv_desc* d, const void* gout, const void* w, int n_in, IgemmParams& p,
                             cudaStream_t st) {
  const int cin = d->c0 + d->c1;
  if (d->kind == 0)
    return run_transposed((const bf16*)gout, d->Cout, d->B, d->Hout, d->Wout, d->Hin, d->Win, d->k, d->stride,
                          (const bf16*)w, cin, n_in, p, st);
  return run_regular((const bf16*)gout, nullptr, d->Cout, 0, d->B, d->Hout, d->Wout, d->Hin, d->Win, d->k, d->stride,
                     (const bf16*)w, cin, n_in, p, st);
}

extern "C" int lmsfc_conv_gdn_fusable(const lmsfc_conv_desc* d, int dgrad, int n_in, int* fusable) {
  int rc = check_desc(d);
  if (rc) return rc;
  LMSFC_CHECK_ARG(fusable != nullptr, "conv_gdn_fusable: null result");
  *fusable = 0;
  const int N = dgrad ? n_in : d->Cout;
  if (N != 256) return LMSFC_OK;
  IgemmParams p;
  memset(&p, 0, sizeof(p));
  p.dry_run = 1;
  static bf16* dummy = nullptr;      // the dry run never dereferences its operands
  rc = dgrad ? conv_dgrad_params(d, dummy, dummy, n_in, p, nullptr) : conv_fwd_params(d, dummy, dummy, dummy, p, nullptr);
  if (rc) return rc;
  *fusable = (p.dry_bn == 256 && p.ksplit == 1) ? 1 : 0;
  return LMSFC_OK;
}

extern "C" int lmsfc_conv_gdn_forward(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* w,
                                      const float* bias, const void* gamma_eff, const float* beta_eff, int inverse,
                                      void* x_out, void* x2_out, void* y_out, void* rn_out, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  LMSFC_CHECK_ARG(src0 && w && (d->c1 == 0 || src1) && gamma_eff && beta_eff && y_out, "conv_gdn_forward: null operand");
  LMSFC_CHECK_ARG(d->Cout == 256, "conv_gdn_forward: the fused GDN needs 256 output channels, got %d", d->Cout);
  IgemmParams p;
  memset(&p, 0, sizeof(p));
  p.epi = EPI_BIAS; p.bias = bias; p.inverse = inverse;
  p.fuse = 1; p.gamma = (const bf16*)gamma_eff; p.beta = beta_eff;
  p.out_x = (bf16*)x_out; p.out_t = (bf16*)x2_out; p.out_bf16 = (bf16*)y_out; p.out_rn = (bf16*)rn_out;
  return conv_fwd_params(d, src0, src1, w, p, ST(stream));
}

extern "C" int lmsfc_conv_dgrad_gdn_backward(const lmsfc_conv_desc* d, const void* gout, const void* w, int n_in,
                                             const void* addend, const void* x, const void* rn, const void* gamma_eff,
                                             int inverse, void* t_out, void* gx_out, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  LMSFC_CHECK_ARG(gout && w && x && rn && gamma_eff && t_out && gx_out, "conv_dgrad_gdn_backward: null operand");
  LMSFC_CHECK_ARG(n_in == 256 && n_in <= d->c0 + d->c1, "conv_dgrad_gdn_backward: the fused GDN needs n_in == 256, got %d", n_in);
  IgemmParams p;
  memset(&p, 0, sizeof(p));
  p.epi = EPI_BIAS; p.inverse = inverse;
  p.fuse = 2; p.gamma = (const bf16*)gamma_eff;
  p.aux_add = (const bf16*)addend; p.aux_x = (const bf16*)x; p.aux_rn = (const bf16*)rn;
  p.out_t = (bf16*)t_out; p.out_bf16 = (bf16*)gx_out;
  return conv_dgrad_params(d, gout, w, n_in, p, ST(stream));
}

extern "C" int lmsfc_conv_wgrad(const lmsfc_conv_desc* d, const void* src0, const void* src1, const void* gout, float* gw,
                                float* gbias, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  LMSFC_CHECK_ARG(src0 && gout && gw && (d->c1 == 0 || src1), "conv_wgrad: null operand");
  WgradParams p;
  memset(&p, 0, sizeof(p));
  p.B = d->B; p.sin = d->stride; p.T = d->k * d->k; p.out = gw;
  taps_regular(d->k, d->k / 2, &p.ntaps, p.dy, p.dx, p.widx);
  if (d->kind == 0) {   // dense = gout, gathered = input
    p.dense = (const bf16*)gout; p.Cd = d->Cout; p.Hg = d->Hout; p.Wg = d->Wout;
    p.g0 = (const bf16*)src0; p.g1 = (const bf16*)src1; p.gc0 = d->c0; p.gc1 = d->c1; p.Cg = d->c0 + d->c1;
    p.Hin = d->Hin; p.Win = d->Win;
  } else {              // dense = input, gathered = gout
    p.dense = (const bf16*)src0; p.Cd = d->c0; p.Hg = d->Hin; p.Wg = d->Win;
    p.g0 = (const bf16*)gout; p.g1 = nullptr; p.gc0 = d->Cout; p.gc1 = 0; p.Cg = d->Cout;
    p.Hin = d->Hout; p.Win = d->Wout;
  }
  p.M = p.B * p.Hg * p.Wg;
  if (gbias && d->Cout == 256) {   // bias gradient by the spare CTAs of the weight-gradient wave
    p.bias_src = (const bf16*)gout; p.bias_P = (long long)d->B * d->Hout * d->Wout; p.gbias = gbias;
    return launch_wgrad(p, ST(stream));
  }
  rc = launch_wgrad(p, ST(stream));
  if (rc) return rc;
  if (gbias) return colsum_bf16((const bf16*)gout, (int64_t)d->B * d->Hout * d->Wout, d->Cout, gbias, ST(stream));
  return LMSFC_OK;
}

// ---------------------------------------------------------------- GDN
static void gdn_gemm_params(IgemmParams& p, const bf16* a, const bf16* w, int64_t P, int C) {
  memset(&p, 0, sizeof(p));
  p.src0 = a; p.c0 = C; p.B = 1; p.Hin = 1; p.Win = (int)P; p.sin = 1;
  p.nphase = 1;
  p.ph[0].Hg = 1; p.ph[0].Wg = (int)P; p.ph[0].M = (int)P; p.ph[0].ntaps = 1;
  p.dy[0] = 0; p.dx[0] = 0; p.widx[0] = 0;
  p.Hout = 1; p.Wout = (int)P; p.sout = 1; p.N = C;
  p.w = w; p.w_rows = C; p.w_cols = C;
}

extern "C" int lmsfc_gdn_forward(const void* x, const void* gamma_eff, const float* beta_eff, int inverse, int64_t P, int C,
                                 void* x2_scratch, void* y, void* rn, void* stream) {
  LMSFC_CHECK_ARG(x && gamma_eff && beta_eff && y, "gdn_forward: null operand");
  LMSFC_CHECK_ARG(C % 128 == 0 && P > 0 && P < (1ll << 31), "gdn_forward: P=%lld C=%d", (long long)P, C);
  IgemmParams p;
  if (!x2_scratch) {   // inference: x^2 is formed by the GEMM's producer warps, nothing is kept
    gdn_gemm_params(p, (const bf16*)x, (const bf16*)gamma_eff, P, C);
    p.a_mode = 1;
  } else {             // training: a streaming kernel writes x^2 (kept for the gamma gradient); A/B-measured 0.5 %
                       // faster in the step than the fused producer (synchronous loads vs. the cp.async pipeline)
    int rc = square_bf16((const bf16*)x, (bf16*)x2_scratch, P * C, ST(stream));
    if (rc) return rc;
    gdn_gemm_params(p, (const bf16*)x2_scratch, (const bf16*)gamma_eff, P, C);
  }
  p.epi = EPI_GDN; p.inverse = inverse; p.bias = beta_eff; p.aux_x = (const bf16*)x;
  p.out_bf16 = (bf16*)y; p.out_rn = (bf16*)rn;
  return launch_igemm(p, ST(stream));
}

extern "C" int lmsfc_gdn_backward(const void* gy, const void* x, const void* rn, const void* gamma_eff, int inverse,
                                  int64_t P, int C, void* x2, int x2_ready, void* t_scratch, void* gx,
                                  float* g_gamma_eff, float* g_beta_eff, void* stream) {
  LMSFC_CHECK_ARG(gy && x && rn && gamma_eff && t_scratch && gx && (x2 || !g_gamma_eff) && (!g_gamma_eff == !g_beta_eff),
                  "gdn_backward: null operand (the two parameter gradients are computed together, or both deferred)");
  LMSFC_CHECK_ARG(C % 256 == 0 && P > 0 && P < (1ll << 31), "gdn_backward: P=%lld C=%d", (long long)P, C);
  int rc;
  if (x2 && !x2_ready) {
    rc = square_bf16((const bf16*)x, (bf16*)x2, P * C, ST(stream));
    if (rc) return rc;
  }
  // t by a plain streaming kernel (its column sums = g_beta ride in the weight-gradient wave).  Forming t inside the
  // GEMM's producer warps (a_mode 2) was measured slower in the step: three synchronous load round trips per K block
  // from HBM-resident x / rn under contention, against one cp.async pipeline over L2-resident t.
  rc = gdn_t((const bf16*)gy, (const bf16*)x, (const bf16*)rn, inverse, P, C, (bf16*)t_scratch, nullptr, ST(stream));
  if (rc) return rc;
  IgemmParams p;
  gdn_gemm_params(p, (const bf16*)t_scratch, (const bf16*)gamma_eff, P, C);
  p.w_mn = 1; p.w_ld_tap = C;      // s_j = sum_i gamma[i][j] t_i: gamma read as an MN-major operand
  p.epi = EPI_GDN_BWD; p.inverse = inverse; p.aux_x = (const bf16*)x; p.aux_gy = (const bf16*)gy; p.aux_rn = (const bf16*)rn;
  p.out_bf16 = (bf16*)gx;
  rc = launch_igemm(p, ST(stream));
  if (rc || !g_gamma_eff) return rc;
  return lmsfc_gdn_wgrad(t_scratch, x2, P, C, g_gamma_eff, g_beta_eff, stream);
}

extern "C" int lmsfc_gdn_wgrad(const void* t, const void* x2, int64_t P, int C, float* g_gamma_eff, float* g_beta_eff,
                               void* stream) {
  LMSFC_CHECK_ARG(t && x2 && g_gamma_eff && g_beta_eff, "gdn_wgrad: null operand");
  LMSFC_CHECK_ARG(C % 256 == 0 && P > 0 && P < (1ll << 31), "gdn_wgrad: P=%lld C=%d", (long long)P, C);
  WgradParams q;
  memset(&q, 0, sizeof(q));
  q.dense = (const bf16*)t; q.Cd = C; q.B = 1; q.Hg = 1; q.Wg = (int)P; q.M = (int)P;
  q.g0 = (const bf16*)x2; q.gc0 = C; q.Cg = C; q.Hin = 1; q.Win = (int)P; q.sin = 1;
  q.ntaps = 1; q.T = 1; q.out = g_gamma_eff;
  if (C == 256) {      // g_beta = column sums of t: by the spare CTAs of the same wave
    q.bias_src = (const bf16*)t; q.bias_P = P; q.gbias = g_beta_eff;
    return launch_wgrad(q, ST(stream));
  }
  int rc = launch_wgrad(q, ST(stream));
  if (rc) return rc;
  return colsum_bf16((const bf16*)t, P, C, g_beta_eff, ST(stream));
}

extern "C" int lmsfc_reparam_forward(const float* p, int64_t n, float bound, float pedestal, float* out_f32, void* out_bf16,
                                     void* stream) {
  return reparam_fwd(p, n, bound, pedestal, out_f32, (bf16*)out_bf16, ST(stream));
}
extern "C" int lmsfc_reparam_backward(const float* g_out, const float* p, int64_t n, float bound, float* g_p, void* stream) {
  return reparam_bwd(g_out, p, n, bound, g_p, ST(stream));
}

extern "C" int lmsfc_reparam_forward_multi(const float* params, const int64_t* table, const float* bounds, int nseg,
                                           int64_t max_n, float pedestal, float* out_f32, void* out_bf16, void* stream) {
  LMSFC_CHECK_ARG(nseg == 0 || (params && table && bounds), "reparam_forward_multi: null operand");
  return reparam_fwd_multi(params, table, bounds, nseg, max_n, pedestal, out_f32, (bf16*)out_bf16, ST(stream));
}
extern "C" int lmsfc_reparam_backward_multi(const float* g_eff, const float* params, const int64_t* table,
                                            const float* bounds, int nseg, int64_t max_n, float* g_params, void* stream) {
  LMSFC_CHECK_ARG(nseg == 0 || (g_eff && params && table && bounds && g_params), "reparam_backward_multi: null operand");
  return reparam_bwd_multi(g_eff, params, table, bounds, nseg, max_n, g_params, ST(stream));
}

// ---------------------------------------------------------------- entropy models
extern "C" int lmsfc_gc_forward(const float* y, const float* scales, const float* means, const float* noise, int64_t n,
                                int mode, float scale_bound, float lik_bound, uint64_t seed, uint64_t offset,
                                const uint64_t* rng_state, float* y_hat, void* y_hat_bf16, float* lik, float* sum_ln,
                                void* stream) {
  LMSFC_CHECK_ARG(mode >= 0 && mode <= 2, "Invalid quantization mode: \"%d\"", mode);   // entropy_models.py:158-159
  LMSFC_CHECK_ARG(n == 0 || (y && (scales ? (lik != nullptr) : (y_hat || y_hat_bf16))), "gc_forward: null operand");
  GcArgs a{y, scales, means, noise, y_hat, lik, sum_ln, (bf16*)y_hat_bf16, n, mode, scale_bound, lik_bound, seed, offset, rng_state};
  return gc_forward(a, ST(stream));
}
extern "C" int lmsfc_gc_backward(const float* y_hat, const float* scales, const float* means, const float* g_lik,
                                 const float* g_yhat, const void* g_yhat_bf16, int64_t n, int mode, float scale_bound,
                                 float lik_bound, float rate_coef, float* g_y, float* g_scales, float* g_means,
                                 void* stream) {
  LMSFC_CHECK_ARG(mode >= 0 && mode <= 2, "Invalid quantization mode: \"%d\"", mode);
  LMSFC_CHECK_ARG(n == 0 || (y_hat && scales && g_y && g_scales), "gc_backward: null operand");
  GcBwdArgs a{y_hat, scales, means, g_lik, g_yhat, (const bf16*)g_yhat_bf16, g_y, g_scales, g_means, n, mode, scale_bound, lik_bound, rate_coef};
  return gc_backward(a, ST(stream));
}
extern "C" int lmsfc_eb_forward(const float* x, const float* params, const float* medians, const float* noise, int64_t P,
                                int C, int mode, float lik_bound, uint64_t seed, uint64_t offset,
                                const uint64_t* rng_state, float* y_hat, void* y_hat_bf16, float* lik, float* sum_ln,
                                void* stream) {
  LMSFC_CHECK_ARG(mode == 0 || mode == 1, "Invalid quantization mode: \"%d\"", mode);
  LMSFC_CHECK_ARG(P == 0 || (x && params && y_hat && lik), "eb_forward: null operand");
  LMSFC_CHECK_ARG(mode == 0 || medians, "eb_forward: dequantize mode needs medians");
  EbArgs a{x, params, medians, noise, y_hat, lik, sum_ln, (bf16*)y_hat_bf16, P, C, mode, lik_bound, seed, offset, rng_state};
  return eb_forward(a, ST(stream));
}
extern "C" int lmsfc_eb_backward(const float* y_hat, const float* params, const float* g_lik, const float* g_yhat,
                                 const void* g_yhat_bf16, int64_t P, int C, int mode, float lik_bound, float rate_coef, float* g_x, float* g_params,
                                 void* stream) {
  LMSFC_CHECK_ARG(P == 0 || (y_hat && params && g_x && g_params), "eb_backward: null operand");
  EbBwdArgs a{y_hat, params, g_lik, g_yhat, (const bf16*)g_yhat_bf16, g_x, g_params, P, C, mode, lik_bound, rate_coef};
  return eb_backward(a, ST(stream));
}
extern "C" int lmsfc_eb_aux_loss(const float* params, const float* quantiles, const float* target, int C, float* loss,
                                 float* g_quantiles, void* stream) {
  LMSFC_CHECK_ARG(params && quantiles && target && loss && g_quantiles && C > 0, "eb_aux_loss: null operand");
  return eb_aux(params, quantiles, target, C, loss, g_quantiles, ST(stream));
}
extern "C" int lmsfc_eb_pmf(const float* params, const float* pmf_start, int C, int max_length, float* pmf, int pmf_stride,
                            float* tail, void* stream) {
  LMSFC_CHECK_ARG(params && pmf_start && pmf && tail && pmf_stride >= max_length, "eb_pmf: null operand / stride");
  return eb_pmf(params, pmf_start, C, max_length, pmf, pmf_stride, tail, ST(stream));
}
extern "C" int lmsfc_gc_pmf(const float* scale_table, const int32_t* pmf_center, int S, int max_length, float* pmf,
                            int pmf_stride, float* tail, void* stream) {
  LMSFC_CHECK_ARG(scale_table && pmf_center && pmf && tail && pmf_stride >= max_length, "gc_pmf: null operand / stride");
  return gc_pmf(scale_table, pmf_center, S, max_length, pmf, pmf_stride, tail, ST(stream));
}
extern "C" int lmsfc_rng_advance(uint64_t* rng_state, void* stream) { return rng_advance(rng_state, ST(stream)); }

// ---------------------------------------------------------------- loss / small ops / optimiser / cdf
extern "C" int lmsfc_mse_forward_backward(const void* xhat, const void* x, int64_t n, float grad_coef, float* sse, void* grad,
                                          void* stream) {
  LMSFC_CHECK_ARG(n == 0 || (xhat && x && sse), "mse: null operand");
  return mse_fwd_bwd((const bf16*)xhat, (const bf16*)x, n, grad_coef, sse, (bf16*)grad, ST(stream));
}
extern "C" int lmsfc_mse_forward_backward_multi(int nlev, const void* const* xhat, const void* const* x, const int64_t* n,
                                                const float* grad_coef, float* sse, void* const* grad, void* stream) {
  LMSFC_CHECK_ARG(xhat && x && n && grad_coef && sse, "mse_multi: null operand");
  return mse_fwd_bwd_multi(nlev, xhat, x, n, grad_coef, sse, grad, ST(stream));
}
extern "C" int lmsfc_rd_loss(int nlev, const void* const* xhat, const void* const* x, const int64_t* n, const float* grad_coef,
                             void* const* grad, int n_rate, const float* const* rate, float rate_coef,
                             const float* dist_coef, float lambda_scale, const float* extra, float* out4, void* stream) {
  LMSFC_CHECK_ARG(xhat && x && n && grad_coef, "rd_loss: null argument array");
  return rd_loss_multi(nlev, xhat, x, n, grad_coef, grad, n_rate, rate, rate_coef, dist_coef, lambda_scale, extra, out4,
                       ST(stream));
}
extern "C" int lmsfc_loss_finalize(const float* accum, const float* coef, int n_rate, int n_dist, float lambda_scale,
                                   float* out, void* stream) {
  return loss_finalize(accum, coef, n_rate, n_dist, lambda_scale, out, ST(stream));
}
extern "C" int lmsfc_relu_backward(void* g, const void* out, int64_t n, void* stream) {
  return relu_bwd_bf16((bf16*)g, (const bf16*)out, n, 0.f, ST(stream));
}
extern "C" int lmsfc_leaky_relu_backward(void* g, const void* out, int64_t n, float slope, void* stream) {
  return relu_bwd_bf16((bf16*)g, (const bf16*)out, n, slope, ST(stream));
}
extern "C" int lmsfc_abs_backward(const float* g_direct, const void* g_abs, const float* y, int64_t n, int relu,
                                  void* g_out, void* stream) {
  return abs_bwd(g_direct, (const bf16*)g_abs, y, n, relu, (bf16*)g_out, ST(stream));
}
extern "C" int lmsfc_colsum(const void* g, int64_t P, int C, float* out, void* stream) {
  return colsum_bf16((const bf16*)g, P, C, out, ST(stream));
}
extern "C" int lmsfc_colsum_static(const void* g, int64_t P, int C, float* out, void* stream) {
  return colsum_bf16_static((const bf16*)g, P, C, out, ST(stream));
}
extern "C" int lmsfc_cast_f32_bf16(const float* in, void* out, int64_t n, void* stream) {
  return cast_f32_bf16(in, (bf16*)out, n, ST(stream));
}
extern "C" int lmsfc_cast_bf16_f32(const void* in, float* out, int64_t n, void* stream) {
  return cast_bf16_f32((const bf16*)in, out, n, ST(stream));
}
extern "C" int lmsfc_nchw_to_nhwc_bf16(const float* in, void* out, int B, int C, int HW, void* stream) {
  return nchw_to_nhwc_bf16(in, (bf16*)out, B, C, HW, ST(stream));
}
extern "C" int lmsfc_sumsq(const float* g, int64_t n, float* accum, int* set_flag, void* stream) {
  return sumsq_f32(g, n, accum, set_flag, 0, ST(stream));
}
extern "C" int lmsfc_sumsq_part_final(const float* g, int64_t n, float* parts, int index, int nparts, float* out,
                                      int* set_flag, void* stream) {
  LMSFC_CHECK_ARG(parts != nullptr && out != nullptr && nparts >= 1 && index >= 0 && index < nparts, "sumsq_part_final: bad arguments");
  return sumsq_f32(g, n, parts + index, set_flag, 1, ST(stream), parts, nparts, out);
}
extern "C" int lmsfc_sumsq_part(const float* g, int64_t n, float* out, void* stream) {
  LMSFC_CHECK_ARG(out != nullptr, "sumsq_part: null result");
  if (n == 0) return sum_parts(nullptr, 0, out, nullptr, ST(stream));
  return sumsq_f32(g, n, out, nullptr, 1, ST(stream));
}
extern "C" int lmsfc_sum_parts(const float* parts, int n, float* out, int* set_flag, void* stream) {
  LMSFC_CHECK_ARG(parts && out && n >= 0, "sum_parts: null operand");
  return sum_parts(parts, n, out, set_flag, ST(stream));
}
extern "C" int lmsfc_adam_step(float* p, float* m, float* v, const float* g, void* p_bf16, int64_t n, float lr, float beta1,
                               float beta2, float eps, float max_norm, const float* sumsq, int* step, float* scratch,
                               void* stream) {
  LMSFC_CHECK_ARG(n == 0 || (p && m && v && g && step && scratch), "adam_step: null operand");
  AdamArgs a{p, m, v, const_cast<float*>(g), (bf16*)p_bf16, n, lr, beta1, beta2, eps, max_norm, sumsq, step, scratch, 0,
             nullptr, nullptr, 0, nullptr};
  return adam_step(a, ST(stream));
}
extern "C" int lmsfc_adam_prep(float lr, const float* lr_dev, float beta1, float beta2, float max_norm, const float* sumsq,
                               int* step, float* scratch, int* pending, int force, const float* gate, void* stream) {
  LMSFC_CHECK_ARG(step && scratch, "adam_prep: null operand");
  AdamArgs a{nullptr, nullptr, nullptr, nullptr, nullptr, 0, lr, beta1, beta2, 0.f, max_norm, sumsq, step, scratch, 0,
             lr_dev, pending, force, gate};
  return adam_prep(a, ST(stream));
}
extern "C" int lmsfc_adam_update(float* p, float* m, float* v, float* g, void* p_bf16, int64_t n, float beta1, float beta2,
                                 float eps, const float* scratch, int zero_g, void* stream) {
  LMSFC_CHECK_ARG(n == 0 || (p && m && v && g && scratch), "adam_update: null operand");
  AdamArgs a{p, m, v, g, (bf16*)p_bf16, n, 0.f, beta1, beta2, eps, 0.f, nullptr, nullptr, const_cast<float*>(scratch), zero_g,
             nullptr, nullptr, 0, nullptr};
  return adam_update(a, ST(stream));
}
extern "C" int lmsfc_pmf_to_quantized_cdf(const float* pmf_host, int n, int precision, uint32_t* cdf_host) {
  LMSFC_CHECK_ARG(pmf_host && cdf_host && n >= 0, "pmf_to_quantized_cdf: null operand");
  return pmf_to_quantized_cdf_host(pmf_host, n, precision, cdf_host);
}
extern "C" int lmsfc_pmf_rows_to_cdf(const float* pmf, int pmf_stride, const float* tail, const int* pmf_length, int rows,
                                     int precision, int32_t* cdf, int cdf_stride, int* status, void* stream) {
  return pmf_rows_to_cdf_device(pmf, pmf_stride, tail, pmf_length, rows, precision, cdf, cdf_stride, status, ST(stream));
}

// ---------------------------------------------------------------- whole-step capture / launch
struct lmsfc_step_graph {
  cudaGraph_t graph;
  cudaGraphExec_t exec;
  int kernels;
};
extern "C" int lmsfc_step_capture_begin(void* stream) {
  LMSFC_CHECK_ARG(stream != nullptr, "step_capture_begin: the legacy default stream cannot be captured");
  // thread-local mode: other host threads (a loader, an autograd worker launching on forked streams) stay legal
  LMSFC_CUDA(cudaStreamBeginCapture(ST(stream), cudaStreamCaptureModeThreadLocal));
  return LMSFC_OK;
}
extern "C" int lmsfc_step_capture_end(void* stream, lmsfc_step_graph** out, int* kernels) {
  LMSFC_CHECK_ARG(out != nullptr, "step_capture_end: null out");
  *out = nullptr;
  cudaGraph_t graph = nullptr;
  {
    cudaError_t e = cudaStreamEndCapture(ST(stream), &graph);
    if (e != cudaSuccess) {
      cudaGetLastError();          // not sticky: leave no stale error behind for the caller's next launch check
      lmsfc::set_error("step_capture_end: cudaStreamEndCapture -> %s", cudaGetErrorString(e));
      return LMSFC_ERR_CUDA;
    }
  }
  LMSFC_CHECK_ARG(graph != nullptr, "step_capture_end: the capture was invalidated (a call inside it is not capturable)");
  size_t n = 0;
  int nk = 0;
  if (cudaGraphGetNodes(graph, nullptr, &n) == cudaSuccess && n > 0) {
    cudaGraphNode_t* nodes = new cudaGraphNode_t[n];
    if (cudaGraphGetNodes(graph, nodes, &n) == cudaSuccess)
      for (size_t i = 0; i < n; ++i) {
        cudaGraphNodeType t;
        if (cudaGraphNodeGetType(nodes[i], &t) == cudaSuccess && t == cudaGraphNodeTypeKernel) ++nk;
      }
    delete[] nodes;
  }
  cudaGraphExec_t exec = nullptr;
  cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
  if (e != cudaSuccess) {
    cudaGraphDestroy(graph);
    lmsfc::set_error("step_capture_end: cudaGraphInstantiate -> %s", cudaGetErrorString(e));
    return LMSFC_ERR_CUDA;
  }
  lmsfc_step_graph* g = new lmsfc_step_graph{graph, exec, nk};
  if (kernels) *kernels = nk;
  *out = g;
  return LMSFC_OK;
}
extern "C" int lmsfc_step_launch(lmsfc_step_graph* g, void* stream) {
  LMSFC_CHECK_ARG(g != nullptr && g->exec != nullptr, "step_launch: null graph");
  LMSFC_CUDA(cudaGraphLaunch(g->exec, ST(stream)));
  lmsfc::g_launch_count += (unsigned long long)g->kernels;
  return LMSFC_OK;
}
extern "C" int lmsfc_step_destroy(lmsfc_step_graph* g) {
  if (!g) return LMSFC_OK;
  cudaError_t e1 = g->exec ? cudaGraphExecDestroy(g->exec) : cudaSuccess;
  cudaError_t e2 = g->graph ? cudaGraphDestroy(g->graph) : cudaSuccess;
  delete g;
  LMSFC_CUDA(e1);
  LMSFC_CUDA(e2);
  return LMSFC_OK;
}
