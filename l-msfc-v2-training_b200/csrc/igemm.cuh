// Parameter blocks of the tcgen05 implicit-GEMM kernels (igemm.cu).
#pragma once
#include "common.cuh"

namespace lmsfc {

constexpr int kMaxTaps = 25;
constexpr float kLeakySlope = 0.01f;   // nn.LeakyReLU's default negative_slope (IgemmParams::relu == 2)

// epilogue modes of the forward/data-gradient kernel
enum : int {
  EPI_BIAS = 0,     // v = acc + bias[n]; optional ReLU; store bf16 / fp32 / |v| bf16
  EPI_GDN = 1,      // nrm = acc + bias[n]; rn = rsqrt(nrm); y = x*rn (GDN) | x/rn (IGDN); store y, rn
  EPI_GDN_BWD = 2,  // s = acc; gx = gy*f + 2*x*s, f = rn (GDN) | 1/rn (IGDN)
};

// Gathered-A x weight-B GEMM:  D[m][n] = sum_{tap,c} A[pixel(m)+tap][c] * W[n][widx(tap)*Ctot + c]
// A launch covers up to 4 sub-pixel phases of a transposed conv (blockIdx.z / ksplit) and, for
// small grids, `ksplit` slices of the K loop (one thread-block cluster) reduced through DSMEM.
constexpr int kMaxPhases = 4;
struct IgemmPhase {
  int Hg, Wg, M;            // GEMM row grid of this phase: m = (b*Hg + hg)*Wg + wg
  int oy, ox;               // output pixel = (hg*sout + oy, wg*sout + ox)
  int tap_begin, ntaps;     // slice of the tap table
  uint32_t wg_mul, wg_shr, hg_mul, hg_shr;   // m -> (b, hg, wg) by multiply-shift (filled by launch_igemm)
};
struct IgemmParams {
  CUtensorMap tmapB;        // filled by launch_igemm, SWIZZLE_128B: box {64 k, BLOCK_N n} or, w_mn = 1, {64 n, 64 k}
  const __nv_bfloat16* w;   // weights, bf16 [w_rows][w_cols]
  int w_rows, w_cols;       //   w_mn = 0: rows = n, cols = k = widx(tap)*Ctot + c     (K-major B operand)
  int w_mn;                 //   w_mn = 1: rows = k-channel c, cols = widx(tap)*w_ld_tap + n  (MN-major B operand:
  int w_ld_tap;             //             the SAME [D0][T][D1] weight block read "transposed", no second copy)
  const __nv_bfloat16* src0;  // NHWC bf16 [B][Hin][Win][c0]
  const __nv_bfloat16* src1;  // optional second source (channel concat), [B][Hin][Win][c1]
  int c0, c1;
  int B, Hin, Win;
  int sin;                  // input pixel = (hg*sin + dy, wg*sin + dx)
  int nphase;
  IgemmPhase ph[kMaxPhases];
  int8_t dy[kMaxTaps], dx[kMaxTaps];
  int16_t widx[kMaxTaps];   // weight tap index of each gathered tap
  int Hout, Wout, sout;
  int N;                    // output channels (row pitch of every output / aux tensor)
  int epi, relu, inverse;
  int ksplit;               // filled by launch_igemm: K slices per output tile = cluster size along z
  int dbg;                  // TIMELINE builds only
  const float* bias;        // [N] or null
  __nv_bfloat16* out_bf16;  // nullable
  float* out_f32;           // nullable
  __nv_bfloat16* out_abs;   // nullable, |v| (feeds h_a, models/google.py:294)
  __nv_bfloat16* out_rn;    // EPI_GDN: rsqrt(norm), saved for backward
  // a_mode 1: the activation operand is COMPUTED by the producer warps instead of copied: A = src0^2 (GDN forward
  // without a scratch buffer); a_out (nullable) also receives it ([M][c0] bf16).  Forming the backward operand
  // t = grad_out * x * f(rn) the same way was measured slower in the step (see lmsfc_gdn_backward).
  int a_mode;
  __nv_bfloat16* a_out;
  const __nv_bfloat16* aux_x;   // EPI_GDN / EPI_GDN_BWD
  const __nv_bfloat16* aux_gy;  // EPI_GDN_BWD
  const __nv_bfloat16* aux_rn;  // EPI_GDN_BWD
  // ---- GDN fused behind the GEMM that produces its input (FUSE kernels: BLOCK_N == N == 256, ksplit == 1).  The CTA
  // owns all 256 channels of its 128 pixels, so after the main loop it forms the second operand (x^2 resp. t) from the
  // accumulator in shared memory and runs the channel contraction as a second tcgen05 GEMM into the other 256 TMEM
  // columns -- no x^2 / t round trip through HBM, no separate launch (layers/gdn.py:77-92 and its backward).
  //   fuse 1 (behind conv/deconv forward):  x = acc0 + bias ; A2 = x^2 ; acc1 = A2 . gamma^T ; nrm = acc1 + beta ;
  //                                         y = x * rsqrt(nrm) | x * sqrt(nrm) ; stores x, x^2, y, rn
  //   fuse 2 (behind a data gradient):      gy = acc0 (+ aux_add) ; A2 = t = gy * x * (-rn^3/2 | rn/2) ; acc1 = A2 . gamma ;
  //                                         gx = gy * (rn | 1/rn) + 2 x acc1 ; stores t, gx
  int fuse;
  int dry_run;                  // host only: launch_igemm just reports the tiling it would use (ksplit, dry_bn)
  int dry_bn;
  CUtensorMap tmapG;            // filled by launch_igemm: gamma_eff bf16 [C][C]; fuse 1 box {64 k, 256 n}, fuse 2 {64 n, 64 k}
  const __nv_bfloat16* gamma;   // gamma_eff (host side: tensor-map source)
  const float* beta;            // fuse 1: beta_eff [C]
  __nv_bfloat16* out_x;         // fuse 1: the convolution's own output x (nullable)
  __nv_bfloat16* out_t;         // fuse 1: x^2 (nullable, kept for gamma's gradient); fuse 2: t = dL/dnorm
  const __nv_bfloat16* aux_add; // fuse 2: gradient w.r.t. y from its other consumer (nullable)
};

// Weight-gradient GEMM (K = pixels):
//   G[d][widx(tap)][g] += sum_pix dense[pix][d] * gath[pixel(pix)+tap][g]
struct WgradParams {
  CUtensorMap tmapA;        // filled by launch_wgrad: box {64 ch, 64 px}, SWIZZLE_128B
  CUtensorMap tmapOut;      // filled by launch_wgrad: fp32 [Cd][T*Cg], box {32 cols, 128 rows}, SWIZZLE_128B
  const __nv_bfloat16* dense;  // dense operand, NHWC bf16 viewed as [M][Cd]
  const __nv_bfloat16* g0;  // gathered NHWC bf16 [B][Hin][Win][gc0]
  const __nv_bfloat16* g1;  // optional concat source [B][Hin][Win][gc1]
  int gc0, gc1;
  int B, Hin, Win;
  int Hg, Wg, M;            // dense pixel grid
  int sin;
  int ntaps;
  int8_t dy[kMaxTaps], dx[kMaxTaps];
  int16_t widx[kMaxTaps];
  float* out;               // fp32 [Cd][T][Cg], accumulated by TMA reduce-add (cp.reduce.async.bulk.tensor ... .add)
  int Cd, Cg, T;
  int kb_total, kb_per_split;
  // pixel index -> (b, hg, wg) by multiply-shift (filled by launch_wgrad): the gather recomputes it for every
  // K block, and hardware integer division there bounded the whole main loop
  uint32_t wg_mul, wg_shr, hg_mul, hg_shr;
  int grid_x, grid_y, gemm_ctas;   // the launch is 1-D: GEMM CTAs (x fastest, then y, then K split) + bias CTAs
  // optional bias gradient (column sums of grad_out [bias_P][256]) computed by the spare CTAs of the same wave
  const __nv_bfloat16* bias_src;
  long long bias_P;
  float* gbias;
  int bias_ctas;
  int dbg;                  // TIMELINE builds only
};

// n / d == umulhi(n, mul) >> shr for 0 <= n < 2^31 (mul == 0: d == 1); the kernels decode pixel indices with it
inline void fast_div_magic(int d, uint32_t* mul, uint32_t* shr) {
  if (d == 1) { *mul = 0; *shr = 0; return; }
  int l = 0;
  while ((1ll << l) < d) ++l;
  const int pw = 31 + l;
  *mul = (uint32_t)(((1ull << pw) + (uint64_t)d - 1) / (uint64_t)d);
  *shr = (uint32_t)(pw - 32);
}

int launch_igemm(IgemmParams& p, cudaStream_t stream);
void set_batch_invariant(int on);   // thread-local: fixed tile width, no split K (see igemm.cu)
int launch_wgrad(WgradParams& p, cudaStream_t stream);

}  // namespace lmsfc
