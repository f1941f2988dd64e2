// Optimiser-side kernels over the flat parameter arena: fused clip_grad_norm_ + Adam with the
// bf16 GEMM-operand shadow written in the same pass, weight repacking, layout conversion.
//   compressai/examples/train.py:100-107 (clip, optimizer.step, aux_optimizer.step)
//   torch.optim.Adam defaults (betas 0.9/0.999, eps 1e-8, no weight decay, no amsgrad)
#include "kernels.cuh"

namespace lmsfc {

namespace {
constexpr int TPB = 256;
inline int grid1d(int64_t work_items) {
  int64_t b = (work_items + TPB - 1) / TPB;
  const int64_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
}  // namespace

// ---------------------------------------------------------------- sum of squares (grad norm)
__global__ void __launch_bounds__(TPB) sumsq_kernel(const float4* __restrict__ g, int64_t n4, const float* tail, int ntail,
                                                    float* accum) {
  pdl_sync();
  float local = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < n4; i += (int64_t)gridDim.x * TPB) {
    const float4 v = __ldg(g + i);
    local += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < ntail) local += tail[threadIdx.x] * tail[threadIdx.x];
  __shared__ float red[32];
  local = block_sum(local, red);
  if (threadIdx.x == 0) atomicAdd(accum, local);
}
int sumsq_f32(const float* g, int64_t n, float* accum, cudaStream_t s) {
  if (n == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG(((uintptr_t)g & 15) == 0, "sumsq: pointer must be 16-byte aligned");
  const int64_t n4 = n / 4;
  LMSFC_LAUNCH(sumsq_kernel, dim3(grid1d(n4 > 0 ? n4 : 1)), dim3(TPB), 0, s, (const float4*)g, n4, g + n4 * 4, (int)(n - n4 * 4), accum);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- Adam
// prep: one thread bumps the step counter and derives {clip_coef, bias_corr1, bias_corr2, norm}
__global__ void adam_prep_kernel(AdamArgs a) {
  pdl_sync();
  const int step = *a.step + 1;
  *a.step = step;
  float clip = 1.f, norm = 0.f;
  if (a.sumsq) {
    norm = sqrtf(*a.sumsq);
    if (a.max_norm > 0.f) clip = fminf(1.f, a.max_norm / (norm + 1e-6f));   // torch clip_grad_norm_
  }
  a.scratch[0] = clip;
  a.scratch[1] = 1.f - powf(a.beta1, (float)step);
  a.scratch[2] = 1.f - powf(a.beta2, (float)step);
  a.scratch[3] = norm;
}
__global__ void __launch_bounds__(TPB) adam_kernel(AdamArgs a) {
  pdl_sync();
  const float clip = a.scratch[0], bc1 = a.scratch[1], bc2 = a.scratch[2];
  const float step_size = a.lr / bc1, rs_bc2 = rsqrtf(bc2);
  for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * TPB) {
    const float g = a.g[i] * clip;
    const float m = a.beta1 * a.m[i] + (1.f - a.beta1) * g;
    const float v = a.beta2 * a.v[i] + (1.f - a.beta2) * g * g;
    const float p = a.p[i] - step_size * m / (sqrtf(v) * rs_bc2 + a.eps);
    a.m[i] = m; a.v[i] = v; a.p[i] = p;
    if (a.p_bf16) a.p_bf16[i] = __float2bfloat16(p);
  }
}
int adam_step(const AdamArgs& a, cudaStream_t s) {
  if (a.n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(adam_prep_kernel, dim3(1), dim3(1), 0, s, a);
  LMSFC_LAUNCH(adam_kernel, dim3(grid1d(a.n)), dim3(TPB), 0, s, a);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- layout
// fp32 NCHW -> bf16 NHWC (feature ingest for callers that hand over the reference's layout)
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int C,
                                                           int HW) {
  pdl_sync();
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int pb = blockIdx.x * 32, cb = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int c = cb + r, p = pb + threadIdx.x;
    tile[r][threadIdx.x] = (c < C && p < HW) ? in[((size_t)b * C + c) * HW + p] : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int p = pb + r, c = cb + threadIdx.x;
    if (c < C && p < HW) out[((size_t)b * HW + p) * C + c] = __float2bfloat16(tile[threadIdx.x][r]);
  }
}
int nchw_to_nhwc_bf16(const float* in, __nv_bfloat16* out, int B, int C, int HW, cudaStream_t s) {
  if (B == 0 || C == 0 || HW == 0) return LMSFC_OK;
  dim3 grid((HW + 31) / 32, (C + 31) / 32, B);
  LMSFC_LAUNCH(nchw_to_nhwc_kernel, dim3(grid), dim3(dim3(32, 8)), 0, s, in, out, C, HW);
  return LMSFC_OK;
}

}  // namespace lmsfc
