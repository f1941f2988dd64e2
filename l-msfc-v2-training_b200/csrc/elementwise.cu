// HBM-bound elementwise / reduction kernels: 16-byte vector accesses, grid sized in multiples of
// the 148 SMs, warp-shuffle block reductions, one atomic per block.
//
//   RateDistortionLoss MSE + its gradient      compressai/losses/rate_distortion.py:47-59
//   GDN helper passes (x^2, t, beta gradient)  compressai/layers/gdn.py:77-92
//   NonNegativeParametrizer fwd/bwd            compressai/ops/parametrizers.py:38-64
//   clip_grad_norm_ + Adam                     compressai/examples/train.py:100-103
#include <mutex>

#include "kernels.cuh"

namespace lmsfc {

namespace {
constexpr int TPB = 256;
inline int grid1d(int64_t work_items) {
  int64_t b = (work_items + TPB - 1) / TPB;
  const int64_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
__device__ __forceinline__ float lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
}  // namespace

// ---------------------------------------------------------------- MSE forward + backward in one pass
// sse += sum (xhat-x)^2 ; grad = grad_coef * (xhat - x).  5 bytes/element at bf16 (2 reads + 1 write).
__global__ void __launch_bounds__(TPB) mse_kernel(const uint4* __restrict__ xh, const uint4* __restrict__ x, int64_t n8,
                                                  float coef, float* sse, uint4* __restrict__ grad) {
  pdl_sync();
  float local = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < n8; i += (int64_t)gridDim.x * TPB) {
    const uint4 a = __ldg(xh + i), b = __ldg(x + i);
    const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
    uint32_t gv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d0 = lo(av[j]) - lo(bv[j]), d1 = hi(av[j]) - hi(bv[j]);
      local += d0 * d0 + d1 * d1;
      gv[j] = pack_bf16(coef * d0, coef * d1);
    }
    if (grad) grad[i] = make_uint4(gv[0], gv[1], gv[2], gv[3]);
  }
  __shared__ float red[32];
  local = block_sum(local, red);
  if (threadIdx.x == 0) atomicAdd(sse, local);
}

int mse_fwd_bwd(const __nv_bfloat16* xhat, const __nv_bfloat16* x, int64_t n, float grad_coef, float* sse,
                __nv_bfloat16* grad, cudaStream_t s) {
  LMSFC_CHECK_ARG(n % 8 == 0, "mse: n=%lld must be a multiple of 8", (long long)n);
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(mse_kernel, dim3(grid1d(n / 8)), dim3(TPB), 0, s, (const uint4*)xhat, (const uint4*)x, n / 8, grad_coef, sse, (uint4*)grad);
  return LMSFC_OK;
}

// all pyramid levels in ONE launch (the loss sits on the step's critical path between forward and backward)
struct MseMultiArgs {
  const uint4* xh[kMseMaxLevels];
  const uint4* x[kMseMaxLevels];
  uint4* grad[kMseMaxLevels];
  int64_t n8[kMseMaxLevels];
  float coef[kMseMaxLevels];
  int blk_end[kMseMaxLevels];     // blocks [blk_end[l-1], blk_end[l]) work on level l
  int nlev;
  float* sse;                     // [nlev]; accumulated (+=) by float atomics, or -- with out4 -- overwritten by the last block
  // ---- the whole RD loss in this launch (out4 != null): per-level sums through a two-stage, fixed-order reduction,
  // then the last block forms bpp / D / loss itself (losses/rate_distortion.py:47-59)
  float* out4;                    // {loss, bpp, D, *extra}
  const float* rate[kMseMaxLevels];   // device scalars sum(ln L_k), reduced by the likelihood kernels
  int n_rate;
  float rate_coef;                // 1 / (-ln 2 * num_pixels)
  float dist_coef[kMseMaxLevels]; // w_l / n_l
  float lambda_scale;
  const float* extra;             // nullable: one more scalar to ride along (the step's aux loss)
  float* partial;                 // [gridDim.x] block partials
  unsigned int* counter;
};
__global__ void __launch_bounds__(TPB) mse_multi_kernel(const __grid_constant__ MseMultiArgs a) {
  pdl_sync();
  int l = 0;
  while (l + 1 < a.nlev && (int)blockIdx.x >= a.blk_end[l]) ++l;
  const int b0 = l ? a.blk_end[l - 1] : 0, nb = a.blk_end[l] - b0;
  const uint4* __restrict__ xh = a.xh[l];
  const uint4* __restrict__ x = a.x[l];
  uint4* __restrict__ grad = a.grad[l];
  const float coef = a.coef[l];
  float local = 0.f;
  for (int64_t i = (int64_t)(blockIdx.x - b0) * TPB + threadIdx.x; i < a.n8[l]; i += (int64_t)nb * TPB) {
    const uint4 p = __ldg(xh + i), q = __ldg(x + i);
    const uint32_t av[4] = {p.x, p.y, p.z, p.w}, bv[4] = {q.x, q.y, q.z, q.w};
    uint32_t gv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d0 = lo(av[j]) - lo(bv[j]), d1 = hi(av[j]) - hi(bv[j]);
      local += d0 * d0 + d1 * d1;
      gv[j] = pack_bf16(coef * d0, coef * d1);
    }
    if (grad) grad[i] = make_uint4(gv[0], gv[1], gv[2], gv[3]);
  }
  __shared__ float red[32];
  local = block_sum(local, red);
  if (!a.out4) {
    if (threadIdx.x == 0) atomicAdd(a.sse + l, local);
    return;
  }
  __shared__ unsigned int is_last;
  if (threadIdx.x == 0) {
    __stcg(a.partial + blockIdx.x, local);
    __threadfence();
    const unsigned int prev = atomicAdd(a.counter, 1u);
    is_last = (prev == gridDim.x - 1) ? 1u : 0u;
    if (is_last) *a.counter = 0;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  float d = 0.f;
  for (int lv = 0; lv < a.nlev; ++lv) {                     // fixed order: the loss is bit-reproducible
    const int lo_b = lv ? a.blk_end[lv - 1] : 0, hi_b = a.blk_end[lv];
    float v = 0.f;
    for (int b = lo_b + (int)threadIdx.x; b < hi_b; b += TPB) v += __ldcg(a.partial + b);
    __syncthreads();                                        // red[] is reused
    v = block_sum(v, red);
    if (threadIdx.x == 0) {
      if (a.sse) a.sse[lv] = v;
      d += v * a.dist_coef[lv];
    }
  }
  if (threadIdx.x == 0) {
    float bpp = 0.f;
    for (int i = 0; i < a.n_rate; ++i) bpp += *a.rate[i] * a.rate_coef;
    a.out4[0] = a.lambda_scale * d + bpp;
    a.out4[1] = bpp;
    a.out4[2] = d;
    a.out4[3] = a.extra ? *a.extra : 0.f;
  }
}
static int mse_multi_fill(MseMultiArgs& a, int nlev, const void* const* xhat, const void* const* x, const int64_t* n,
                          const float* grad_coef, void* const* grad, int* grid);
// RateDistortionLoss in ONE launch: all levels' squared errors + gradients, then bpp / D / loss by the last block
int rd_loss_multi(int nlev, const void* const* xhat, const void* const* x, const int64_t* n, const float* grad_coef,
                  void* const* grad, int n_rate, const float* const* rate, float rate_coef, const float* dist_coef,
                  float lambda_scale, const float* extra, float* out4, cudaStream_t s) {
  LMSFC_CHECK_ARG(out4 != nullptr && n_rate >= 0 && n_rate <= kMseMaxLevels && (n_rate == 0 || rate) && dist_coef,
                  "rd_loss: bad arguments (at most %d likelihood terms)", kMseMaxLevels);
  MseMultiArgs a;
  int grid = 0;
  if (int rc = mse_multi_fill(a, nlev, xhat, x, n, grad_coef, grad, &grid)) return rc;
  if (int rc = ensure_reduce_scratch()) return rc;
  for (int i = 0; i < n_rate; ++i) {
    LMSFC_CHECK_ARG(rate[i] != nullptr, "rd_loss: null likelihood sum %d", i);
    a.rate[i] = rate[i];
  }
  for (int l = 0; l < nlev; ++l) a.dist_coef[l] = dist_coef[l];
  a.out4 = out4; a.n_rate = n_rate; a.rate_coef = rate_coef; a.lambda_scale = lambda_scale; a.extra = extra;
  a.partial = red_scratch(s); a.counter = red_count(s) + 62;
  LMSFC_LAUNCH(mse_multi_kernel, dim3(grid), dim3(TPB), 0, s, a);
  return LMSFC_OK;
}
int mse_fwd_bwd_multi(int nlev, const void* const* xhat, const void* const* x, const int64_t* n, const float* grad_coef,
                      float* sse, void* const* grad, cudaStream_t s) {
  MseMultiArgs a;
  int grid = 0;
  if (int rc = mse_multi_fill(a, nlev, xhat, x, n, grad_coef, grad, &grid)) return rc;
  a.sse = sse;
  LMSFC_LAUNCH(mse_multi_kernel, dim3(grid), dim3(TPB), 0, s, a);
  return LMSFC_OK;
}
static int mse_multi_fill(MseMultiArgs& a, int nlev, const void* const* xhat, const void* const* x, const int64_t* n,
                          const float* grad_coef, void* const* grad, int* grid) {
  LMSFC_CHECK_ARG(nlev >= 1 && nlev <= kMseMaxLevels, "mse_multi: %d levels (max %d)", nlev, kMseMaxLevels);
  memset(&a, 0, sizeof(a));
  int64_t total8 = 0;
  for (int l = 0; l < nlev; ++l) {
    LMSFC_CHECK_ARG(n[l] > 0 && n[l] % 8 == 0 && xhat[l] && x[l], "mse_multi: level %d n=%lld", l, (long long)n[l]);
    total8 += n[l] / 8;
  }
  const int budget = 148 * 8;      // blocks shared out in proportion to the level sizes, at least one each
  int end = 0;
  for (int l = 0; l < nlev; ++l) {
    a.xh[l] = (const uint4*)xhat[l]; a.x[l] = (const uint4*)x[l]; a.grad[l] = (uint4*)(grad ? grad[l] : nullptr);
    a.n8[l] = n[l] / 8; a.coef[l] = grad_coef[l];
    int64_t want = (a.n8[l] + TPB - 1) / TPB, share = (int64_t)budget * a.n8[l] / total8;
    end += (int)max((int64_t)1, min(want, max(share, (int64_t)1)));
    a.blk_end[l] = end;
  }
  a.nlev = nlev;
  *grid = end;
  return LMSFC_OK;
}

// out = {bpp, distortion, loss}: bpp = sum_{i<n_rate} accum[i]*coef[i];
// D = sum_{n_rate<=i<n_rate+n_dist} accum[i]*coef[i]; loss = lambda_scale*D + bpp.
__global__ void loss_finalize_kernel(const float* accum, const float* coef, int n_rate, int n_dist, float lambda_scale,
                                     float* out) {
  pdl_sync();
  float bpp = 0.f, d = 0.f;
  for (int i = 0; i < n_rate; ++i) bpp += accum[i] * coef[i];
  for (int i = 0; i < n_dist; ++i) d += accum[n_rate + i] * coef[n_rate + i];
  out[0] = bpp; out[1] = d; out[2] = lambda_scale * d + bpp;
}
int loss_finalize(const float* accum, const float* coef, int n_rate, int n_dist, float lambda_scale, float* out,
                  cudaStream_t s) {
  LMSFC_LAUNCH(loss_finalize_kernel, dim3(1), dim3(1), 0, s, accum, coef, n_rate, n_dist, lambda_scale, out);
  return LMSFC_OK;
}


// Two-stage column reduction (same-address float atomics run at ~1 lane-op/cycle chip-wide, far
// below HBM speed): every block parks its 256 partial sums in `scratch`, the last block to arrive
// (device counter, self-resetting) adds the column totals to `out`.
__device__ __forceinline__ void colsum_finish(const float (&acc)[8], int c8, int C8, float* out, float* scratch,
                                              unsigned int* counter) {
  __shared__ __align__(16) float red[8][32][9];
  __shared__ unsigned int is_last;
  const int C = C8 * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = acc[j];
  __syncthreads();
  if (threadIdx.y == 0 && c8 < C8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float v = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) v += red[k][threadIdx.x][j];
      __stcg(scratch + (size_t)blockIdx.y * C + c8 * 8 + j, v);
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    const unsigned int prev = atomicAdd(counter + blockIdx.x, 1u);
    is_last = (prev == gridDim.y - 1) ? 1u : 0u;
    if (is_last) counter[blockIdx.x] = 0;
  }
  __syncthreads();
  if (is_last) {
    // 256 threads = 4 row slices x 64 float4 column groups of this block's 256 columns; 8 independent
    // L2 loads in flight per thread (a serial loop over the partial rows was latency-bound)
    __threadfence();
    const int tid = threadIdx.y * 32 + threadIdx.x;
    const int q = tid & 63, sl = tid >> 6;
    const int col = blockIdx.x * 256 + q * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (col < C) {
      const int nb = (int)gridDim.y;
#pragma unroll 8
      for (int b2 = sl; b2 < nb; b2 += 4) {
        const float4 t = __ldcg(reinterpret_cast<const float4*>(scratch + (size_t)b2 * C + col));
        v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
      }
    }
    float4* fin = reinterpret_cast<float4*>(&red[0][0][0]);      // reuse (all reads of red are done)
    fin[tid] = v;
    __syncthreads();
    if (sl == 0 && col < C) {
#pragma unroll
      for (int k = 1; k < 4; ++k) {
        const float4 t = fin[k * 64 + q];
        v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
      }
      float4* o = reinterpret_cast<float4*>(out + col);
      float4 cur = *o;
      cur.x += v.x; cur.y += v.y; cur.z += v.z; cur.w += v.w;
      *o = cur;
    }
  }
}

// One scratch slot per CUDA stream: kernels of one stream serialise, kernels of different streams (the
// weight-gradient side streams, the hyperprior side stream) may overlap and must not share partial sums.
// The table is guarded by a mutex (the autograd worker thread and the main thread both launch); when every slot
// is taken, the least recently used one is handed over -- its counters reset themselves after every launch, so a
// slot whose stream is idle is reusable, and 64 slots cover torch's whole stream pool (32 handles per priority).
// All slots are allocated up front: cudaMalloc is not allowed while a stream is being captured.
constexpr int kRedBlocksY = 296, kRedMaxC = 1024, kRedSlots = 64;
float* g_red_scratch = nullptr;
unsigned int* g_red_count = nullptr;
cudaStream_t g_red_streams[kRedSlots];
unsigned long long g_red_stamp[kRedSlots];
unsigned long long g_red_clock = 0;
int g_red_nstreams = 0;
std::mutex g_red_mutex;
int ensure_reduce_scratch() {
  std::lock_guard<std::mutex> lock(g_red_mutex);
  if (g_red_scratch) return LMSFC_OK;
  float* scratch = nullptr;
  LMSFC_CUDA(cudaMalloc(&scratch, sizeof(float) * kRedSlots * kRedBlocksY * kRedMaxC));
  LMSFC_CUDA(cudaMalloc(&g_red_count, sizeof(unsigned int) * kRedSlots * 64));
  LMSFC_CUDA(cudaMemset(g_red_count, 0, sizeof(unsigned int) * kRedSlots * 64));
  LMSFC_CUDA(cudaDeviceSynchronize());
  g_red_scratch = scratch;
  return LMSFC_OK;
}
int red_slot(cudaStream_t s) {
  std::lock_guard<std::mutex> lock(g_red_mutex);
  ++g_red_clock;
  for (int i = 0; i < g_red_nstreams; ++i)
    if (g_red_streams[i] == s) { g_red_stamp[i] = g_red_clock; return i; }
  int slot = g_red_nstreams;
  if (g_red_nstreams < kRedSlots) {
    ++g_red_nstreams;
  } else {
    slot = 0;
    for (int i = 1; i < kRedSlots; ++i)
      if (g_red_stamp[i] < g_red_stamp[slot]) slot = i;
  }
  g_red_streams[slot] = s;
  g_red_stamp[slot] = g_red_clock;
  return slot;
}
float* red_scratch(cudaStream_t s) { return g_red_scratch + (size_t)red_slot(s) * kRedBlocksY * kRedMaxC; }
unsigned int* red_count(cudaStream_t s) { return g_red_count + red_slot(s) * 64; }

// ---------------------------------------------------------------- GDN helper passes
__global__ void __launch_bounds__(TPB) square_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int64_t n8) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < n8; i += (int64_t)gridDim.x * TPB) {
    const uint4 a = __ldg(x + i);
    const uint32_t av[4] = {a.x, a.y, a.z, a.w};
    uint32_t r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = pack_bf16(lo(av[j]) * lo(av[j]), hi(av[j]) * hi(av[j]));
    y[i] = make_uint4(r[0], r[1], r[2], r[3]);
  }
}
int square_bf16(const __nv_bfloat16* x, __nv_bfloat16* x2, int64_t n, cudaStream_t s) {
  LMSFC_CHECK_ARG(n % 8 == 0, "square: n must be a multiple of 8");
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(square_kernel, dim3(grid1d(n / 8)), dim3(TPB), 0, s, (const uint4*)x, (uint4*)x2, n / 8);
  return LMSFC_OK;
}

// t = -1/2 gy x rn^3 (GDN) | +1/2 gy x rn (IGDN), bf16; g_beta[c] += sum_pix t.
// Block = 32 x 8 threads: threadIdx.x covers 8 channels (one uint4), threadIdx.y strides pixels.
__global__ void __launch_bounds__(256) gdn_t_kernel(const uint4* __restrict__ gy, const uint4* __restrict__ x,
                                                    const uint4* __restrict__ rn, int inverse, int64_t P, int C8,
                                                    uint4* __restrict__ t, float* g_beta, float* scratch,
                                                    unsigned int* counter) {
  pdl_sync();
  const int c8 = blockIdx.x * 32 + threadIdx.x;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c8 < C8) {
#pragma unroll 4
    for (int64_t p = (int64_t)blockIdx.y * 8 + threadIdx.y; p < P; p += (int64_t)gridDim.y * 8) {
      const int64_t i = p * C8 + c8;
      const uint4 g = __ldg(gy + i), xv = __ldg(x + i), r = __ldg(rn + i);
      const uint32_t gv[4] = {g.x, g.y, g.z, g.w}, xx[4] = {xv.x, xv.y, xv.z, xv.w}, rv[4] = {r.x, r.y, r.z, r.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float r0 = lo(rv[j]), r1 = hi(rv[j]);
        const float f0 = inverse ? 0.5f * r0 : -0.5f * r0 * r0 * r0;
        const float f1 = inverse ? 0.5f * r1 : -0.5f * r1 * r1 * r1;
        const float t0 = lo(gv[j]) * lo(xx[j]) * f0, t1 = hi(gv[j]) * hi(xx[j]) * f1;
        acc[2 * j] += t0; acc[2 * j + 1] += t1;
        o[j] = pack_bf16(t0, t1);
      }
      t[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
  if (g_beta) colsum_finish(acc, c8, C8, g_beta, scratch, counter);   // null: the caller sums t's columns elsewhere
}
int gdn_t(const __nv_bfloat16* gy, const __nv_bfloat16* x, const __nv_bfloat16* rn, int inverse, int64_t P, int C,
          __nv_bfloat16* t, float* g_beta, cudaStream_t s) {
  LMSFC_CHECK_ARG(C % 8 == 0, "gdn_t: C must be a multiple of 8");
  LMSFC_CHECK_ARG(C <= kRedMaxC, "column reduction: C=%d too large", C);
  if (int rc = ensure_reduce_scratch()) return rc;
  if (P == 0) return LMSFC_OK;
  const int C8 = C / 8, bx = (C8 + 31) / 32;
  int64_t by = (148 * 2 + bx - 1) / bx;   // ~2 blocks per SM: few same-address atomics, 4 loads in flight per thread
  const int64_t maxy = (P + 7) / 8;
  if (by > maxy) by = maxy;
  LMSFC_LAUNCH(gdn_t_kernel, dim3(dim3(bx, (unsigned)by)), dim3(dim3(32, 8)), 0, s, (const uint4*)gy, (const uint4*)x, (const uint4*)rn,
                                                                inverse, P, C8, (uint4*)t, g_beta, red_scratch(s),
                                                                red_count(s));
  return LMSFC_OK;
}

// column sums of a bf16 [P][C] matrix (conv bias gradients)
__global__ void __launch_bounds__(256) colsum_kernel(const uint4* __restrict__ g, int64_t P, int C8, float* out,
                                                     float* scratch, unsigned int* counter) {
  pdl_sync();
  const int c8 = blockIdx.x * 32 + threadIdx.x;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c8 < C8) {
#pragma unroll 8
    for (int64_t p = (int64_t)blockIdx.y * 8 + threadIdx.y; p < P; p += (int64_t)gridDim.y * 8) {
      const uint4 a = __ldg(g + p * C8 + c8);
      const uint32_t av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) { acc[2 * j] += lo(av[j]); acc[2 * j + 1] += hi(av[j]); }
    }
  }
  colsum_finish(acc, c8, C8, out, scratch, counter);
}
int colsum_bf16(const __nv_bfloat16* g, int64_t P, int C, float* out, cudaStream_t s) {
  LMSFC_CHECK_ARG(C % 8 == 0, "colsum: C must be a multiple of 8");
  LMSFC_CHECK_ARG(C <= kRedMaxC, "column reduction: C=%d too large", C);
  LMSFC_CHECK_ARG(((uintptr_t)out & 15) == 0, "colsum: out must be 16-byte aligned");
  if (int rc = ensure_reduce_scratch()) return rc;
  if (P == 0) return LMSFC_OK;
  const int C8 = C / 8, bx = (C8 + 31) / 32;
  int64_t by = (148 * 2 + bx - 1) / bx;   // ~2 blocks per SM, 8 x 16 B loads in flight per thread
  const int64_t maxy = (P + 7) / 8;
  if (by > maxy) by = maxy;
  LMSFC_LAUNCH(colsum_kernel, dim3(dim3(bx, (unsigned)by)), dim3(dim3(32, 8)), 0, s, (const uint4*)g, P, C8, out, red_scratch(s), red_count(s));
  return LMSFC_OK;
}

// ---------------------------------------------------------------- column sums without any allocation
// lmsfc_colsum_static (include/b200lmsfc.h): the colsum_kernel above with its partial-sum scratch and arrival counters
// in file-scope device arrays instead of the cudaMalloc'ed per-stream slots, for callers that may not allocate (stream
// capture, serving loops).  Every GPU that loads the library holds the 1.16 MiB + 256 B until exit.  In host code the
// array names are host-side stand-ins, not device addresses: the addresses the kernel gets come from
// cudaGetSymbolAddress, looked up once per device with that device current.  All calls share the arrays.
constexpr int kStaticBlocksY = kRedBlocksY, kStaticMaxC = kRedMaxC, kStaticMaxDevices = 64;
static __device__ __align__(256) float g_static_colsum_scratch[kStaticBlocksY * kStaticMaxC];
static __device__ __align__(256) unsigned int g_static_colsum_count[64];
namespace {
struct StaticColsumAddr { float* scratch; unsigned int* count; };
StaticColsumAddr g_static_addr[kStaticMaxDevices] = {};
std::mutex g_static_mutex;
}  // namespace
int colsum_bf16_static(const __nv_bfloat16* g, int64_t P, int C, float* out, cudaStream_t s) {
  LMSFC_CHECK_ARG(g != nullptr && out != nullptr, "colsum_static: null pointer");
  LMSFC_CHECK_ARG(P >= 0, "colsum_static: P=%lld is negative", (long long)P);
  LMSFC_CHECK_ARG(C > 0 && C % 8 == 0, "colsum_static: C=%d must be a positive multiple of 8", C);
  LMSFC_CHECK_ARG(C <= kStaticMaxC, "colsum_static: C=%d exceeds the static scratch (at most %d columns)", C, kStaticMaxC);
  LMSFC_CHECK_ARG(((uintptr_t)out & 15) == 0, "colsum_static: out must be 16-byte aligned");
  if (P == 0) return LMSFC_OK;
  int dev = -1;
  LMSFC_CUDA(cudaGetDevice(&dev));
  LMSFC_CHECK_ARG(dev >= 0 && dev < kStaticMaxDevices, "colsum_static: device ordinal %d not supported", dev);
  StaticColsumAddr a;
  {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[dev].scratch) {            // set-up, once per device, with that device current
      float* sp = nullptr;
      unsigned int* cp = nullptr;
      LMSFC_CUDA(cudaGetSymbolAddress((void**)&sp, g_static_colsum_scratch));
      LMSFC_CUDA(cudaGetSymbolAddress((void**)&cp, g_static_colsum_count));
      g_static_addr[dev].count = cp;
      g_static_addr[dev].scratch = sp;
    }
    a = g_static_addr[dev];
  }
  const int C8 = C / 8, bx = (C8 + 31) / 32;
  int64_t by = (148 * 2 + bx - 1) / bx;
  const int64_t maxy = (P + 7) / 8;
  if (by > maxy) by = maxy;
  // the counters must start at zero whatever an earlier (possibly failed) call left behind; the partial sums need no
  // clearing, every row the last block reads is written by this launch first
  LMSFC_CUDA(cudaMemsetAsync(a.count, 0, sizeof(unsigned int) * 64, s));
  LMSFC_LAUNCH(colsum_kernel, dim3(dim3(bx, (unsigned)by)), dim3(dim3(32, 8)), 0, s, (const uint4*)g, P, C8, out, a.scratch,
               a.count);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- NonNegativeParametrizer
__global__ void reparam_fwd_kernel(const float* p, int64_t n, float bound, float ped, float* o32, __nv_bfloat16* o16) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = fmaxf(p[i], bound);
    const float r = v * v - ped;
    if (o32) o32[i] = r;
    if (o16) o16[i] = __float2bfloat16(r);
  }
}
int reparam_fwd(const float* p, int64_t n, float bound, float pedestal, float* out_f32, __nv_bfloat16* out_bf16,
                cudaStream_t s) {
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(reparam_fwd_kernel, dim3(grid1d(n)), dim3(TPB), 0, s, p, n, bound, pedestal, out_f32, out_bf16);
  return LMSFC_OK;
}
__global__ void reparam_bwd_kernel(const float* g_out, const float* p, int64_t n, float bound, float* g_p) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = p[i];
    const float g = g_out[i] * 2.f * fmaxf(v, bound);
    g_p[i] += (v >= bound || g < 0.f) ? g : 0.f;
  }
}
int reparam_bwd(const float* g_out, const float* p, int64_t n, float bound, float* g_p, cudaStream_t s) {
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(reparam_bwd_kernel, dim3(grid1d(n)), dim3(TPB), 0, s, g_out, p, n, bound, g_p);
  return LMSFC_OK;
}

// Batched form over a table of segments (all GDN layers of a model in ONE launch per direction).
// table row s = {param offset, n, out_f32 offset | -1, out_bf16 offset | -1, g_eff offset}.
__global__ void reparam_fwd_multi_kernel(const float* params, const int64_t* table, const float* bounds, float ped,
                                         float* o32, __nv_bfloat16* o16) {
  pdl_sync();
  const int64_t* t = table + 5 * blockIdx.y;
  const float* p = params + t[0];
  const int64_t n = t[1], a32 = t[2], a16 = t[3];
  const float bound = bounds[blockIdx.y];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = fmaxf(p[i], bound);
    const float r = v * v - ped;
    if (a32 >= 0) o32[a32 + i] = r;
    if (a16 >= 0) o16[a16 + i] = __float2bfloat16(r);
  }
}
__global__ void reparam_bwd_multi_kernel(const float* g_eff, const float* params, const int64_t* table,
                                         const float* bounds, float* g_params) {
  pdl_sync();
  const int64_t* t = table + 5 * blockIdx.y;
  const float* p = params + t[0];
  float* gp = g_params + t[0];
  const float* ge = g_eff + t[4];
  const int64_t n = t[1];
  const float bound = bounds[blockIdx.y];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = p[i];
    const float g = ge[i] * 2.f * fmaxf(v, bound);
    gp[i] += (v >= bound || g < 0.f) ? g : 0.f;
  }
}
int reparam_fwd_multi(const float* params, const int64_t* table, const float* bounds, int nseg, int64_t max_n,
                      float pedestal, float* out_f32, __nv_bfloat16* out_bf16, cudaStream_t s) {
  if (nseg == 0) return LMSFC_OK;
  LMSFC_LAUNCH(reparam_fwd_multi_kernel, dim3(dim3(grid1d(max_n), nseg)), dim3(TPB), 0, s, params, table, bounds, pedestal, out_f32, out_bf16);
  return LMSFC_OK;
}
int reparam_bwd_multi(const float* g_eff, const float* params, const int64_t* table, const float* bounds, int nseg,
                      int64_t max_n, float* g_params, cudaStream_t s) {
  if (nseg == 0) return LMSFC_OK;
  LMSFC_LAUNCH(reparam_bwd_multi_kernel, dim3(dim3(grid1d(max_n), nseg)), dim3(TPB), 0, s, g_eff, params, table, bounds, g_params);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- small activation gradients
// slope 0: ReLU; slope > 0: LeakyReLU (the sign of the output is the sign of the input)
__global__ void relu_bwd_kernel(__nv_bfloat16* g, const __nv_bfloat16* out, int64_t n, float slope) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    if (!(__bfloat162float(out[i]) > 0.f)) g[i] = __float2bfloat16(slope * __bfloat162float(g[i]));
}
int relu_bwd_bf16(__nv_bfloat16* g, const __nv_bfloat16* out, int64_t n, float slope, cudaStream_t s) {
  LMSFC_CHECK_ARG(slope >= 0.f && slope < 1.f, "relu_backward: slope %g outside [0, 1)", (double)slope);
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(relu_bwd_kernel, dim3(grid1d(n)), dim3(TPB), 0, s, g, out, n, slope);
  return LMSFC_OK;
}
// g_out = bf16(g_direct + sign(y) * g_abs): gradient of y feeding both GaussianConditional and h_a(|y|)
// relu != 0: y is a ReLU output, both branches are masked by (y > 0)
__global__ void abs_bwd_kernel(const float* gd, const __nv_bfloat16* ga, const float* y, int64_t n, int relu,
                               __nv_bfloat16* o) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float yv = y[i];
    const float sg = (relu & 2) ? 1.f : ((yv > 0.f) ? 1.f : ((yv < 0.f) ? -1.f : 0.f));   // bit 1: second output was y, not |y|
    const float d = ((relu & 1) && !(yv > 0.f)) ? 0.f : (gd ? gd[i] : 0.f);
    o[i] = __float2bfloat16(d + (ga ? sg * __bfloat162float(ga[i]) : 0.f));
  }
}
int abs_bwd(const float* g_direct, const __nv_bfloat16* g_abs, const float* y, int64_t n, int relu, __nv_bfloat16* g_out,
            cudaStream_t s) {
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(abs_bwd_kernel, dim3(grid1d(n)), dim3(TPB), 0, s, g_direct, g_abs, y, n, relu, g_out);
  return LMSFC_OK;
}

__global__ void cast_f32_bf16_kernel(const float* in, __nv_bfloat16* out, int64_t n) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16(in[i]);
}
int cast_f32_bf16(const float* in, __nv_bfloat16* out, int64_t n, cudaStream_t s) {
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(cast_f32_bf16_kernel, dim3(grid1d(n)), dim3(TPB), 0, s, in, out, n);
  return LMSFC_OK;
}
__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* in, float* out, int64_t n) {
  pdl_sync();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = __bfloat162float(in[i]);
}
int cast_bf16_f32(const __nv_bfloat16* in, float* out, int64_t n, cudaStream_t s) {
  if (n == 0) return LMSFC_OK;
  LMSFC_LAUNCH(cast_bf16_f32_kernel, dim3(grid1d(n)), dim3(TPB), 0, s, in, out, n);
  return LMSFC_OK;
}

}  // namespace lmsfc
