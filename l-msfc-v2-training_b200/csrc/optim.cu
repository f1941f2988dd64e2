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
// Deterministic: block partials are parked in scratch and the LAST block to arrive adds them up in index order, so two
// ranks holding the same (all-reduced) gradient compute bit-identical norms -> identical clip coefficients -> replicas
// never drift apart (float atomics made the norm differ in the last bits from run to run and rank to rank).
__global__ void __launch_bounds__(TPB) sumsq_kernel(const float4* __restrict__ g, int64_t n4, const float* tail, int ntail,
                                                    float* accum, float* partial, unsigned int* counter, int* set_flag,
                                                    int overwrite, const float* parts, int nparts, float* total) {
  pdl_sync();
  float local = 0.f;
#pragma unroll 8
  for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < n4; i += (int64_t)gridDim.x * TPB) {
    const float4 v = __ldg(g + i);
    local += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < ntail) local += tail[threadIdx.x] * tail[threadIdx.x];
  __shared__ float red[32];
  __shared__ unsigned int is_last;
  local = block_sum(local, red);
  if (threadIdx.x == 0) {
    __stcg(partial + blockIdx.x, local);
    __threadfence();
    const unsigned int prev = atomicAdd(counter, 1u);
    is_last = (prev == gridDim.x - 1) ? 1u : 0u;
    if (is_last) *counter = 0;
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float v = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += TPB) v += __ldcg(partial + i);   // fixed order per thread
    v = block_sum(v, red);                                                               // fixed tree
    if (threadIdx.x == 0) {
      if (overwrite) *accum = v; else *accum += v;
      if (parts) {                      // the last slice of a sliced norm also adds the parts up, in index order
        float t = 0.f;
        for (int i = 0; i < nparts; ++i) t += (parts + i == accum) ? *accum : __ldcg(parts + i);
        *total = t;
      }
      if (set_flag) *set_flag = 1;      // "an update is pending" (pipelined train step), consumed by adam_prep
    }
  }
}
// out = parts[0] + parts[1] + ... in index order (the slices' squared norms -> the whole gradient's)
__global__ void sum_parts_kernel(const float* parts, int n, float* out, int* set_flag) {
  pdl_sync();
  float s = 0.f;
  for (int i = 0; i < n; ++i) s += parts[i];
  *out = s;
  if (set_flag) *set_flag = 1;
}
int sum_parts(const float* parts, int n, float* out, int* set_flag, cudaStream_t s) {
  LMSFC_LAUNCH(sum_parts_kernel, dim3(1), dim3(1), 0, s, parts, n, out, set_flag);
  return LMSFC_OK;
}
int sumsq_f32(const float* g, int64_t n, float* accum, int* set_flag, int overwrite, cudaStream_t s, const float* parts,
              int nparts, float* total) {
  LMSFC_CHECK_ARG(n > 0 || parts == nullptr, "sumsq: the closing slice of a sliced norm cannot be empty");
  if (n == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG(((uintptr_t)g & 15) == 0, "sumsq: pointer must be 16-byte aligned");
  if (int rc = ensure_reduce_scratch()) return rc;
  const int64_t n4 = n / 4;
  // >= 8 float4 per thread: the slices of the sliced norm (2-3 M gradients each) sit on the step's tail, and a grid of a
  // thousand blocks with one or two loads each spends its time in the last-block hand-over, not in HBM
  const int grid = grid1d((n4 + 7) / 8 > 0 ? (n4 + 7) / 8 : 1);
  LMSFC_LAUNCH(sumsq_kernel, dim3(grid), dim3(TPB), 0, s, (const float4*)g, n4, g + n4 * 4, (int)(n - n4 * 4), accum,
               red_scratch(s), red_count(s) + 63, set_flag, overwrite, parts, nparts, total);
  return LMSFC_OK;
}

// ---------------------------------------------------------------- Adam
// prep: one thread bumps the step counter and derives {clip_coef, bias_corr1, bias_corr2, norm, lr, active}.
// The learning rate is read from device memory when lr_dev is given, so a captured graph follows an LR schedule
// (torch.optim.lr_scheduler writes param_groups[..]["lr"]; the host mirrors it into lr_dev between replays).
// `pending` (nullable): consume the "an update is pending" flag -- when it is 0 nothing is applied (no step bump, the
// element-wise pass returns at once): the replay that follows a flush() of the pipelined train step.  `gate`
// (nullable): take "active" from another optimiser's scratch[5] instead (the aux optimiser follows the net one).
__global__ void adam_prep_kernel(AdamArgs a) {
  pdl_sync();
  float active = 1.f;
  if (a.pending) {
    active = (a.force || *a.pending != 0) ? 1.f : 0.f;     // force: an eager step applies its own backward pass
    *a.pending = 0;
  } else if (a.gate) {
    active = *a.gate;
  }
  a.scratch[4] = a.lr_dev ? *a.lr_dev : a.lr;
  a.scratch[5] = active;
  if (active == 0.f) return;
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
__device__ __forceinline__ void adam_one(float g, float& m, float& v, float& p, float clip, float b1, float b2,
                                         float step_size, float rs_bc2, float eps) {
  g *= clip;
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  p = p - step_size * m / (sqrtf(v) * rs_bc2 + eps);
}
// 30 B of HBM traffic per parameter (34 with zero_g).  2 CTAs of 256 threads per SM, 2 x 4 independent 16-byte
// loads per thread in flight: a small footprint on purpose -- in the pipelined step this kernel shares the SMs with
// the forward convolutions (a grid that fills every thread slot would hold their CTAs back until it ends).
__device__ __forceinline__ void adam_quad(const float4& g, float4& m, float4& v, float4& p, float clip, float b1, float b2,
                                          float step_size, float rs_bc2, float eps) {
  adam_one(g.x, m.x, v.x, p.x, clip, b1, b2, step_size, rs_bc2, eps);
  adam_one(g.y, m.y, v.y, p.y, clip, b1, b2, step_size, rs_bc2, eps);
  adam_one(g.z, m.z, v.z, p.z, clip, b1, b2, step_size, rs_bc2, eps);
  adam_one(g.w, m.w, v.w, p.w, clip, b1, b2, step_size, rs_bc2, eps);
}
template <int TPB>
__device__ __forceinline__ void adam_body(const AdamArgs& a) {
  pdl_sync();
  if (a.scratch[5] == 0.f) return;        // no update pending (see adam_prep_kernel)
  const float clip = a.scratch[0], bc1 = a.scratch[1], bc2 = a.scratch[2];
  const float step_size = a.scratch[4] / bc1, rs_bc2 = rsqrtf(bc2);
  const int64_t n4 = a.n >> 2, stride = (int64_t)gridDim.x * TPB;
  float4* __restrict__ p4 = reinterpret_cast<float4*>(a.p);
  float4* __restrict__ m4 = reinterpret_cast<float4*>(a.m);
  float4* __restrict__ v4 = reinterpret_cast<float4*>(a.v);
  float4* __restrict__ g4 = reinterpret_cast<float4*>(a.g);
  uint2* __restrict__ s4 = reinterpret_cast<uint2*>(a.p_bf16);
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  for (; i + stride < n4; i += 2 * stride) {
    const int64_t j = i + stride;
    const float4 g0 = __ldcs(g4 + i), g1 = __ldcs(g4 + j);
    float4 m0 = __ldcs(m4 + i), m1 = __ldcs(m4 + j);
    float4 v0 = __ldcs(v4 + i), v1 = __ldcs(v4 + j);
    float4 p0 = __ldcs(p4 + i), p1 = __ldcs(p4 + j);
    adam_quad(g0, m0, v0, p0, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
    adam_quad(g1, m1, v1, p1, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
    __stcs(m4 + i, m0); __stcs(v4 + i, v0); __stcs(p4 + i, p0);
    __stcs(m4 + j, m1); __stcs(v4 + j, v1); __stcs(p4 + j, p1);
    if (a.p_bf16) {                                            // read again by the next forward: default caching
      s4[i] = make_uint2(pack_bf16(p0.x, p0.y), pack_bf16(p0.z, p0.w));
      s4[j] = make_uint2(pack_bf16(p1.x, p1.y), pack_bf16(p1.z, p1.w));
    }
    if (a.zero_g) { g4[i] = zero; g4[j] = zero; }
  }
  if (i < n4) {
    const float4 g0 = __ldcs(g4 + i);
    float4 m0 = __ldcs(m4 + i), v0 = __ldcs(v4 + i), p0 = __ldcs(p4 + i);
    adam_quad(g0, m0, v0, p0, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
    __stcs(m4 + i, m0); __stcs(v4 + i, v0); __stcs(p4 + i, p0);
    if (a.p_bf16) s4[i] = make_uint2(pack_bf16(p0.x, p0.y), pack_bf16(p0.z, p0.w));
    if (a.zero_g) g4[i] = zero;
  }
  if (blockIdx.x == 0 && (int64_t)threadIdx.x < a.n - n4 * 4) {   // scalar tail
    const int64_t i = n4 * 4 + threadIdx.x;
    float m = a.m[i], v = a.v[i], p = a.p[i];
    adam_one(a.g[i], m, v, p, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
    a.m[i] = m; a.v[i] = v; a.p[i] = p;
    if (a.p_bf16) a.p_bf16[i] = __float2bfloat16(p);
    if (a.zero_g) a.g[i] = 0.f;
  }
}
__global__ void __launch_bounds__(256, 4) adam_kernel(AdamArgs a) { adam_body<256>(a); }
// ---------------------------------------------------------------- Adam as a bulk-copy pipeline (default for large slices)
// One persistent CTA per SM whose memory traffic is asynchronous end to end: one thread issues cp.async.bulk loads of
// the next chunks of p, m, v, g into a ring of shared-memory stages (mbarrier complete_tx), all threads update a landed
// chunk in place, and the same thread sends it back with cp.async.bulk stores (p, m, v, the bf16 shadow, zeros for g) --
// nothing waits on a load or a store except through the ring.  Chunk = 2048 parameters: 32 KB in, 36 KB out; 4 stages =
// 144 KB + 8 KB of zeros.  Measured on a B200 over 27.9 M parameters (profiles/r2_adam_sms.txt): 5.68 TB/s on 148 SMs
// (adam_kernel: 5.51), and the per-SM ceiling is ~80 GB/s whichever way the bytes move (a register-staged 1024-thread
// CTA: 76): 96 SMs saturate HBM, 20 SMs reach 1.6 TB/s -- confining the optimiser to the SMs a 128-CTA GEMM leaves idle
// is what the step effectively does already, and giving it fewer SMs on purpose only stretches it (LMSFC_ADAM_STREAM=n).
namespace stream_adam {
constexpr int kThreads = 512, kChunk = 2048, kStages = 4;
constexpr int kStageBytes = 4 * kChunk * 4 + kChunk * 2;           // p, m, v, g fp32 + bf16 shadow
constexpr int kSmemBytes = kStages * kStageBytes + kChunk * 4 + 64 + 1024;
}  // namespace stream_adam
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__global__ void __launch_bounds__(stream_adam::kThreads, 1) adam_stream_kernel(AdamArgs a) {
  using namespace stream_adam;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  float* zeros = (float*)(smem + kStages * kStageBytes);
  uint64_t* full = (uint64_t*)(smem + kStages * kStageBytes + kChunk * 4);
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  for (int i = tid; i < kChunk; i += kThreads) zeros[i] = 0.f;
  fence_proxy_async_smem();
  __syncthreads();
  pdl_sync();
  if (a.scratch[5] == 0.f) return;        // no update pending (see adam_prep_kernel)
  const float clip = a.scratch[0], bc1 = a.scratch[1], bc2 = a.scratch[2];
  const float step_size = a.scratch[4] / bc1, rs_bc2 = rsqrtf(bc2);
  const int64_t nchunks = a.n / kChunk;                        // whole chunks go through the ring
  const int64_t mine = blockIdx.x < nchunks ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const uint32_t sbase = smem_u32(smem);
  auto issue = [&](int64_t k) {                                // thread 0: loads of this CTA's k-th chunk
    const int s = (int)(k % kStages);
    const int64_t e = (blockIdx.x + k * gridDim.x) * kChunk;
    const uint32_t d = sbase + s * kStageBytes;
    mbar_arrive_expect_tx(&full[s], 4 * kChunk * 4);
    bulk_load(d, a.p + e, kChunk * 4, &full[s]);
    bulk_load(d + kChunk * 4, a.m + e, kChunk * 4, &full[s]);
    bulk_load(d + 2 * kChunk * 4, a.v + e, kChunk * 4, &full[s]);
    bulk_load(d + 3 * kChunk * 4, a.g + e, kChunk * 4, &full[s]);
  };
  if (tid == 0)
    for (int64_t k = 0; k < kStages - 1 && k < mine; ++k) issue(k);
  for (int64_t k = 0; k < mine; ++k) {
    const int s = (int)(k % kStages);
    mbar_wait(&full[s], (uint32_t)((k / kStages) & 1));
    float4* sp = reinterpret_cast<float4*>(smem + s * kStageBytes);
    float4* sm = sp + kChunk / 4;
    float4* sv = sm + kChunk / 4;
    const float4* sg = sv + kChunk / 4;
    uint2* sh = reinterpret_cast<uint2*>(smem + s * kStageBytes + 4 * kChunk * 4);
#pragma unroll
    for (int q = tid; q < kChunk / 4; q += kThreads) {
      const float4 g = sg[q];
      float4 m = sm[q], v = sv[q], p = sp[q];
      adam_quad(g, m, v, p, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
      sm[q] = m; sv[q] = v; sp[q] = p;
      sh[q] = make_uint2(pack_bf16(p.x, p.y), pack_bf16(p.z, p.w));
    }
    fence_proxy_async_smem();              // generic-proxy writes -> visible to the bulk stores
    __syncthreads();
    if (tid == 0) {
      const int64_t e = (blockIdx.x + k * gridDim.x) * kChunk;
      const uint32_t d = sbase + s * kStageBytes;
      bulk_store(a.p + e, d, kChunk * 4);
      bulk_store(a.m + e, d + kChunk * 4, kChunk * 4);
      bulk_store(a.v + e, d + 2 * kChunk * 4, kChunk * 4);
      if (a.p_bf16) bulk_store(a.p_bf16 + e, d + 4 * kChunk * 4, kChunk * 2);
      if (a.zero_g) bulk_store(a.g + e, smem_u32(zeros), kChunk * 4);
      bulk_commit_group();
      if (k + kStages - 1 < mine) {
        bulk_wait_group_read<1>();         // the stage consumed in the previous iteration has been read out
        issue(k + kStages - 1);
      }
    }
  }
  if (tid == 0) bulk_wait_group_read<0>();
  // what is left of the slice (< one chunk): the last CTA, plain loads and stores
  if (blockIdx.x == gridDim.x - 1) {
    for (int64_t i = nchunks * kChunk + tid; i < a.n; i += kThreads) {
      float m = a.m[i], v = a.v[i], p = a.p[i];
      adam_one(a.g[i], m, v, p, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
      a.m[i] = m; a.v[i] = v; a.p[i] = p;
      if (a.p_bf16) a.p_bf16[i] = __float2bfloat16(p);
      if (a.zero_g) a.g[i] = 0.f;
    }
  }
}
int adam_prep(const AdamArgs& a, cudaStream_t s) {
  LMSFC_LAUNCH(adam_prep_kernel, dim3(1), dim3(1), 0, s, a);
  return LMSFC_OK;
}
int adam_update(const AdamArgs& a, cudaStream_t s) {
  if (a.n == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG((((uintptr_t)a.p | (uintptr_t)a.m | (uintptr_t)a.v | (uintptr_t)a.g) & 15) == 0 &&
                      ((uintptr_t)a.p_bf16 & 7) == 0, "adam: arena slices must be 16-byte aligned");
  // short-lived CTAs (8 quads per thread): in the pipelined step the forward convolutions run on a higher-priority
  // stream and take every SM an Adam CTA leaves (a GEMM CTA cannot join an SM in the L1-heavy split)
  const int64_t quads = (a.n + 3) / 4, want = (quads + 8 * TPB - 1) / (8 * TPB);
  const int grid = (int)max((int64_t)1, want);
  // LMSFC_ADAM_STREAM: number of SMs (= CTAs) of the bulk-copy pipeline; unset: all of them; 0: adam_kernel only
  static const int stream_sms = [] {
    const char* e = getenv("LMSFC_ADAM_STREAM");
    if (e) return atoi(e);
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return sms;
  }();
  // bulk copies want 16-byte aligned global addresses (arena slices are 256-byte aligned; a caller's odd bf16 pointer
  // takes the register-staged kernel)
  if (stream_sms > 0 && a.n >= (int64_t)stream_sms * stream_adam::kChunk && ((uintptr_t)a.p_bf16 & 15) == 0) {
    static const cudaError_t attr = cudaFuncSetAttribute(adam_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                         stream_adam::kSmemBytes);
    LMSFC_CUDA(attr);
    LMSFC_LAUNCH(adam_stream_kernel, dim3(stream_sms), dim3(stream_adam::kThreads), stream_adam::kSmemBytes, s, a);
    return LMSFC_OK;
  }
  LMSFC_LAUNCH(adam_kernel, dim3(grid), dim3(TPB), 0, s, a);
  return LMSFC_OK;
}
int adam_step(const AdamArgs& a, cudaStream_t s) {
  if (a.n == 0) return LMSFC_OK;
  if (int rc = adam_prep(a, s)) return rc;
  return adam_update(a, s);
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
