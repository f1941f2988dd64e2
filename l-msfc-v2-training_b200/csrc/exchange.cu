// Data-parallel gradient exchange as this library's own kernels over NVLink / NVSwitch peer memory (no NCCL on the
// data path).  Replaces nn.DataParallel's per-step replicate / gather (compressai/examples/train.py:64-71, 275-276) and
// the clip + Adam that follows (examples/train.py:100-103) for N ranks:
//
//   every rank                       owner of a sub-slice (1/N of each arena slice)
//   --------------------------------------------------------------------------------------------------------------
//   backward -> local fp32 grads  -> xchg_reduce:  wait until all ranks finished the slice, g = mean_k grads_k (P2P loads,
//                                     fixed rank order), kept in the owner's arena, sum(g^2) reduced deterministically
//                                    xchg_norm_publish / xchg_norm_wait: the N per-rank sums visit every rank, summed in the
//                                     same order everywhere -> one clip coefficient, bit-identical on all ranks
//                                    xchg_adam_push: Adam on the OWNED sub-slice only (1/N of the optimiser's HBM traffic),
//                                     the new bf16 GEMM operands (fp32 for the parameters kernels read in fp32: GDN, entropy
//                                     bottleneck, biases) stored straight into every rank's arena (P2P stores)
//   forward waits per slice       <- xchg_wait on the slice's flag
//
// Synchronisation is by flags in peer-mapped memory: flags[slot][sender] of the RECEIVING rank holds the epoch of the
// sender's latest signal (st.release.sys after __threadfence_system; pollers ld.acquire.sys their own memory).  The epoch
// (`tick`) is a device counter bumped once per step by xchg_norm_publish, so captured CUDA graphs replay the protocol
// without host arguments.  Every spin is bounded (trap after ~4 s) so that a lost rank fails the step instead of hanging.
#include "../../include/b200lmsfc.h"

#include "kernels.cuh"

namespace lmsfc {

namespace {
constexpr int TPB = 256;

struct XP {                      // device-side copy of lmsfc_xchg_peers
  int world, rank;
  float* grads[LMSFC_MAX_RANKS];
  float* params[LMSFC_MAX_RANKS];
  __nv_bfloat16* shadow[LMSFC_MAX_RANKS];
  unsigned long long* flags[LMSFC_MAX_RANKS];
  float* norm_all[LMSFC_MAX_RANKS];
  unsigned long long* tick;
  float* mc_grads;               // NVSwitch multicast views (nullable)
  float* mc_params;
  __nv_bfloat16* mc_shadow;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_tick(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// sum over all ranks of one float4, formed inside the NVSwitch (NVLS): one 16-byte response instead of N
__device__ __forceinline__ float4 mc_ld_reduce(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_b64(void* mc, uint2 v) {          // one store, delivered to every rank
  const unsigned long long b = ((unsigned long long)v.y << 32) | v.x;
  asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(mc), "l"(b) : "memory");
}
__device__ __forceinline__ void mc_st_f4(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// thread k < world waits until rank k's signal in `slot` of this rank's flag block reaches `want`
__device__ __forceinline__ void wait_all(const XP& P, int slot, unsigned long long want) {
  if ((int)threadIdx.x < P.world) {
    const unsigned long long* f = P.flags[P.rank] + slot * LMSFC_MAX_RANKS + threadIdx.x;
    const long long t0 = clock64();
    while (ld_acquire_sys(f) < want) {
      if (clock64() - t0 > (8ll << 30)) {       // ~4 s at 2 GHz
        printf("lmsfc exchange: rank %d timed out in slot %d waiting for rank %d (want %llu, have %llu)\n", P.rank, slot,
               (int)threadIdx.x, want, ld_acquire_sys(f));
        __trap();
      }
      __nanosleep(200);
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void signal_all(const XP& P, int slot, unsigned long long v) {   // threads k < world
  if ((int)threadIdx.x < P.world) st_release_sys(P.flags[threadIdx.x] + slot * LMSFC_MAX_RANKS + P.rank, v);
}

// ---------------------------------------------------------------- reduce-scatter of one slice (owner side)
// "my gradients of this slice are complete" to every peer, then wait for theirs.  ONE block polls (a handful of
// system-scope loads every 200 ns): the reduction grid behind it never spins -- hundreds of polling blocks were measured
// to slow the concurrent host-to-device feature copies by 0.5 ms per step.
__global__ void xchg_ready_kernel(XP P, int slot) {
  pdl_sync();
  const unsigned long long want = ld_tick(P.tick) + 1;
  if ((int)threadIdx.x < P.world) __threadfence_system();   // this rank's gradients of the slice are complete (stream order)
  signal_all(P, slot, want);
  wait_all(P, slot, want);
}

// MC (NVLS) variant: <= 64 registers, so that its blocks fit beside a 256-thread, ~190-register GEMM CTA on the same SM
// (the peer-load variant keeps world x 4 float4 in flight per thread and does not).
template <bool MC>
__global__ void __launch_bounds__(TPB, MC ? 4 : 1) xchg_reduce_kernel(XP P, int slot, long long lo4, long long n4, float inv_world,
                                                                      float* partial, unsigned int* counter, float* norm_part) {
  pdl_sync();                                    // behind xchg_ready_kernel: every rank's gradients of the slice are complete
  float local = 0.f;
  float4* mine = reinterpret_cast<float4*>(P.grads[P.rank]) + lo4;
  // NVLink / HBM latency is hidden by loads in flight: U quads per thread and iteration, world loads each, all issued
  // before the first use (world = 8: 32 independent 16-byte loads per thread)
  constexpr int U = 4;
  const long long stride = (long long)gridDim.x * TPB;
  if (MC) {
    // NVLS: the switch adds the N ranks' values; this rank receives only the sums of its own part
    const float4* mc = reinterpret_cast<const float4*>(P.mc_grads) + lo4;
    constexpr int UM = 8;
    for (long long i0 = (long long)blockIdx.x * TPB + threadIdx.x; i0 < n4; i0 += UM * stride) {
      float4 v[UM];
#pragma unroll
      for (int u = 0; u < UM; ++u) {
        const long long i = i0 + u * stride;
        if (i < n4) v[u] = mc_ld_reduce(mc + i);
      }
#pragma unroll
      for (int u = 0; u < UM; ++u) {
        const long long i = i0 + u * stride;
        if (i >= n4) break;
        float4 a = v[u];
        a.x *= inv_world; a.y *= inv_world; a.z *= inv_world; a.w *= inv_world;
        mine[i] = a;
        local += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
      }
    }
  } else
  for (long long i0 = (long long)blockIdx.x * TPB + threadIdx.x; i0 < n4; i0 += U * stride) {
    float4 v[U][LMSFC_MAX_RANKS];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
#pragma unroll
      for (int k = 0; k < LMSFC_MAX_RANKS; ++k)
        if (k < P.world && i < n4) v[u][k] = __ldcg(reinterpret_cast<const float4*>(P.grads[k]) + lo4 + i);   // L1 is not coherent with peers
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= n4) break;
      float4 a = v[u][0];
#pragma unroll
      for (int k = 1; k < LMSFC_MAX_RANKS; ++k)
        if (k < P.world) { a.x += v[u][k].x; a.y += v[u][k].y; a.z += v[u][k].z; a.w += v[u][k].w; }   // fixed rank order
      a.x *= inv_world; a.y *= inv_world; a.z *= inv_world; a.w *= inv_world;
      mine[i] = a;
      local += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
    }
  }
  // deterministic sum(g^2): block partials parked, the last block adds them in index order
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
    for (int i = threadIdx.x; i < (int)gridDim.x; i += TPB) v += __ldcg(partial + i);
    v = block_sum(v, red);
    if (threadIdx.x == 0) *norm_part = v;
  }
}

// ---------------------------------------------------------------- gradient norm: N numbers visit every rank
__global__ void xchg_norm_publish_kernel(XP P, const float* parts, int nparts, int slot, int* set_flag) {
  pdl_sync();
  const unsigned long long next = ld_tick(P.tick) + 1;
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < nparts; ++i) s += parts[i];
    for (int k = 0; k < P.world; ++k) __stcg(P.norm_all[k] + P.rank, s);
    __threadfence_system();
  }
  __syncthreads();
  signal_all(P, slot, next);
  __syncthreads();
  if (threadIdx.x == 0) {
    *P.tick = next;                              // the step's epoch: every later kernel of this step reads it
    if (set_flag) *set_flag = 1;                 // "an update is pending" (pipelined train step), consumed by adam_prep
  }
}
__global__ void xchg_norm_wait_kernel(XP P, int slot, float* sumsq_out) {
  pdl_sync();
  wait_all(P, slot, ld_tick(P.tick));
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int k = 0; k < P.world; ++k) s += __ldcg(P.norm_all[P.rank] + k);   // same order on every rank
    *sumsq_out = s;
  }
}
__global__ void xchg_wait_kernel(XP P, int slot) {
  pdl_sync();
  wait_all(P, slot, ld_tick(P.tick));
}

// ---------------------------------------------------------------- Adam on the owned sub-slice + all-gather by P2P stores
__device__ __forceinline__ void adam1(float g, float& m, float& v, float& p, float clip, float b1, float b2, float step_size,
                                      float rs_bc2, float eps) {
  g *= clip;
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  p = p - step_size * m / (sqrtf(v) * rs_bc2 + eps);
}
struct PushArgs {
  long long unit_lo4, unit_hi4, own_lo4, own_hi4, fp32_end4;   // in float4 units
  float* m;
  float* v;
  float beta1, beta2, eps;
  const float* scratch;            // {clip, bc1, bc2, norm, lr, active} from adam_prep
  unsigned int* counter;
  int slot;
};
__global__ void __launch_bounds__(TPB, 4) xchg_adam_push_kernel(XP P, PushArgs a) {   // <= 64 registers: fits beside a GEMM CTA
  pdl_sync();
  const unsigned long long ep = ld_tick(P.tick);
  const bool active = a.scratch[5] != 0.f;
  if (active) {
    const float clip = a.scratch[0], bc1 = a.scratch[1], bc2 = a.scratch[2];
    const float step_size = a.scratch[4] / bc1, rs_bc2 = rsqrtf(bc2);
    float4* p4 = reinterpret_cast<float4*>(P.params[P.rank]);
    float4* g4 = reinterpret_cast<float4*>(P.grads[P.rank]);
    float4* m4 = reinterpret_cast<float4*>(a.m);
    float4* v4 = reinterpret_cast<float4*>(a.v);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long stride = (long long)gridDim.x * TPB, t0 = (long long)blockIdx.x * TPB + threadIdx.x;
    auto one = [&](long long i, const float4& g, float4 m, float4 v, float4 p) {
      adam1(g.x, m.x, v.x, p.x, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
      adam1(g.y, m.y, v.y, p.y, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
      adam1(g.z, m.z, v.z, p.z, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
      adam1(g.w, m.w, v.w, p.w, clip, a.beta1, a.beta2, step_size, rs_bc2, a.eps);
      __stcs(m4 + i, m); __stcs(v4 + i, v); __stcs(p4 + i, p);
      g4[i] = zero;
      const uint2 b = make_uint2(pack_bf16(p.x, p.y), pack_bf16(p.z, p.w));
      if (P.mc_shadow) {                         // one multicast store reaches every rank's arena
        mc_st_b64(reinterpret_cast<uint2*>(P.mc_shadow) + i, b);
        if (i < a.fp32_end4) mc_st_f4(reinterpret_cast<float4*>(P.mc_params) + i, p);
        return;
      }
#pragma unroll
      for (int k = 0; k < LMSFC_MAX_RANKS; ++k) {
        if (k >= P.world) break;
        reinterpret_cast<uint2*>(P.shadow[k])[i] = b;                                   // bf16 GEMM operands, every rank
        if (i < a.fp32_end4 && k != P.rank) reinterpret_cast<float4*>(P.params[k])[i] = p;   // parameters read in fp32
      }
    };
    long long i = a.own_lo4 + t0;
    for (; i + stride < a.own_hi4; i += 2 * stride) {          // 2 x 4 independent 16-byte loads in flight per thread
      const long long j = i + stride;
      const float4 g0 = __ldcs(g4 + i), g1 = __ldcs(g4 + j);
      const float4 m0 = __ldcs(m4 + i), m1 = __ldcs(m4 + j);
      const float4 v0 = __ldcs(v4 + i), v1 = __ldcs(v4 + j);
      const float4 q0 = __ldcs(p4 + i), q1 = __ldcs(p4 + j);
      one(i, g0, m0, v0, q0);
      one(j, g1, m1, v1, q1);
    }
    if (i < a.own_hi4) one(i, __ldcs(g4 + i), __ldcs(m4 + i), __ldcs(v4 + i), __ldcs(p4 + i));
    // the rest of the slice belongs to other ranks: they have read this rank's gradients (norm wait), clear them
    for (long long i = a.unit_lo4 + t0; i < a.own_lo4; i += stride) g4[i] = zero;
    for (long long i = a.own_hi4 + t0; i < a.unit_hi4; i += stride) g4[i] = zero;
  }
  // ONE system-scope fence per kernel: blocks publish their stores at GPU scope (barrier + fence + counter), the last
  // block to arrive -- which thereby observes all of them -- issues the cumulative fence.sys in front of the flag stores.
  // A fence.sys per block (let alone per thread) stalls for tens of microseconds while the copy engines keep PCIe busy:
  // measured +0.54 ms per step in the end-to-end loop (profiles/tools/e2e_probe.py).
  __syncthreads();
  __shared__ unsigned int is_last;
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int prev = atomicAdd(a.counter, 1u);
    is_last = (prev == gridDim.x - 1) ? 1u : 0u;
    if (is_last) *a.counter = 0;
  }
  __syncthreads();
  if (is_last) {
    __threadfence_system();
    signal_all(P, a.slot, ep);                   // "my part of this slice has landed everywhere"
  }
}

int to_xp(const lmsfc_xchg_peers* p, XP& x) {
  LMSFC_CHECK_ARG(p && p->world >= 1 && p->world <= LMSFC_MAX_RANKS && p->rank >= 0 && p->rank < p->world && p->tick,
                  "exchange: bad peer table");
  x.world = p->world; x.rank = p->rank; x.tick = (unsigned long long*)p->tick;
  for (int k = 0; k < LMSFC_MAX_RANKS; ++k) {
    const bool in = k < p->world;
    x.grads[k] = in ? p->grads[k] : nullptr;
    x.params[k] = in ? p->params[k] : nullptr;
    x.shadow[k] = in ? (__nv_bfloat16*)p->shadow[k] : nullptr;
    x.flags[k] = in ? (unsigned long long*)p->flags[k] : nullptr;
    x.norm_all[k] = in ? p->norm_all[k] : nullptr;
    LMSFC_CHECK_ARG(!in || (x.grads[k] && x.params[k] && x.shadow[k] && x.flags[k] && x.norm_all[k]), "exchange: rank %d has a null buffer", k);
  }
  x.mc_grads = p->mc_grads;
  x.mc_params = p->mc_params;
  x.mc_shadow = (__nv_bfloat16*)p->mc_shadow;
  LMSFC_CHECK_ARG(!x.mc_shadow == !x.mc_params, "exchange: the multicast views of params and shadow come together");
  return LMSFC_OK;
}

// exchange kernels share their SMs with the GEMMs of the running backward / forward pass: ask for the shared-memory-heavy
// carve-out so that a 200 KB GEMM CTA can join an SM one of these blocks occupies (see common.cuh)
template <typename K>
void prefer_shared(K kernel) {
  static bool done = false;
  if (!done) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    done = true;
  }
}
}  // namespace
}  // namespace lmsfc

using namespace lmsfc;
#define ST(s) ((cudaStream_t)(s))

// Load every exchange kernel now: with lazy module loading the first launch of a kernel can synchronise with kernels
// already running on the device -- fatal when one of those is spinning on a flag that only a later launch can raise.
extern "C" int lmsfc_xchg_init(void) {
  cudaFuncAttributes a;
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_reduce_kernel<false>));
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_reduce_kernel<true>));
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_ready_kernel));
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_norm_publish_kernel));
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_norm_wait_kernel));
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_wait_kernel));
  LMSFC_CUDA(cudaFuncGetAttributes(&a, xchg_adam_push_kernel));
  prefer_shared(xchg_reduce_kernel<false>);
  prefer_shared(xchg_reduce_kernel<true>);
  prefer_shared(xchg_adam_push_kernel);
  return ensure_reduce_scratch();
}

extern "C" int lmsfc_xchg_reduce(const lmsfc_xchg_peers* peers, int slot, int64_t lo, int64_t hi, float* norm_part,
                                 int max_ctas, void* stream) {
  XP P;
  if (int rc = to_xp(peers, P)) return rc;
  LMSFC_CHECK_ARG(slot >= 0 && slot < LMSFC_XCHG_SLOTS && lo % 4 == 0 && hi % 4 == 0 && hi >= lo && norm_part,
                  "xchg_reduce: slot %d range [%lld, %lld)", slot, (long long)lo, (long long)hi);
  if (int rc = ensure_reduce_scratch()) return rc;
  const long long n4 = (hi - lo) / 4;
  long long ctas = (n4 + 4 * TPB - 1) / (4 * TPB);
  if (ctas < 1) ctas = 1;
  if (max_ctas <= 0) max_ctas = 2 * 148;         // two blocks per SM keep ~1.2 MB of 16-byte loads in flight
  if (ctas > max_ctas) ctas = max_ctas;
  LMSFC_LAUNCH(xchg_ready_kernel, dim3(1), dim3(32), 0, ST(stream), P, slot);
  if (P.mc_grads) {
    LMSFC_LAUNCH(xchg_reduce_kernel<true>, dim3((unsigned)ctas), dim3(TPB), 0, ST(stream), P, slot, (long long)(lo / 4), n4,
                 1.0f / (float)P.world, red_scratch(ST(stream)), red_count(ST(stream)) + 62, norm_part);
  } else {
    LMSFC_LAUNCH(xchg_reduce_kernel<false>, dim3((unsigned)ctas), dim3(TPB), 0, ST(stream), P, slot, (long long)(lo / 4), n4,
                 1.0f / (float)P.world, red_scratch(ST(stream)), red_count(ST(stream)) + 62, norm_part);
  }
  return LMSFC_OK;
}
extern "C" int lmsfc_xchg_norm_publish(const lmsfc_xchg_peers* peers, const float* norm_parts, int nparts, int slot,
                                       int* set_flag, void* stream) {
  XP P;
  if (int rc = to_xp(peers, P)) return rc;
  LMSFC_CHECK_ARG(norm_parts && nparts >= 1 && slot >= 0 && slot < LMSFC_XCHG_SLOTS, "xchg_norm_publish: bad argument");
  LMSFC_LAUNCH(xchg_norm_publish_kernel, dim3(1), dim3(32), 0, ST(stream), P, norm_parts, nparts, slot, set_flag);
  return LMSFC_OK;
}
extern "C" int lmsfc_xchg_norm_wait(const lmsfc_xchg_peers* peers, int slot, float* sumsq_out, void* stream) {
  XP P;
  if (int rc = to_xp(peers, P)) return rc;
  LMSFC_CHECK_ARG(sumsq_out && slot >= 0 && slot < LMSFC_XCHG_SLOTS, "xchg_norm_wait: bad argument");
  LMSFC_LAUNCH(xchg_norm_wait_kernel, dim3(1), dim3(32), 0, ST(stream), P, slot, sumsq_out);
  return LMSFC_OK;
}
extern "C" int lmsfc_xchg_wait(const lmsfc_xchg_peers* peers, int slot, void* stream) {
  XP P;
  if (int rc = to_xp(peers, P)) return rc;
  LMSFC_CHECK_ARG(slot >= 0 && slot < LMSFC_XCHG_SLOTS, "xchg_wait: slot %d", slot);
  LMSFC_LAUNCH(xchg_wait_kernel, dim3(1), dim3(32), 0, ST(stream), P, slot);
  return LMSFC_OK;
}
extern "C" int lmsfc_xchg_adam_push(const lmsfc_xchg_peers* peers, int slot, int64_t unit_lo, int64_t unit_hi,
                                    int64_t own_lo, int64_t own_hi, int64_t fp32_end, float* m, float* v, float beta1,
                                    float beta2, float eps, const float* scratch, int max_ctas, void* stream) {
  XP P;
  if (int rc = to_xp(peers, P)) return rc;
  LMSFC_CHECK_ARG(slot >= 0 && slot < LMSFC_XCHG_SLOTS && m && v && scratch && unit_lo <= own_lo && own_lo <= own_hi &&
                      own_hi <= unit_hi && ((unit_lo | unit_hi | own_lo | own_hi | fp32_end) % 4) == 0,
                  "xchg_adam_push: slot %d unit [%lld, %lld) own [%lld, %lld)", slot, (long long)unit_lo, (long long)unit_hi,
                  (long long)own_lo, (long long)own_hi);
  if (int rc = ensure_reduce_scratch()) return rc;
  PushArgs a{unit_lo / 4, unit_hi / 4, own_lo / 4, own_hi / 4, fp32_end / 4, m, v, beta1, beta2, eps, scratch,
             red_count(ST(stream)) + 61, slot};
  const long long work4 = (unit_hi - unit_lo) / 4;
  long long ctas = (work4 + 8 * TPB - 1) / (8 * TPB);
  if (ctas < 1) ctas = 1;
  if (max_ctas <= 0) max_ctas = 4 * 148;
  if (ctas > max_ctas) ctas = max_ctas;
  prefer_shared(xchg_adam_push_kernel);
  LMSFC_LAUNCH(xchg_adam_push_kernel, dim3((unsigned)ctas), dim3(TPB), 0, ST(stream), P, a);
  return LMSFC_OK;
}
