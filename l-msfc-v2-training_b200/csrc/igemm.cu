// tcgen05 / TMEM implicit-GEMM kernels for sm_100a.
//
// Replaces, for the L-MSFC path, what the reference gets from cuDNN through
//   compressai/models/utils.py:128-146   conv / deconv (k5 s2, k3 s1)
//   compressai/layers/gdn.py:77-92       the 1x1 channel contraction of GDN / IGDN
// and their autograd backward passes.
//
// One CTA = one 128 x BLOCK_N output tile of one sub-pixel phase and one K slice.
// Warp roles (192 threads):
//   warps 0-3  gather the activation operand with cp.async (zero-filled padding, strided /
//              sub-pixel-phase taps, two-source channel concat) into SWIZZLE_128B shared
//              memory, then run the epilogue out of TMEM (tcgen05.ld)
//   warp 4     TMA producer for the weight operand (cp.async.bulk.tensor, SWIZZLE_128B)
//   warp 5     TMEM allocator + single-thread tcgen05.mma issuer
// Pipelines: smem full/empty mbarrier ring (STAGES deep) and one TMEM-full barrier.
// Small grids (latent-resolution layers) split K over a thread-block cluster (cluster dims 1x1xS,
// S <= 8): every CTA keeps its partial tile in TMEM, then scatters it column-slice-wise into the
// owning CTA's shared memory (st.shared::cluster), and after one cluster barrier each CTA sums the S
// partials of its own BLOCK_N/S columns and runs the epilogue for them -- no global atomics, no
// workspace, no counters (the atomic version queued ~10 us of same-address reductions in L2).
#include <stdlib.h>

#include "igemm.cuh"

namespace lmsfc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;            // 64 bf16 = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
constexpr int NUM_GATHER_THREADS = 128;
constexpr int NUM_THREADS = 256;   // warps 0-3 gather, 4 weight TMA, 5 MMA issue; all 8 run the direct epilogue

template <int BLOCK_N>
struct FwdCfg {
  static constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (BLOCK_N == 256) ? 4 : 6;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
  // fused GDN: one 32 rows x (64 + 16) B staging block per warp behind the barriers (x leaves through it in phase 1)
  static constexpr int XS_ROW = 80, XS_WARP = 32 * XS_ROW;
  static constexpr int SMEM_BYTES_FUSED = SMEM_BYTES + 8 * XS_WARP;
};

__device__ __forceinline__ int fast_div(int n, uint32_t mul, uint32_t shr) {   // n / d, d's magic from fast_div_magic
  return mul ? (int)(__umulhi((uint32_t)n, mul) >> shr) : n;
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
template <int W>
__device__ __forceinline__ void load_bf16(const __nv_bfloat16* p, uint32_t (&r)[W / 2]) {
  const uint4* s = reinterpret_cast<const uint4*>(p);
#pragma unroll
  for (int j = 0; j < W / 8; ++j) {
    const uint4 q = __ldg(s + j);
    r[4 * j] = q.x; r[4 * j + 1] = q.y; r[4 * j + 2] = q.z; r[4 * j + 3] = q.w;
  }
}
template <int W>
__device__ __forceinline__ void load_f32(const float* p, float (&r)[W]) {   // 16-byte aligned, warp-uniform address
  const float4* s = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int j = 0; j < W / 4; ++j) {
    const float4 q = __ldg(s + j);
    r[4 * j] = q.x; r[4 * j + 1] = q.y; r[4 * j + 2] = q.z; r[4 * j + 3] = q.w;
  }
}
template <int W>
__device__ __forceinline__ void store_bf16(__nv_bfloat16* p, const float (&v)[W]) {
  uint4* d = reinterpret_cast<uint4*>(p);
#pragma unroll
  for (int j = 0; j < W / 8; ++j)
    d[j] = make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                      pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7]));
}
template <int H>
__device__ __forceinline__ float pick(const uint32_t (&r)[H], int j) {
  return (j & 1) ? bf16_hi(r[j >> 1]) : bf16_lo(r[j >> 1]);
}

// GDN forward / backward maths on W channels of one pixel; xr / gr / rr = x, grad_out, rsqrt(norm) as packed bf16
template <int W>
__device__ __forceinline__ void gdn_math(const IgemmParams& p, int epi, const float* __restrict__ beta_ptr, float (&v)[W],
                                         const uint32_t (&xr)[W / 2], const uint32_t (&gr)[W / 2],
                                         const uint32_t (&rr)[W / 2], float (&rn)[W], int nb) {
  // norm >= beta_min > 0 and rn is a bf16 value: the approximate rsqrt / reciprocal (2 ulp) lose nothing here,
  // while the IEEE division of the inverse backward was a call per element
  if (epi == EPI_GDN) {
    float beta[W];
    load_f32<W>(beta_ptr + nb, beta);
#pragma unroll
    for (int j = 0; j < W; ++j) {
      const float nrm = v[j] + beta[j];
      asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rn[j]) : "f"(nrm));
      v[j] = nrm;
    }
    if (p.inverse) {
#pragma unroll
      for (int j = 0; j < W; ++j) v[j] = pick(xr, j) * (v[j] * rn[j]);
    } else {
#pragma unroll
      for (int j = 0; j < W; ++j) v[j] = pick(xr, j) * rn[j];
    }
  } else if (p.inverse) {  // EPI_GDN_BWD of y = x * sqrt(norm)
#pragma unroll
    for (int j = 0; j < W; ++j) {
      float f;
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(f) : "f"(pick(rr, j)));
      v[j] = pick(gr, j) * f + 2.f * pick(xr, j) * v[j];
    }
  } else {                 // EPI_GDN_BWD of y = x * rsqrt(norm)
#pragma unroll
    for (int j = 0; j < W; ++j) v[j] = pick(gr, j) * pick(rr, j) + 2.f * pick(xr, j) * v[j];
  }
}
// 32 bf16 of this thread's staged row (shared memory, 16-byte aligned)
__device__ __forceinline__ void load_bf16_smem32(uint32_t saddr, uint32_t (&r)[16]) {
#pragma unroll
  for (int j = 0; j < 4; ++j)
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[4 * j]), "=r"(r[4 * j + 1]), "=r"(r[4 * j + 2]), "=r"(r[4 * j + 3])
                 : "r"(saddr + 16 * j));
}

// 32 values of this thread's row -> its staging row in shared memory, as bf16 (64 B) or fp32 (128 B)
__device__ __forceinline__ void stage_bf16_32(uint32_t saddr, const float (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 4; ++j)
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr + 16 * j), "r"(pack_bf16(v[8 * j], v[8 * j + 1])),
                 "r"(pack_bf16(v[8 * j + 2], v[8 * j + 3])), "r"(pack_bf16(v[8 * j + 4], v[8 * j + 5])),
                 "r"(pack_bf16(v[8 * j + 6], v[8 * j + 7]))
                 : "memory");
}
__device__ __forceinline__ void stage_f32_32(uint32_t saddr, const float (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(saddr + 16 * j), "f"(v[4 * j]), "f"(v[4 * j + 1]),
                 "f"(v[4 * j + 2]), "f"(v[4 * j + 3])
                 : "memory");
}
// epilogue maths for W consecutive output channels [nb, nb+W) of one output pixel at offset o
template <int W>
__device__ __forceinline__ void epilogue_chunk(const IgemmParams& p, float (&v)[W], size_t o, int nb) {
  if (p.epi == EPI_BIAS) {
    if (p.bias) {
      float bv[W];
      load_f32<W>(p.bias + nb, bv);
#pragma unroll
      for (int j = 0; j < W; ++j) v[j] += bv[j];
    }
    if (p.relu == 1) {
#pragma unroll
      for (int j = 0; j < W; ++j) v[j] = fmaxf(v[j], 0.f);
    } else if (p.relu == 2) {            // LeakyReLU(0.01): the mean-scale hyperprior's activation (models/google.py:377-391)
#pragma unroll
      for (int j = 0; j < W; ++j) v[j] = v[j] > 0.f ? v[j] : kLeakySlope * v[j];
    }
    if (p.out_f32) {
      float4* d = reinterpret_cast<float4*>(p.out_f32 + o);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) d[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    if (p.out_abs) {
      float a[W];
#pragma unroll
      for (int j = 0; j < W; ++j) a[j] = fabsf(v[j]);
      store_bf16<W>(p.out_abs + o, a);
    }
  } else {
    uint32_t xr[W / 2], gr[W / 2], rr[W / 2];
    load_bf16<W>(p.aux_x + o, xr);
    if (p.epi == EPI_GDN_BWD) {
      load_bf16<W>(p.aux_gy + o, gr);
      load_bf16<W>(p.aux_rn + o, rr);
    }
    float rn[W];
    gdn_math<W>(p, p.epi, p.bias, v, xr, gr, rr, rn, nb);
    if (p.epi == EPI_GDN && p.out_rn) store_bf16<W>(p.out_rn + o, rn);
  }
  if (p.out_bf16) store_bf16<W>(p.out_bf16 + o, v);
}

// ---- thread-block cluster helpers (split-K reduction through distributed shared memory)
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

#ifdef LMSFC_TIMELINE
// debug build only (make TIMELINE=1): %globaltimer stamps of CTA 0 / last CTA, read back by profiles/timeline.py
__device__ unsigned long long g_timeline[64 * 16];
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TL(slot)                                                                                   \
  do {                                                                                             \
    const int _c = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                 \
    if (_c < 64) g_timeline[_c * 16 + (slot)] = gtimer();                                          \
    if (_c < 64 && ((slot) == 9 || (slot) == 10)) g_timeline[_c * 16 + (slot) + 2] = (unsigned long long)clock64();  \
  } while (0)
#else
#define TL(slot) do {} while (0)
#endif

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <int BLOCK_N, bool B_MN, bool FUSE>
__global__ void __launch_bounds__(NUM_THREADS, 1) igemm_fwd_kernel(const __grid_constant__ IgemmParams p) {
  static_assert(!FUSE || BLOCK_N == 256, "the fused GDN needs all 256 channels of a pixel in one CTA");
  using Cfg = FwdCfg<BLOCK_N>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int TMEM_COLS = FUSE ? 512 : BLOCK_N;      // fused: accumulator 0 (conv / dgrad) and 1 (GDN contraction)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint64_t* tmem_full2_bar = tmem_full_bar + 1;        // FUSE: the GDN contraction has retired
  uint64_t* stage_free_bar = tmem_full_bar + 2;        // FUSE (backward): phase-1 staging inside the B slots is released
  uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 3);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) TL(0);
  const int phase_id = blockIdx.z / p.ksplit, ks = blockIdx.z % p.ksplit;   // ks == %cluster_ctarank (cluster 1x1xksplit)
  const IgemmPhase ph = p.ph[phase_id];
  const int m0 = blockIdx.x * BLOCK_M;
  if (m0 >= ph.M) return;                       // phases differ in size: whole CTA exits together
  const int n0 = blockIdx.y * BLOCK_N;
  const int ctot = p.c0 + p.c1;
  const int kc_per_tap = ctot / BLOCK_K;
  const int nkb_total = ph.ntaps * kc_per_tap;
  const int kb_per = (nkb_total + p.ksplit - 1) / p.ksplit;
  const int kb_begin = ks * kb_per;
  const int nkb = min(kb_per, nkb_total - kb_begin);   // host guarantees >= 1
  const uint32_t smem_base = smem_u32(smem);

  // Row decode (no global memory): before the barrier / pdl_sync, so it overlaps the previous kernel's tail.
  // output row of this thread (warp w and w + 4 share a row quarter in the epilogue)
  const int row = (warp & 3) * 32 + lane;
  bool ok = false;
  size_t obase = 0;
  {
    const int m = m0 + row;
    if (m < ph.M) {
      const int t = fast_div(m, ph.wg_mul, ph.wg_shr), wg = m - t * ph.Wg;
      const int b = fast_div(t, ph.hg_mul, ph.hg_shr), hg = t - b * ph.Hg;
      const int ho = hg * p.sout + ph.oy, wo = wg * p.sout + ph.ox;
      ok = ho < p.Hout && wo < p.Wout;
      obase = ((size_t)(b * p.Hout + ho) * p.Wout + wo) * p.N + n0;
    }
  }
  // gather rows of this thread (producer warps): r0 + 16*i
  const int chunk = tid & 7, r0 = (tid >> 3) & 15;
  int pixbase[8], hi0[8], wi0[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + r0 + 16 * i;
    if (m < ph.M) {
      const int t = fast_div(m, ph.wg_mul, ph.wg_shr), wg = m - t * ph.Wg;
      const int b = fast_div(t, ph.hg_mul, ph.hg_shr), hg = t - b * ph.Hg;
      pixbase[i] = b * p.Hin * p.Win;
      hi0[i] = hg * p.sin;
      wi0[i] = wg * p.sin;
    } else {
      pixbase[i] = 0; hi0[i] = -100000; wi0[i] = -100000;
    }
  }

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&p.tmapB);
    if (FUSE && p.fuse) tma_prefetch_desc(&p.tmapG);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], NUM_GATHER_THREADS + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    if (FUSE) {
      mbar_init(tmem_full2_bar, 1);
      mbar_init(stage_free_bar, NUM_THREADS);
    }
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();                            // everything above touched no global memory
#ifdef LMSFC_TIMELINE
  const int dbg = p.dbg;                 // bit0: skip the gather, bit1: skip the TMA loads (timing experiments)
#else
  constexpr int dbg = 0;
#endif
  if (tid == 0) TL(1);

  const uint32_t trow = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  const bool fused = FUSE && p.fuse != 0;
  int ring_stage = 0, ring_phase = 0;    // FUSE: where the TMA / MMA thread's ring stands after the main loop

  if (warp < 4) {
    // ================= activation gather (producer) =================
    const uint32_t dst_thread = smem_base + r0 * 128 + ((chunk ^ (r0 & 7)) << 4);
    int stage = 0, phase = 0;
    int tap = ph.tap_begin + kb_begin / kc_per_tap, kc = kb_begin % kc_per_tap;
    if (p.a_mode != 0) {
      // ---- computed operand (GDN forward): load x -> x^2 in registers -> swizzled st.shared (+ optional global copy)
      const bool writer = p.a_out != nullptr && blockIdx.y == 0;
      for (int kb = 0; kb < nkb; ++kb) {
        const int coff = kc * BLOCK_K + chunk * 8;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        const uint32_t dst = dst_thread + stage * Cfg::STAGE_BYTES;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint4 xv[4];
          size_t off[4];
          bool okr[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int hi = hi0[4 * h + i], wi = wi0[4 * h + i];
            okr[i] = (unsigned)hi < (unsigned)p.Hin && (unsigned)wi < (unsigned)p.Win;
            off[i] = (size_t)(pixbase[4 * h + i] + hi * p.Win + wi) * p.c0 + coff;
            xv[i] = okr[i] ? __ldg(reinterpret_cast<const uint4*>(p.src0 + off[i])) : make_uint4(0u, 0u, 0u, 0u);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t xx[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float x0 = bf16_lo(xx[j]), x1 = bf16_hi(xx[j]);
              o[j] = pack_bf16(x0 * x0, x1 * x1);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (4 * h + i) * 2048), "r"(o[0]), "r"(o[1]),
                         "r"(o[2]), "r"(o[3])
                         : "memory");
            if (writer && okr[i]) *reinterpret_cast<uint4*>(p.a_out + off[i]) = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
        fence_proxy_async_smem();          // generic-proxy stores -> visible to the tensor core's async proxy
        mbar_arrive(&full_bar[stage]);
        if (++kc == kc_per_tap) { kc = 0; ++tap; }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
    for (int kb = (p.a_mode != 0 ? nkb : 0); kb < nkb; ++kb) {
      const int cbase = kc * BLOCK_K;
      const bool first = cbase < p.c0;
      const __nv_bfloat16* src = first ? p.src0 : p.src1;
      const int C = first ? p.c0 : p.c1;
      const int coff = (first ? cbase : cbase - p.c0) + chunk * 8;
      const int dy = p.dy[tap], dx = p.dx[tap];
      mbar_wait(&empty_bar[stage], phase ^ 1);
      const uint32_t dst = dst_thread + stage * Cfg::STAGE_BYTES;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int hi = hi0[i] + dy, wi = wi0[i] + dx;
        const bool ok = (unsigned)hi < (unsigned)p.Hin && (unsigned)wi < (unsigned)p.Win;
        const __nv_bfloat16* g = ok ? src + ((size_t)(pixbase[i] + hi * p.Win + wi) * C + coff) : src;
        if (!(dbg & 1)) cp_async16(dst + i * 2048, g, ok ? 16u : 0u);
      }
      cp_async_mbar_arrive_noinc(&full_bar[stage]);
      if (++kc == kc_per_tap) { kc = 0; ++tap; }
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    if (tid == 0) TL(2);
    cp_async_wait_all();
    if (tid == 0) TL(3);

    // ================= epilogue: TMEM -> registers -> global =================
    if (nkb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    if (tid == 0) TL(4);
  } else if (warp == 4) {
    // ================= weight TMA producer =================
    if (lane == 0) {
      int stage = 0, phase = 0;
      int tap = ph.tap_begin + kb_begin / kc_per_tap, kc = kb_begin % kc_per_tap;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], (dbg & 2) ? 0 : Cfg::B_STAGE_BYTES);
        const uint32_t sb = smem_base + stage * Cfg::STAGE_BYTES + A_STAGE_BYTES;
        if (dbg & 2) {
        } else if (!B_MN) {
          const int kcoord = (int)p.widx[tap] * ctot + kc * BLOCK_K;
          tma_load_2d(&p.tmapB, &full_bar[stage], sb, kcoord, n0);
        } else {
          // rows = k channel, cols = tap*N + n: BLOCK_N/64 boxes of [64 k][64 n], one 8 KB segment each
          const int ncoord = (int)p.widx[tap] * p.w_ld_tap + n0;
#pragma unroll
          for (int seg = 0; seg < BLOCK_N / 64; ++seg)
            tma_load_2d(&p.tmapB, &full_bar[stage], sb + seg * (BLOCK_K * 128), ncoord + 64 * seg, kc * BLOCK_K);
        }
        if (++kc == kc_per_tap) { kc = 0; ++tap; }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (fused) {
        // gamma_eff follows the weights through the same ring: 4 K blocks of 64 channels into the B slots.  Backward
        // (fuse 2) defers blocks 2 and 3: their B slots stage the aux tiles of phase 1 first.
        const int early = p.fuse == 1 ? 4 : 2;
        for (int i = 0; i < early; ++i) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::B_STAGE_BYTES);
          const uint32_t sb = smem_base + stage * Cfg::STAGE_BYTES + A_STAGE_BYTES;
          if (p.fuse == 1) {
            tma_load_2d(&p.tmapG, &full_bar[stage], sb, i * BLOCK_K, 0);
          } else {
#pragma unroll
            for (int seg = 0; seg < BLOCK_N / 64; ++seg)
              tma_load_2d(&p.tmapG, &full_bar[stage], sb + seg * (BLOCK_K * 128), 64 * seg, i * BLOCK_K);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        ring_stage = stage; ring_phase = phase;
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer (one thread) =================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BLOCK_N, 0, B_MN ? 1 : 0);
      int stage = 0, phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (kb == 0) TL(9);
        const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
        const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < BLOCK_K / 16; ++k) {
          const uint64_t ad = make_sdesc_sw128(sa + k * 32, 16, 1024);
          // K-major: 32 B per 16-k step inside the 128 B row; MN-major: 16 k rows = 2 KB per step,
          // LBO = next 64-n segment, SBO = next 8 k rows (same image as wgrad.cu's gathered operand)
          const uint64_t bd = B_MN ? make_sdesc_sw128(sb + k * 2048, BLOCK_K * 128, 1024)
                                   : make_sdesc_sw128(sb + k * 32, 16, 1024);
          umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (nkb > 0) umma_commit(tmem_full_bar);
      ring_stage = stage; ring_phase = phase;
      TL(10);
    }
  }
  if (p.ksplit == 1) {
    // ================= direct epilogue: all 8 warps =================
    // Warp w reads TMEM lanes 32*(w%4)...: warps w and w+4 share a row quarter and split the columns in halves
    // (two warps per scheduler instead of one: the epilogue is instruction / latency bound).  Thread-per-row
    // GLOBAL accesses (16-byte pieces at a 512-byte stride) are avoided altogether: all tile traffic goes through
    // the idle pipeline memory in 64-column passes -- a warp copies the aux rows of ITS 32 pixels in with coalesced
    // cp.async (all in flight at once), every thread computes its own row from / to shared memory (row stride
    // 128 + 16 B resp. 256 + 16 B: conflict-free for 16-byte accesses), the warp copies the results out as whole
    // row segments.  Only __syncwarp is needed: a warp touches its own rows of its own column half.
    __syncwarp();
    if (nkb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    if (tid == 0) TL(4);
    constexpr int PW = 64, ROWB = PW * 2 + 16, ROWF = PW * 4 + 16, HALF = BLOCK_N / 2, PASSES = HALF / PW;
    constexpr int SLOT_B = BLOCK_M * ROWB, SLOT_F = BLOCK_M * ROWF, GROUP = 3 * SLOT_B + SLOT_F;
    static_assert(2 * GROUP <= Cfg::STAGES * Cfg::STAGE_BYTES, "staging slots");
    const int half = warp >> 2, wrow0 = (warp & 3) * 32;
    const long long orow = ok ? (long long)obase : -1ll;      // element offset of this row (column n0), -1: not stored

    if constexpr (FUSE) {
      if (fused) {
        // ================= fused GDN, phase 1: accumulator 0 -> second operand (x^2 | t) in the A slots =================
        // Thread = one pixel row x the 128 channels of its warp's column half, in four 32-channel chunks.  Channels
        // [64 i, 64 i + 64) are K block i of the contraction: its A tile is written by the four warps of column half
        // i / 2 (128 threads = the gather threads' arrival count of full_bar), its B tile (gamma) arrives by TMA.
        constexpr int XS_ROW = Cfg::XS_ROW, XS_WARP = Cfg::XS_WARP;
        const bool fwd = p.fuse == 1;
        // phase-1 staging: forward -- x on its way out, one 32 x 80 B block per warp behind the barriers;
        // backward -- the aux tiles x, rn, addend (3 blocks per warp) inside the B slots of K blocks 2 / 3
        const uint32_t xs_warp = fwd ? smem_base + STAGES * Cfg::STAGE_BYTES + 256 + (uint32_t)warp * XS_WARP
                                     : smem_base + (uint32_t)((nkb + 2 + half) % STAGES) * Cfg::STAGE_BYTES + A_STAGE_BYTES +
                                           (uint32_t)(warp & 3) * (3 * XS_WARP);
        const uint32_t xs_row = xs_warp + (uint32_t)lane * XS_ROW;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          const int col = half * 128 + c * 32;
          if (!fwd) {                    // aux rows of the warp's 32 pixels, 64 B each: 4 lanes per row, coalesced
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int rr = 8 * i + (lane >> 2);
              const long long o = __shfl_sync(0xffffffffu, orow, rr);
              if (o >= 0) {
                const uint32_t d = xs_warp + (uint32_t)rr * XS_ROW + (lane & 3) * 16;
                const long long g = o + col + (lane & 3) * 8;
                cp_async16(d, p.aux_x + g, 16u);
                cp_async16(d + XS_WARP, p.aux_rn + g, 16u);
                if (p.aux_add) cp_async16(d + 2 * XS_WARP, p.aux_add + g, 16u);
              }
            }
          }
          uint32_t r[32];
          tmem_ld32(trow + col, r);
          tmem_ld_wait();
          uint32_t a2[16];                // the second GEMM's operand for this row: x^2 (forward) | t (backward), bf16
          if (fwd) {
            float bv[32];
            if (p.bias) {
              load_f32<32>(p.bias + col, bv);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) bv[j] = 0.f;
            }
            uint32_t xb[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              xb[j] = pack_bf16(__uint_as_float(r[2 * j]) + bv[2 * j], __uint_as_float(r[2 * j + 1]) + bv[2 * j + 1]);
              const float x0 = bf16_lo(xb[j]), x1 = bf16_hi(xb[j]);      // x^2 of the bf16 value every later reader sees
              a2[j] = pack_bf16(x0 * x0, x1 * x1);
            }
            if (p.out_x) {               // x leaves as 64-byte row pieces through the warp's staging block
#pragma unroll
              for (int j = 0; j < 4; ++j)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(xs_row + 16 * j), "r"(xb[4 * j]),
                             "r"(xb[4 * j + 1]), "r"(xb[4 * j + 2]), "r"(xb[4 * j + 3])
                             : "memory");
              __syncwarp();
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int rr = 8 * i + (lane >> 2);
                const long long o = __shfl_sync(0xffffffffu, orow, rr);
                if (o >= 0) {
                  uint32_t q0, q1, q2, q3;
                  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                               : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3)
                               : "r"(xs_warp + (uint32_t)rr * XS_ROW + (lane & 3) * 16));
                  *reinterpret_cast<uint4*>(p.out_x + o + col + (lane & 3) * 8) = make_uint4(q0, q1, q2, q3);
                }
              }
              __syncwarp();
            }
          } else {
            cp_async_wait_all();
            __syncwarp();
            if (ok) {
              uint32_t xr[16], rr16[16], ar[16];
              load_bf16_smem32(xs_row, xr);
              load_bf16_smem32(xs_row + XS_WARP, rr16);
              if (p.aux_add) {
                load_bf16_smem32(xs_row + 2 * XS_WARP, ar);
              } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) ar[j] = 0u;
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                // gy = data gradient (+ the other consumer's), rounded to the bf16 value phase 2 recomputes
                const uint32_t gb = pack_bf16(__uint_as_float(r[2 * j]) + bf16_lo(ar[j]), __uint_as_float(r[2 * j + 1]) + bf16_hi(ar[j]));
                const float r0 = bf16_lo(rr16[j]), r1 = bf16_hi(rr16[j]);
                const float f0 = p.inverse ? 0.5f * r0 : -0.5f * r0 * r0 * r0;
                const float f1 = p.inverse ? 0.5f * r1 : -0.5f * r1 * r1 * r1;
                a2[j] = pack_bf16(bf16_lo(gb) * bf16_lo(xr[j]) * f0, bf16_hi(gb) * bf16_hi(xr[j]) * f1);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) a2[j] = 0u;
            }
            __syncwarp();                // the staging blocks are refilled by the next chunk's loads
          }
          // this row's 64 bytes of K block kbg, SWIZZLE_128B K-major: 16-byte chunk q of row r sits at r*128 + ((q ^ (r & 7)) << 4)
          const int kbg = col >> 6;
          const uint32_t sA = smem_base + (uint32_t)((nkb + kbg) % STAGES) * Cfg::STAGE_BYTES;
          {
            const uint32_t arow = sA + (uint32_t)row * 128;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int q = (c & 1) * 4 + j;
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(arow + (uint32_t)((q ^ (row & 7)) << 4)),
                           "r"(a2[4 * j]), "r"(a2[4 * j + 1]), "r"(a2[4 * j + 2]), "r"(a2[4 * j + 3])
                           : "memory");
            }
          }
          if (c & 1) {
            fence_proxy_async_smem();    // generic-proxy stores -> visible to the tensor core's async proxy
            mbar_arrive(&full_bar[(nkb + kbg) % STAGES]);
            if (p.out_t) {               // x^2 / t also leave for HBM (gamma's gradient reads them): straight from the A slot
              __syncwarp();              // the warp's own 32 rows of this K block are complete
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int rr = 4 * i + (lane >> 3), q = lane & 7, R = wrow0 + rr;
                const long long o = __shfl_sync(0xffffffffu, orow, rr);
                if (o >= 0) {
                  uint32_t q0, q1, q2, q3;
                  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                               : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3)
                               : "r"(sA + (uint32_t)R * 128 + (uint32_t)((q ^ (R & 7)) << 4)));
                  *reinterpret_cast<uint4*>(p.out_t + o + kbg * 64 + q * 8) = make_uint4(q0, q1, q2, q3);
                }
              }
            }
          }
        }
        if (!fwd) {
          fence_proxy_async_smem();      // staging reads (generic proxy) before the TMA writes of gamma blocks 2, 3
          mbar_arrive(stage_free_bar);
          if (warp == 4 && lane == 0) {
            mbar_wait(stage_free_bar, 0);
            int stage = ring_stage, phase = ring_phase;
            for (int i = 2; i < 4; ++i) {
              mbar_wait(&empty_bar[stage], phase ^ 1);
              mbar_arrive_expect_tx(&full_bar[stage], Cfg::B_STAGE_BYTES);
              const uint32_t sb = smem_base + stage * Cfg::STAGE_BYTES + A_STAGE_BYTES;
#pragma unroll
              for (int seg = 0; seg < BLOCK_N / 64; ++seg)
                tma_load_2d(&p.tmapG, &full_bar[stage], sb + seg * (BLOCK_K * 128), 64 * seg, i * BLOCK_K);
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
          }
        }
        if (warp == 5 && lane == 0) {
          // ---- the contraction: 4 K blocks x 4 MMAs into TMEM columns [256, 512)
          const uint32_t idesc2 = make_idesc_bf16(BLOCK_M, BLOCK_N, 0, fwd ? 0 : 1);
          int stage = ring_stage, phase = ring_phase;
          for (int i = 0; i < 4; ++i) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
            const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) {
              const uint64_t ad = make_sdesc_sw128(sa + k * 32, 16, 1024);
              const uint64_t bd = fwd ? make_sdesc_sw128(sb + k * 32, 16, 1024)
                                      : make_sdesc_sw128(sb + k * 2048, BLOCK_K * 128, 1024);
              umma_bf16(tmem_base + BLOCK_N, ad, bd, idesc2, (i | k) != 0);
            }
            umma_commit(&empty_bar[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          umma_commit(tmem_full2_bar);
        }
        __syncwarp();
        mbar_wait(tmem_full2_bar, 0);
        tc_fence_after();
        __syncthreads();                 // every warp's x^2 / t copy-out has left the A slots: phase 2 stages over them
      }
    }
    // fused phase 2 == the GDN epilogues below with the GEMM result in accumulator 1 and x (forward) resp. grad_out
    // (backward) recomputed from accumulator 0 instead of loaded
    const bool gdn_f = fused ? p.fuse == 1 : p.epi == EPI_GDN, gdn_b = fused ? p.fuse == 2 : p.epi == EPI_GDN_BWD;
    const uint32_t tacc = fused ? trow + BLOCK_N : trow;
    // per column half: bf16 slots 0..2 and one fp32 slot.  GDN forward {x -> y in place, rn}; GDN backward
    // {x -> grad_in in place, grad_out, rn}; conv {y, |y|, -, fp32}
    const uint32_t sgroup = smem_base + (uint32_t)half * GROUP;
    const uint32_t srow_b = sgroup + (uint32_t)row * ROWB, swarp_b = sgroup + (uint32_t)wrow0 * ROWB;
    const uint32_t srow_f = sgroup + 3 * SLOT_B + (uint32_t)row * ROWF, swarp_f = sgroup + 3 * SLOT_B + (uint32_t)wrow0 * ROWF;
    auto issue = [&](const __nv_bfloat16* src, int col, int slot) {   // 4 rows x 128 B per instruction
      const uint32_t sdst = swarp_b + (uint32_t)slot * SLOT_B + (lane & 7) * 16;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int rr = 4 * i + (lane >> 3);
        const long long o = __shfl_sync(0xffffffffu, orow, rr);
        if (o >= 0) cp_async16(sdst + (uint32_t)rr * ROWB, src + o + col + (lane & 7) * 8, 16u);
      }
    };
    auto copy_out = [&](uint32_t swarp, int stride, void* dst, int col, int esize) {
      const int lpr = (PW * esize) >> 4, sub = lane / lpr, cb16 = (lane % lpr) * 16;   // lanes per row: 8 (bf16) / 16 (fp32)
      for (int i = 0; i < 32; i += 32 / lpr) {
        const int rr = i + sub;
        const long long o = __shfl_sync(0xffffffffu, orow, rr);
        if (o >= 0) {
          uint32_t a, b, c, d;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(swarp + rr * stride + cb16));
          *reinterpret_cast<uint4*>((uint8_t*)dst + (size_t)(o + col) * esize + cb16) = make_uint4(a, b, c, d);
        }
      }
    };
#pragma unroll 1
    for (int ps = 0; ps < PASSES; ++ps) {
      const int col = half * HALF + ps * PW;                // first column of this pass inside the tile
      if ((gdn_f && !fused) || gdn_b) {
        issue(p.aux_x, col, 0);
        if (gdn_b) {
          if (!fused) issue(p.aux_gy, col, 1);
          else if (p.aux_add) issue(p.aux_add, col, 1);
          issue(p.aux_rn, col, 2);
        }
        cp_async_wait_all();
        __syncwarp();
      }
      if (tid == 0 && ps == 0) TL(13);
#pragma unroll 1
      for (int cb = 0; cb < PW; cb += 32) {
        uint32_t r[32];
        if (nkb > 0) {                   // warp-uniform: tcgen05.ld is .sync.aligned
          tmem_ld32(tacc + col + cb, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0u;
        }
        uint32_t r0[FUSE ? 32 : 1];      // fused: accumulator 0 (x resp. grad_out before rounding)
        if constexpr (FUSE) {
          if (fused) {
            tmem_ld32(trow + col + cb, r0);
            tmem_ld_wait();
          }
        }
        if (!ok) continue;               // rows past M / cropped outputs are never stored
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        const uint32_t coff = (uint32_t)cb * 2;
        if (gdn_f || gdn_b) {
          uint32_t xr[16], gr[16], rr[16];
          float rn[32];
          bool from_acc = false;
          if constexpr (FUSE) {
            if (fused && gdn_f) {        // x = bf16(acc0 + bias), exactly the value phase 1 stored and squared
              float bv[32];
              if (p.bias) {
                load_f32<32>(p.bias + col + cb, bv);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) bv[j] = 0.f;
              }
#pragma unroll
              for (int j = 0; j < 16; ++j)
                xr[j] = pack_bf16(__uint_as_float(r0[2 * j]) + bv[2 * j], __uint_as_float(r0[2 * j + 1]) + bv[2 * j + 1]);
              from_acc = true;
            }
          }
          if (!from_acc) load_bf16_smem32(srow_b + coff, xr);
          if (gdn_b) {
            load_bf16_smem32(srow_b + 2 * SLOT_B + coff, rr);
            bool gy_from_acc = false;
            if constexpr (FUSE) {
              if (fused) {               // grad_out = bf16(acc0 + addend), as in phase 1
                if (p.aux_add) {
                  load_bf16_smem32(srow_b + SLOT_B + coff, gr);
                } else {
#pragma unroll
                  for (int j = 0; j < 16; ++j) gr[j] = 0u;
                }
#pragma unroll
                for (int j = 0; j < 16; ++j)
                  gr[j] = pack_bf16(__uint_as_float(r0[2 * j]) + bf16_lo(gr[j]), __uint_as_float(r0[2 * j + 1]) + bf16_hi(gr[j]));
                gy_from_acc = true;
              }
            }
            if (!gy_from_acc) load_bf16_smem32(srow_b + SLOT_B + coff, gr);
          }
          gdn_math<32>(p, gdn_f ? EPI_GDN : EPI_GDN_BWD, fused ? p.beta : p.bias, v, xr, gr, rr, rn, n0 + col + cb);
          if (gdn_f && p.out_rn) stage_bf16_32(srow_b + SLOT_B + coff, rn);
          if (p.out_bf16) stage_bf16_32(srow_b + coff, v);
        } else {
          if (p.bias) {
            float bv[32];
            load_f32<32>(p.bias + n0 + col + cb, bv);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += bv[j];
          }
          if (p.relu == 1) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          } else if (p.relu == 2) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : kLeakySlope * v[j];
          }
          if (p.out_bf16) stage_bf16_32(srow_b + coff, v);
          if (p.out_f32) stage_f32_32(srow_f + coff * 2, v);
          if (p.out_abs) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fabsf(v[j]);
            stage_bf16_32(srow_b + SLOT_B + coff, v);
          }
        }
      }
      __syncwarp();
      if (tid == 0 && ps == 0) TL(14);
      if (p.out_bf16) copy_out(swarp_b, ROWB, p.out_bf16, col, 2);
      if (gdn_f && p.out_rn) copy_out(swarp_b + SLOT_B, ROWB, p.out_rn, col, 2);
      if (!gdn_f && !gdn_b) {
        if (p.out_f32) copy_out(swarp_f, ROWF, p.out_f32, col, 4);
        if (p.out_abs) copy_out(swarp_b + SLOT_B, ROWB, p.out_abs, col, 2);
      }
      if (tid == 0 && ps == 0) TL(15);
      __syncwarp();                      // staging rows are reused by the next pass
    }
  }
  if (p.ksplit > 1) {
    // ================= split K: reduce the cluster's partial tiles through DSMEM =================
    // recv[src rank][w/4][row] float4 overlays the (now idle) pipeline stages of every CTA; CTA r owns
    // output columns [r*w, (r+1)*w), w = BLOCK_N / ksplit.  Row-fastest, so the 32 lanes (= rows) of a
    // warp write / read 512 contiguous bytes per instruction.
    const int S = p.ksplit, w = BLOCK_N / S;
    __syncwarp();
    cluster_sync_all();                 // every CTA of the cluster has retired its MMAs: stage smem is free
    if (tid == 0) TL(5);
    if (warp < 4) {
#pragma unroll 1
      for (int cb = 0; cb < BLOCK_N; cb += 32) {
        uint32_t r[32];
        if (nkb > 0) {                   // warp-uniform: tcgen05.ld is .sync.aligned
          tmem_ld32(trow + cb, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0u;
        }
        if (!ok) continue;               // rows past M / cropped outputs are never stored
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int col = cb + 16 * h;
          const int owner = col / w, cin = col - owner * w;
          const uint32_t dst = map_to_cta(smem_base + (uint32_t)(((ks * (w >> 2) + (cin >> 2)) * BLOCK_M + row) * 16),
                                          (uint32_t)owner);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            st_cluster_v4(dst + j * (BLOCK_M * 16), __uint_as_float(r[16 * h + 4 * j]), __uint_as_float(r[16 * h + 4 * j + 1]),
                          __uint_as_float(r[16 * h + 4 * j + 2]), __uint_as_float(r[16 * h + 4 * j + 3]));
        }
      }
    }
    cluster_sync_all();                 // all partials have landed
    if (tid == 0) TL(6);
    if (warp < 4 && ok) {
      const float4* mine = reinterpret_cast<const float4*>(smem) + row;     // [src][w/4][row]
      const int w4 = w >> 2;
      if (w >= 32) {
#pragma unroll 1
        for (int c = 0; c < w; c += 32) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
          for (int src = 0; src < S; ++src) {
            const float4* q = mine + (size_t)(src * w4 + (c >> 2)) * BLOCK_M;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 t = q[j * BLOCK_M];
              v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
            }
          }
          epilogue_chunk<32>(p, v, obase + ks * w + c, n0 + ks * w + c);
        }
      } else {                          // w == 16
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0.f;
        for (int src = 0; src < S; ++src) {
          const float4* q = mine + (size_t)(src * w4) * BLOCK_M;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 t = q[j * BLOCK_M];
            v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
          }
        }
        epilogue_chunk<16>(p, v, obase + ks * w, n0 + ks * w);
      }
    }
  }
  if (tid == 0) TL(7);
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
  if (tid == 0) TL(8);
}

#ifdef LMSFC_TIMELINE
extern "C" int lmsfc_debug_read_timeline(unsigned long long* host64x16) {
  return (int)cudaMemcpyFromSymbol(host64x16, g_timeline, sizeof(unsigned long long) * 64 * 16);
}
#endif

// ---------------------------------------------------------------- host side
// how many clusters of `ks` CTAs of this kernel can be resident at once (1 CTA per SM; a cluster
// must fit inside one GPC), queried once per configuration
template <int BLOCK_N, bool B_MN>
static int max_active_clusters_t(int ks) {
  static int cache[9] = {0};
  if (cache[ks]) return cache[ks];
  using Cfg = FwdCfg<BLOCK_N>;
  cudaFuncSetAttribute(igemm_fwd_kernel<BLOCK_N, B_MN, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1, 1, ks * 64);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = (unsigned)ks;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, igemm_fwd_kernel<BLOCK_N, B_MN, false>, &cfg) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    n = 148 / ks / 2;                      // conservative guess
  }
  return cache[ks] = n;
}
static int max_active_clusters(int bn, bool mn, int ks) {
  if (bn == 256) return mn ? max_active_clusters_t<256, true>(ks) : max_active_clusters_t<256, false>(ks);
  return mn ? max_active_clusters_t<128, true>(ks) : max_active_clusters_t<128, false>(ks);
}

template <int BLOCK_N, bool B_MN, bool FUSE>
static int launch_fwd_t(const IgemmParams& p, dim3 grid, cudaStream_t stream) {
  using Cfg = FwdCfg<BLOCK_N>;
  constexpr int SMEM = FUSE ? Cfg::SMEM_BYTES_FUSED : Cfg::SMEM_BYTES;
  static bool attr_done = false;
  if (!attr_done) {
    LMSFC_CUDA(cudaFuncSetAttribute(igemm_fwd_kernel<BLOCK_N, B_MN, FUSE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_done = true;
  }
  // the K slices of one output tile form a cluster along z
  LMSFC_CUDA(launch_pdl(igemm_fwd_kernel<BLOCK_N, B_MN, FUSE>, grid, dim3(NUM_THREADS), (size_t)SMEM, stream,
                        (unsigned)p.ksplit, p));
  ++g_launch_count;
  return LMSFC_OK;
}

// Entropy-coded path: h_s(z_hat) indexes the Gaussian CDF tables through hard thresholds, so encoder and decoder must
// compute it with the SAME fp32 summation order whatever the batch size or the GPU's occupancy.  In this mode every
// output element is one CTA's sequential K loop (128-wide tiles, no split K): independent of M, B and the device.
static thread_local int g_batch_invariant = 0;
void set_batch_invariant(int on) { g_batch_invariant = on; }

int launch_igemm(IgemmParams& p, cudaStream_t stream) {
  LMSFC_CHECK_ARG(p.N > 0 && p.N % 128 == 0, "igemm: N=%d must be a multiple of 128", p.N);
  LMSFC_CHECK_ARG(p.c0 % 64 == 0 && p.c1 % 64 == 0 && p.c0 > 0, "igemm: channels (%d,%d) must be multiples of 64",
                  p.c0, p.c1);
  LMSFC_CHECK_ARG(p.nphase >= 1 && p.nphase <= kMaxPhases, "igemm: nphase=%d", p.nphase);
  LMSFC_CHECK_ARG(p.w_cols % 64 == 0 && (p.w_mn ? p.w_rows >= p.c0 + p.c1 : p.w_rows >= p.N),
                  "igemm: weight matrix %dx%d", p.w_rows, p.w_cols);
  int maxM = 0, min_taps = kMaxTaps, max_taps = 0;
  int64_t tiles128 = 0;
  for (int i = 0; i < p.nphase; ++i) {
    LMSFC_CHECK_ARG(p.ph[i].M > 0 && p.ph[i].ntaps >= 1 && p.ph[i].tap_begin + p.ph[i].ntaps <= kMaxTaps,
                    "igemm: phase %d M=%d ntaps=%d", i, p.ph[i].M, p.ph[i].ntaps);
    fast_div_magic(p.ph[i].Wg, &p.ph[i].wg_mul, &p.ph[i].wg_shr);
    fast_div_magic(p.ph[i].Hg, &p.ph[i].hg_mul, &p.ph[i].hg_shr);
    maxM = max(maxM, p.ph[i].M);
    min_taps = min(min_taps, p.ph[i].ntaps);
    max_taps = max(max_taps, p.ph[i].ntaps);
    tiles128 += ceil_div(p.ph[i].M, BLOCK_M);
  }
  // Tile width / split-K choice from a small cost model (us): waves * (K blocks * t_kb + fixed).
  // 256-wide tiles halve the gather traffic per flop; splitting K over a cluster fills more of the
  // 148 SMs when the output has few tiles (latent-resolution layers).  t_kb: tensor-pipe time of one
  // 64-deep K block (0.38 / 0.19 us), stretched to 0.5 / 0.3 us when a full-chip grid saturates L2.
  const int kc = (p.c0 + p.c1) / BLOCK_K;
  const int grid_x = ceil_div(maxM, BLOCK_M);
  int bn = 128, ksplit = 1;
  double best = 1e30;
  static const int ks_opts[] = {1, 2, 4, 8};
  for (int cand_bn = 128; cand_bn <= 256; cand_bn += 128) {
    if (p.N % cand_bn != 0 || (g_batch_invariant && cand_bn != 128)) continue;
    const int64_t tiles = tiles128 * (p.N / cand_bn);
    for (int ks : ks_opts) {
      if (ks > 1 && g_batch_invariant) continue;
      if (ks > 1) {
        if (cand_bn / ks < 16) continue;
        const int kb_per = ceil_div(min_taps * kc, ks);
        if ((ks - 1) * kb_per >= min_taps * kc) continue;      // no empty K slice in any phase
      }
      const int64_t ctas = tiles * ks;
      const double waves = ks == 1 ? (double)ceil_div(ctas, 148)
                                   : (double)ceil_div(tiles, max_active_clusters(cand_bn, p.w_mn != 0, ks));
      const bool busy = ctas >= 100;
      const double t_kb = cand_bn == 256 ? (busy ? 0.5 : 0.38) : (busy ? 0.3 : 0.19);
      const double t = waves * (ceil_div(max_taps * kc, ks) * t_kb + 4.0) + (ks > 1 ? 1.5 : 0.0);
      if (t < best) { best = t; bn = cand_bn; ksplit = ks; }
    }
  }
  if (p.dry_run) {                       // the caller only asks which tiling the cost model picks (fusion decision)
    p.ksplit = ksplit;
    p.dry_bn = bn;
    return LMSFC_OK;
  }
  if (p.fuse) {
    // the fused GDN needs a CTA to own all 256 channels of its pixels and the whole K loop
    LMSFC_CHECK_ARG(p.N == 256 && p.gamma && (p.fuse == 1 ? p.beta != nullptr : (p.aux_x && p.aux_rn && p.out_t)) && p.out_bf16,
                    "igemm: fused GDN needs N == 256 (got %d) and its operands", p.N);
    bn = 256;
    ksplit = 1;
    int rcg = p.fuse == 1 ? make_tmap_2d_bf16(&p.tmapG, p.gamma, 256, 256, 256, 256, 64)
                          : make_tmap_2d_bf16(&p.tmapG, p.gamma, 256, 256, 256, 64, 64);
    if (rcg) return rcg;
  }
  dim3 grid(grid_x, p.N / bn, p.nphase * ksplit);
  p.ksplit = ksplit;
#ifdef LMSFC_TIMELINE
  { const char* e = getenv("LMSFC_WG_DBG"); p.dbg = e ? atoi(e) : 0; }
#endif
  int rc = make_tmap_2d_bf16(&p.tmapB, p.w, (uint64_t)p.w_rows, (uint64_t)p.w_cols, (uint64_t)p.w_cols,
                             p.w_mn ? 64 : bn, 64);
  if (rc) return rc;
  if (p.fuse) return p.w_mn ? launch_fwd_t<256, true, true>(p, grid, stream) : launch_fwd_t<256, false, true>(p, grid, stream);
  if (p.w_mn) return bn == 256 ? launch_fwd_t<256, true, false>(p, grid, stream) : launch_fwd_t<128, true, false>(p, grid, stream);
  return bn == 256 ? launch_fwd_t<256, false, false>(p, grid, stream) : launch_fwd_t<128, false, false>(p, grid, stream);
}

}  // namespace lmsfc
