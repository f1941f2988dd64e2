// tcgen05 / TMEM implicit-GEMM kernels for sm_100a.
//
// Replaces, for the L-MSFC path, what the reference gets from cuDNN through
//   compressai/models/utils.py:128-146   conv / deconv (k5 s2, k3 s1)
//   compressai/layers/gdn.py:77-92       the 1x1 channel contraction of GDN / IGDN
// and their autograd backward passes.
//
// One CTA = one 128 x BLOCK_N output tile.  Warp roles (192 threads):
//   warps 0-3  gather the activation operand with cp.async (zero-filled padding, strided /
//              sub-pixel-phase taps, two-source channel concat) into SWIZZLE_128B shared
//              memory, then run the epilogue out of TMEM (tcgen05.ld)
//   warp 4     TMA producer for the weight operand (cp.async.bulk.tensor, SWIZZLE_128B)
//   warp 5     TMEM allocator + single-thread tcgen05.mma issuer
// Pipelines: smem full/empty mbarrier ring (STAGES deep) and one TMEM-full barrier.
#include "igemm.cuh"

namespace lmsfc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;            // 64 bf16 = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
constexpr int NUM_GATHER_THREADS = 128;
constexpr int NUM_THREADS = 192;
constexpr int GATHER_LAG = 2;          // cp.async groups kept in flight per thread

template <int BLOCK_N>
struct FwdCfg {
  static constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (BLOCK_N == 256) ? 4 : 6;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

template <int BLOCK_N>
__global__ void __launch_bounds__(NUM_THREADS, 1) igemm_fwd_kernel(const __grid_constant__ IgemmParams p) {
  using Cfg = FwdCfg<BLOCK_N>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * BLOCK_M;
  const int n0 = blockIdx.y * BLOCK_N;
  const int ctot = p.c0 + p.c1;
  const int kc_per_tap = ctot / BLOCK_K;
  const int nkb = p.ntaps * kc_per_tap;
  const uint32_t smem_base = smem_u32(smem);

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&p.tmapB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], NUM_GATHER_THREADS + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc<BLOCK_N>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ================= activation gather (producer) =================
    const int chunk = tid & 7, r0 = tid >> 3;
    int pixbase[8], hi0[8], wi0[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = m0 + r0 + 16 * i;
      if (m < p.M) {
        const int wg = m % p.Wg, t = m / p.Wg;
        const int hg = t % p.Hg, b = t / p.Hg;
        pixbase[i] = b * p.Hin * p.Win;
        hi0[i] = hg * p.sin;
        wi0[i] = wg * p.sin;
      } else {
        pixbase[i] = 0; hi0[i] = -100000; wi0[i] = -100000;
      }
    }
    const uint32_t dst_thread = smem_base + r0 * 128 + ((chunk ^ (r0 & 7)) << 4);
    int stage = 0, phase = 0, tap = 0, kc = 0;
    for (int kb = 0; kb < nkb; ++kb) {
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
        const __nv_bfloat16* g =
            ok ? src + ((size_t)(pixbase[i] + hi * p.Win + wi) * C + coff) : src;
        cp_async16(dst + i * 2048, g, ok ? 16u : 0u);
      }
      cp_async_commit();
      if (kb >= GATHER_LAG) {
        cp_async_wait<GATHER_LAG>();
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[(kb - GATHER_LAG) % STAGES]);
      }
      if (++kc == kc_per_tap) { kc = 0; ++tap; }
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    if (nkb >= 2) {
      cp_async_wait<1>();
      fence_proxy_async_smem();
      mbar_arrive(&full_bar[(nkb - 2) % STAGES]);
    }
    cp_async_wait<0>();
    fence_proxy_async_smem();
    mbar_arrive(&full_bar[(nkb - 1) % STAGES]);

    // ================= epilogue: TMEM -> registers -> global =================
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    const int m = m0 + row;
    bool ok = m < p.M;
    size_t obase = 0;
    if (ok) {
      const int wg = m % p.Wg, t = m / p.Wg;
      const int hg = t % p.Hg, b = t / p.Hg;
      const int ho = hg * p.sout + p.oy, wo = wg * p.sout + p.ox;
      ok = ho < p.Hout && wo < p.Wout;
      obase = ((size_t)(b * p.Hout + ho) * p.Wout + wo) * p.N + n0;
    }
#pragma unroll 1
    for (int cb = 0; cb < BLOCK_N; cb += 32) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + cb, r);
      tmem_ld_wait();
      if (!ok) continue;
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
      const size_t o = obase + cb;
      if (p.epi == EPI_BIAS) {
        if (p.bias) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + n0 + cb + j);
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (p.out_f32) {
          float4* d = reinterpret_cast<float4*>(p.out_f32 + o);
#pragma unroll
          for (int j = 0; j < 8; ++j) d[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
        if (p.out_abs) {
          uint4* d = reinterpret_cast<uint4*>(p.out_abs + o);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            d[j] = make_uint4(pack_bf16(fabsf(v[8 * j]), fabsf(v[8 * j + 1])),
                              pack_bf16(fabsf(v[8 * j + 2]), fabsf(v[8 * j + 3])),
                              pack_bf16(fabsf(v[8 * j + 4]), fabsf(v[8 * j + 5])),
                              pack_bf16(fabsf(v[8 * j + 6]), fabsf(v[8 * j + 7])));
        }
      } else {
        // GDN epilogues read x (and gy, rn) at the output coordinates
        uint32_t xr[16];
        {
          const uint4* s = reinterpret_cast<const uint4*>(p.aux_x + o);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint4 q = __ldg(s + j);
            xr[4 * j] = q.x; xr[4 * j + 1] = q.y; xr[4 * j + 2] = q.z; xr[4 * j + 3] = q.w;
          }
        }
        if (p.epi == EPI_GDN) {
          float rn[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float nrm = v[j] + __ldg(p.bias + n0 + cb + j);
            rn[j] = rsqrtf(nrm);
            const float x = (j & 1) ? bf16_hi(xr[j >> 1]) : bf16_lo(xr[j >> 1]);
            v[j] = p.inverse ? x * (nrm * rn[j]) : x * rn[j];
          }
          if (p.out_rn) {
            uint4* d = reinterpret_cast<uint4*>(p.out_rn + o);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              d[j] = make_uint4(pack_bf16(rn[8 * j], rn[8 * j + 1]), pack_bf16(rn[8 * j + 2], rn[8 * j + 3]),
                                pack_bf16(rn[8 * j + 4], rn[8 * j + 5]), pack_bf16(rn[8 * j + 6], rn[8 * j + 7]));
          }
        } else {  // EPI_GDN_BWD
          uint32_t gr[16], rr[16];
          const uint4* sg = reinterpret_cast<const uint4*>(p.aux_gy + o);
          const uint4* sr = reinterpret_cast<const uint4*>(p.aux_rn + o);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint4 q = __ldg(sg + j);
            gr[4 * j] = q.x; gr[4 * j + 1] = q.y; gr[4 * j + 2] = q.z; gr[4 * j + 3] = q.w;
            const uint4 w = __ldg(sr + j);
            rr[4 * j] = w.x; rr[4 * j + 1] = w.y; rr[4 * j + 2] = w.z; rr[4 * j + 3] = w.w;
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float x = (j & 1) ? bf16_hi(xr[j >> 1]) : bf16_lo(xr[j >> 1]);
            const float gy = (j & 1) ? bf16_hi(gr[j >> 1]) : bf16_lo(gr[j >> 1]);
            const float rn = (j & 1) ? bf16_hi(rr[j >> 1]) : bf16_lo(rr[j >> 1]);
            const float f = p.inverse ? 1.f / rn : rn;
            v[j] = gy * f + 2.f * x * v[j];
          }
        }
      }
      if (p.out_bf16) {
        uint4* d = reinterpret_cast<uint4*>(p.out_bf16 + o);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          d[j] = make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                            pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7]));
      }
    }
  } else if (warp == 4) {
    // ================= weight TMA producer =================
    if (lane == 0) {
      int stage = 0, phase = 0, tap = 0, kc = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::B_STAGE_BYTES);
        const int kcoord = (int)p.widx[tap] * ctot + kc * BLOCK_K;
        tma_load_2d(&p.tmapB, &full_bar[stage], smem_base + stage * Cfg::STAGE_BYTES + A_STAGE_BYTES,
                    kcoord, n0);
        if (++kc == kc_per_tap) { kc = 0; ++tap; }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ================= MMA issuer (one thread) =================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BLOCK_N, 0, 0);
      int stage = 0, phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
        const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < BLOCK_K / 16; ++k) {
          const uint64_t ad = make_sdesc_sw128(sa + k * 32, 16, 1024);
          const uint64_t bd = make_sdesc_sw128(sb + k * 32, 16, 1024);
          umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(tmem_full_bar);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc<BLOCK_N>(tmem_base);
  }
}

template <int BLOCK_N>
static int launch_fwd_t(const IgemmParams& p, cudaStream_t stream) {
  using Cfg = FwdCfg<BLOCK_N>;
  static bool attr_done = false;
  if (!attr_done) {
    LMSFC_CUDA(cudaFuncSetAttribute(igemm_fwd_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::SMEM_BYTES));
    attr_done = true;
  }
  dim3 grid(ceil_div(p.M, BLOCK_M), p.N / BLOCK_N);
  igemm_fwd_kernel<BLOCK_N><<<grid, NUM_THREADS, Cfg::SMEM_BYTES, stream>>>(p);
  LMSFC_LAUNCH_CHECK();
  return LMSFC_OK;
}

int igemm_block_n(int M, int N) {
  // 256-wide tiles halve the gather traffic per flop; use them once the grid fills the 148 SMs.
  if (N % 256 == 0 && (int64_t)ceil_div(M, BLOCK_M) * (N / 256) >= 120) return 256;
  return 128;
}

int launch_igemm(IgemmParams& p, cudaStream_t stream) {
  LMSFC_CHECK_ARG(p.M > 0 && p.N > 0 && p.N % 128 == 0, "igemm: N=%d must be a multiple of 128", p.N);
  LMSFC_CHECK_ARG(p.c0 % 64 == 0 && p.c1 % 64 == 0 && p.c0 > 0, "igemm: channels (%d,%d) must be multiples of 64",
                  p.c0, p.c1);
  LMSFC_CHECK_ARG(p.ntaps >= 1 && p.ntaps <= kMaxTaps, "igemm: ntaps=%d", p.ntaps);
  LMSFC_CHECK_ARG(p.w_rows >= p.N && p.w_cols % 64 == 0, "igemm: weight matrix %dx%d", p.w_rows, p.w_cols);
  const int bn = igemm_block_n(p.M, p.N);
  int rc = make_tmap_2d_bf16(&p.tmapB, p.w, (uint64_t)p.w_rows, (uint64_t)p.w_cols, (uint64_t)p.w_cols, bn, 64);
  if (rc) return rc;
  if (bn == 256) return launch_fwd_t<256>(p, stream);
  return launch_fwd_t<128>(p, stream);
}

}  // namespace lmsfc
