// tcgen05 weight-gradient GEMM (K = pixels) for sm_100a.
//
//   G[d][widx(tap)][g] += sum_pix dense[pix][d] * gath[pixel(pix) + tap][g]
//
// Backward-by-weights of compressai/models/utils.py:128-146 (conv: dense = grad_out, gathered =
// input; deconv: dense = input, gathered = grad_out) and of GDN's gamma (layers/gdn.py:86).
// Both operands are channel-contiguous NHWC, i.e. MN-major for the MMA: the dense side arrives
// by TMA (box 64 ch x 64 px, SWIZZLE_128B), the shifted/strided side by a cp.async gather into
// the same 128-byte-row swizzled image.  One CTA = 128 dense channels x 256 gathered channels of
// one filter tap over one K split; fp32 partials leave TMEM through shared-memory boxes and TMA
// reduce-add (cp.reduce.async.bulk.tensor.2d ... .add.f32).
#include <stdlib.h>

#include "igemm.cuh"

namespace lmsfc {

namespace {
constexpr int WG_BM = 128, WG_BN = 256, WG_BK = 64;
constexpr int WG_B_BYTES = WG_BK * WG_BN * 2;   // 32 KB: 4 segments of [64 px][64 ch]
constexpr int WG_SEG_BYTES = WG_BK * 128;       // 8 KB
constexpr int WG_THREADS = 192;
// MT = number of 128-channel dense tiles per CTA.  MT = 2 keeps two 128x256 fp32 accumulators in
// the 512 TMEM columns and feeds both from ONE gathered operand tile, halving the cp.async
// traffic per flop (the gather, not the tensor pipe, bounds MT = 1).
template <int MT>
struct WgCfg {
  static constexpr int A_BYTES = MT * WG_BK * WG_BM * 2;          // 16 KB per dense tile
  static constexpr int STAGE_BYTES = A_BYTES + WG_B_BYTES;
  static constexpr int STAGES = (MT == 2) ? 3 : 4;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
  static constexpr int TMEM_COLS = MT * WG_BN;
};

__device__ __forceinline__ void wg_epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

#ifdef LMSFC_TIMELINE
// debug build only (make TIMELINE=1): %globaltimer stamps of the first 64 CTAs, read back by profiles/timeline.py
__device__ unsigned long long g_wg_timeline[64 * 16];
#define WTL(slot)                                                                                  \
  do {                                                                                             \
    const int _c = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                 \
    unsigned long long _t;                                                                         \
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(_t));                                          \
    if (_c < 64) g_wg_timeline[_c * 16 + (slot)] = _t;                                             \
    if (_c < 64 && (slot) >= 9) g_wg_timeline[_c * 16 + (slot) + 2] = (unsigned long long)clock64(); \
  } while (0)
#else
#define WTL(slot) do {} while (0)
#endif
}  // namespace

__device__ __forceinline__ int fast_div(int n, uint32_t mul, uint32_t shr) {
  return mul ? (int)(__umulhi((uint32_t)n, mul) >> shr) : n;
}

// Bias gradient = column sums of grad_out [P][256] (bf16), by CTA `idx` of `n` spare CTAs riding in the GEMM's wave:
// 32 channel groups x 6 row lanes of 16-byte loads, lanes combined in shared memory, 256 float atomics per CTA.
__device__ __forceinline__ void wg_bias_cta(const WgradParams& p, int idx, int n, float* red) {
  const int tid = threadIdx.x, c8 = tid & 31, rl = tid >> 5;
  const long long per = (p.bias_P + n - 1) / n;
  const long long r0 = (long long)idx * per, r1 = min(p.bias_P, r0 + per);
  const uint4* __restrict__ g = reinterpret_cast<const uint4*>(p.bias_src);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 16
  for (long long r = r0 + rl; r < r1; r += WG_THREADS / 32) {
    const uint4 a = __ldg(g + r * 32 + c8);
    const uint32_t av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      acc[2 * j] += __uint_as_float(av[j] << 16);
      acc[2 * j + 1] += __uint_as_float(av[j] & 0xffff0000u);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[rl * 256 + c8 * 8 + j] = acc[j];
  __syncthreads();
  for (int c = tid; c < 256; c += WG_THREADS) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < WG_THREADS / 32; ++k) v += red[k * 256 + c];
    atomicAdd(p.gbias + c, v);
  }
}

template <int MT>
__global__ void __launch_bounds__(WG_THREADS, 1) wgrad_kernel(const __grid_constant__ WgradParams p) {
  using Cfg = WgCfg<MT>;
  constexpr int WG_STAGES = Cfg::STAGES;
  constexpr int WG_STAGE_BYTES = Cfg::STAGE_BYTES;
  constexpr int WG_A_BYTES = Cfg::A_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + WG_STAGES * WG_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + WG_STAGES;
  uint64_t* tmem_full_bar = empty_bar + WG_STAGES;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

  const uint32_t smem_base = smem_u32(smem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if ((int)blockIdx.x >= p.gemm_ctas) {    // spare CTA of the wave: bias gradient
    pdl_sync();
    wg_bias_cta(p, (int)blockIdx.x - p.gemm_ctas, p.bias_ctas, reinterpret_cast<float*>(smem));
    return;
  }
  const int bx = (int)blockIdx.x % p.grid_x, byz = (int)blockIdx.x / p.grid_x;
  const int by = byz % p.grid_y, bz = byz / p.grid_y;
  const int ntile_per_tap = p.Cg / WG_BN;
  const int tap = bx / ntile_per_tap;
  const int n0 = (bx % ntile_per_tap) * WG_BN;            // gathered-channel offset
  const int d0 = by * WG_BM * MT;                         // dense-channel offset
  const int kb_begin = bz * p.kb_per_split;
  const int kb_end = min(kb_begin + p.kb_per_split, p.kb_total);
  const int nkb = kb_end - kb_begin;
  if (nkb <= 0) return;
  if (tid == 0) WTL(0);

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&p.tmapA);
    tma_prefetch_desc(&p.tmapOut);
    for (int s = 0; s < WG_STAGES; ++s) {
      mbar_init(&full_bar[s], 128 + 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();                            // everything above touched no global memory
  if (tid == 0) WTL(1);
#ifdef LMSFC_TIMELINE
  const int dbg = p.dbg;                 // bit0: skip the gather, bit1: skip the TMA loads (timing experiments)
#else
  constexpr int dbg = 0;
#endif

  if (warp < 4) {
    // ============ gather of the shifted operand: 64 pixel rows x 256 channels per K block ============
    const int chunk = tid & 7, r0 = tid >> 3;   // rows r0 + 16*i, i < 4
    const bool first = n0 < p.gc0;
    const __nv_bfloat16* src = first ? p.g0 : p.g1;
    const int C = first ? p.gc0 : p.gc1;
    const int coff = (first ? n0 : n0 - p.gc0) + chunk * 8;
    const int dy = p.dy[tap], dx = p.dx[tap];
    const uint32_t dst_thread = smem_base + WG_A_BYTES + r0 * 128 + ((chunk ^ (r0 & 7)) << 4);
    int stage = 0, phase = 0;
    for (int kb = 0; kb < nkb; ++kb) {
      const int pix0 = (kb_begin + kb) * WG_BK;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      const uint32_t dst = dst_thread + stage * WG_STAGE_BYTES;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = pix0 + r0 + 16 * i;
        bool ok = m < p.M;
        const __nv_bfloat16* g = src;
        if (ok) {
          const int t = fast_div(m, p.wg_mul, p.wg_shr), wg = m - t * p.Wg;
          const int b = fast_div(t, p.hg_mul, p.hg_shr), hg = t - b * p.Hg;
          const int hi = hg * p.sin + dy, wi = wg * p.sin + dx;
          ok = (unsigned)hi < (unsigned)p.Hin && (unsigned)wi < (unsigned)p.Win;
          if (ok) g = src + ((size_t)((b * p.Hin + hi) * p.Win + wi) * C + coff);
        }
        if (dbg & 1) continue;
#pragma unroll
        for (int seg = 0; seg < 4; ++seg)
          cp_async16(dst + seg * WG_SEG_BYTES + i * 2048, g + seg * 64, ok ? 16u : 0u);
      }
      cp_async_mbar_arrive_noinc(&full_bar[stage]);
      if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
    }
    cp_async_wait_all();

    // ============ epilogue: fp32 partial tile -> smem box -> TMA reduce-add ============
    // thread = accumulator row (dense channel); a 128 x 32 fp32 chunk is staged as a SWIZZLE_128B box in
    // the (now idle) pipeline memory and added to the gradient by ONE bulk tensor operation: coalesced
    // L2 traffic instead of 32 scattered 16-byte atomics per warp instruction.  4 boxes rotate.
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    if (tid == 0) WTL(4);
    const int row = warp * 32 + lane;
    const int out_col0 = (int)p.widx[tap] * p.Cg + n0;
    int it = 0;
#pragma unroll 1
    for (int mt = 0; mt < MT; ++mt) {
#pragma unroll 1
      for (int cb = 0; cb < WG_BN; cb += 32, ++it) {
        const uint32_t box = smem_base + (uint32_t)(it & 3) * (WG_BM * 128);
        if (it >= 4) {                      // the bulk op that last read this box has finished reading
          if (tid == 0) bulk_wait_group_read<3>();
          wg_epi_bar();
        }
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + mt * WG_BN + cb, r);
        tmem_ld_wait();
        const uint32_t dst = box + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + ((j ^ (row & 7)) << 4)), "r"(r[4 * j]),
                       "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                       : "memory");
        fence_proxy_async_smem();
        wg_epi_bar();
        if (tid == 0) {
          tma_reduce_add_2d(&p.tmapOut, box, out_col0 + cb, d0 + mt * WG_BM);
          bulk_commit_group();
        }
      }
    }
    if (tid == 0) WTL(5);
    if (tid == 0) bulk_wait_group<0>();
    if (tid == 0) WTL(7);
  } else if (warp == 4) {
    // ============ dense operand by TMA: 2*MT [64 px][64 ch] boxes per K block ============
    if (lane == 0) {
      int stage = 0, phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], (dbg & 2) ? 0 : WG_A_BYTES);
        if (dbg & 2) { if (++stage == WG_STAGES) { stage = 0; phase ^= 1; } continue; }
        const int pix0 = (kb_begin + kb) * WG_BK;
        const uint32_t sa = smem_base + stage * WG_STAGE_BYTES;
#pragma unroll
        for (int seg = 0; seg < 2 * MT; ++seg)
          tma_load_2d(&p.tmapA, &full_bar[stage], sa + seg * WG_SEG_BYTES, d0 + 64 * seg, pix0);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ============ MMA issuer: both operands MN-major (channels contiguous, K = pixel rows) ============
    if (lane == 0) {
#ifdef LMSFC_TIMELINE
      // timing experiment (results are garbage): bit2 / bit3 read the A / B stage image as a K-major tile
      const uint32_t idesc = make_idesc_bf16(WG_BM, WG_BN, (dbg & 4) ? 0 : 1, (dbg & 8) ? 0 : 1);
#else
      constexpr uint32_t idesc = make_idesc_bf16(WG_BM, WG_BN, 1, 1);
#endif
      int stage = 0, phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (kb == 0) WTL(9);
        const uint32_t sa = smem_base + stage * WG_STAGE_BYTES;
        const uint32_t sb = sa + WG_A_BYTES;
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
          for (int k = 0; k < WG_BK / 16; ++k) {
            // K step of 16 pixel rows = 2 KB; LBO = next 64-channel segment, SBO = next 8 rows
            uint64_t bd = make_sdesc_sw128(sb + k * 2048, WG_SEG_BYTES, 1024);
            if (dbg & 8) bd = make_sdesc_sw128(sb + k * 32, 16, 1024);
            uint64_t ad = make_sdesc_sw128(sa + mt * 2 * WG_SEG_BYTES + k * 2048, WG_SEG_BYTES, 1024);
            if (dbg & 4) ad = make_sdesc_sw128(sa + mt * 2 * WG_SEG_BYTES + k * 32, 16, 1024);
            umma_bf16(tmem_base + ((dbg & 16) ? 0 : mt * WG_BN), ad, bd, idesc, (kb | k) != 0);
          }
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(tmem_full_bar);
      WTL(10);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  }
  if (tid == 0) WTL(8);
}

#ifdef LMSFC_TIMELINE
extern "C" int lmsfc_debug_read_wg_timeline(unsigned long long* host64x16) {
  return (int)cudaMemcpyFromSymbol(host64x16, g_wg_timeline, sizeof(unsigned long long) * 64 * 16);
}
#endif

int launch_wgrad(WgradParams& p, cudaStream_t stream) {
  LMSFC_CHECK_ARG(p.Cd % 128 == 0 && p.Cg % 256 == 0, "wgrad: Cd=%d must be a multiple of 128, Cg=%d of 256", p.Cd,
                  p.Cg);
  LMSFC_CHECK_ARG(p.gc0 + p.gc1 == p.Cg && p.gc0 % 256 == 0, "wgrad: gathered channels %d+%d != %d", p.gc0, p.gc1,
                  p.Cg);
  LMSFC_CHECK_ARG(p.ntaps >= 1 && p.ntaps <= kMaxTaps && p.M > 0, "wgrad: ntaps=%d M=%d", p.ntaps, p.M);
  static bool attr_done = false;
  if (!attr_done) {
    LMSFC_CUDA(cudaFuncSetAttribute(wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, WgCfg<1>::SMEM_BYTES));
    LMSFC_CUDA(cudaFuncSetAttribute(wgrad_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, WgCfg<2>::SMEM_BYTES));
    attr_done = true;
  }
  int rc = make_tmap_2d_bf16(&p.tmapA, p.dense, (uint64_t)p.M, (uint64_t)p.Cd, (uint64_t)p.Cd, 64, 64);
  if (rc) return rc;
  rc = make_tmap_2d_f32(&p.tmapOut, p.out, (uint64_t)p.Cd, (uint64_t)p.T * p.Cg, (uint64_t)p.T * p.Cg, WG_BM, 32);
  if (rc) return rc;
  int mt = (p.Cd % 256 == 0) ? 2 : 1;
#ifdef LMSFC_TIMELINE
  { const char* e = getenv("LMSFC_WG_MT"); if (e) mt = atoi(e); }
#endif
  p.kb_total = ceil_div(p.M, WG_BK);
#ifdef LMSFC_TIMELINE
  { const char* e = getenv("LMSFC_WG_DBG"); p.dbg = e ? atoi(e) : 0; }
#endif
  const int tiles = p.ntaps * (p.Cg / WG_BN) * (p.Cd / (WG_BM * mt));
  fast_div_magic(p.Wg, &p.wg_mul, &p.wg_shr);
  fast_div_magic(p.Hg, &p.hg_mul, &p.hg_shr);
  // bias CTAs: about one per 512 KB of grad_out (a lone SM streams ~0.1 TB/s), at most 16, inside the same wave
  int spare = 0;
  if (p.gbias) spare = (int)max(1ll, min(16ll, (p.bias_P * 512 + (512 << 10) - 1) / (512 << 10)));
  // split K (pixels) so that the grid fills, but does not exceed, one wave of the 148 SMs
  int ksplit = max(1, (148 - spare) / tiles);
  ksplit = max(1, min(ksplit, ceil_div(p.kb_total, 4)));   // >= 4 K blocks per CTA: amortise its 256 KB reduce-add
#ifdef LMSFC_TIMELINE
  { const char* e = getenv("LMSFC_WG_KSPLIT"); if (e) ksplit = atoi(e); }
#endif
  p.kb_per_split = ceil_div(p.kb_total, ksplit);
  ksplit = ceil_div(p.kb_total, p.kb_per_split);
  p.grid_x = p.ntaps * (p.Cg / WG_BN);
  p.grid_y = p.Cd / (WG_BM * mt);
  p.gemm_ctas = p.grid_x * p.grid_y * ksplit;
  p.bias_ctas = p.gbias ? max(spare, min(148 - p.gemm_ctas, 32)) : 0;
  dim3 grid(p.gemm_ctas + p.bias_ctas);
  if (mt == 2) LMSFC_CUDA(launch_pdl(wgrad_kernel<2>, grid, dim3(WG_THREADS), (size_t)WgCfg<2>::SMEM_BYTES, stream, 1u, p));
  else         LMSFC_CUDA(launch_pdl(wgrad_kernel<1>, grid, dim3(WG_THREADS), (size_t)WgCfg<1>::SMEM_BYTES, stream, 1u, p));
  ++g_launch_count;
  return LMSFC_OK;
}

}  // namespace lmsfc
