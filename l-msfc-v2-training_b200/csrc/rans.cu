// Range-ANS entropy coding on the device, bit-exact with the reference's CPU coder
//   compressai/cpp_exts/rans/rans_interface.cpp:108-288 (encode_with_indexes / flush / decode_with_indexes)
//   compressai/third_party/ryg_rans/rans64.h            (64-bit state, 32-bit words, 16-bit probabilities)
//   compressai/entropy_models/entropy_models.py:235-325,684-689 (compress / decompress / build_indexes)
//
// A string is inherently serial (every record depends on the coder state), so parallelism comes from strings:
// one THREAD codes one lane, lane k of K of an item = symbols[k::K] of that item coded exactly like the reference
// codes a whole item.  K = 1 reproduces the reference's bytes; K > 1 is this library's interleaved container
// (K strings + a length table per item, see coder.py).  Consecutive lanes sit in consecutive threads, so the
// strided symbol / index reads of a warp coalesce.  Per record the state-dependent chain is one 64-bit division
// (encode) or one table search (decode); symbol, index and CDF loads do not depend on the state and are issued a
// chunk ahead.  The division is done exactly with an fp64 reciprocal estimate and an integer fix-up
// (the quotient is < 2^47, so the estimate is off by at most one).
#include "../../include/b200lmsfc.h"
#include "kernels.cuh"

namespace lmsfc {

static constexpr uint32_t kPrecision = 16, kBypassBits = 4, kBypassMax = 15;
static constexpr uint64_t kRansLow = 1ull << 31;
static constexpr int kChunk = 4;

struct RansTables {
  const int32_t* cdfs;     // [rows][cdf_stride]
  const int32_t* sizes;    // [rows]  entries used in each row (entropy_models.py `_cdf_length`)
  const int32_t* offsets;  // [rows]
  int rows, cdf_stride;
};

// x / freq for x < 2^47 * freq, freq < 2^17 (rans64.h:89), exact
__device__ __forceinline__ void divmod(uint64_t x, uint32_t freq, double rcp, uint64_t& q, uint32_t& r) {
  uint64_t qe = (uint64_t)__double2ull_rz(__ull2double_rz(x) * rcp);
  int64_t rem = (int64_t)(x - qe * freq);
  while (rem < 0) { --qe; rem += freq; }
  while (rem >= (int64_t)freq) { ++qe; rem -= freq; }
  q = qe;
  r = (uint32_t)rem;
}

struct Enc {
  uint64_t x;
  uint32_t* p;
  __device__ __forceinline__ void put(uint32_t start, uint32_t freq, double rcp) {      // rans64.h:75-90
    const uint64_t lim = ((kRansLow >> kPrecision) << 32) * freq;
    if (x >= lim) { *--p = (uint32_t)x; x >>= 32; }
    uint64_t q; uint32_t r;
    divmod(x, freq, rcp, q, r);
    x = (q << kPrecision) + r + start;
  }
  __device__ __forceinline__ void put_bits(uint32_t val) {                              // rans_interface.cpp:68-86
    if (x >= (1ull << 59)) { *--p = (uint32_t)x; x >>= 32; }
    x = (x << kBypassBits) | val;
  }
};

// status: 0 ok, 1 index outside the table, 2 scratch exhausted
__global__ void __launch_bounds__(128) rans_encode_kernel(const int32_t* __restrict__ symbols,
                                                          const int32_t* __restrict__ indexes, int64_t n, int items,
                                                          int lanes, RansTables tb, uint32_t* scratch, int64_t lane_cap,
                                                          uint32_t* lane_words, int* status) {
  pdl_sync();
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)items * lanes) return;
  const int item = (int)(t / lanes), lane = (int)(t % lanes);
  const int64_t count = lane < n ? (n - lane + lanes - 1) / lanes : 0;
  const int32_t* sym = symbols + (int64_t)item * n;
  const int32_t* idx = indexes + (int64_t)item * n;
  uint32_t* const base = scratch + t * lane_cap;
  Enc e{kRansLow, base + lane_cap};
  int bad = 0;
  for (int64_t hi = count; hi > 0; hi -= kChunk) {        // records are coded last-to-first (flush, :189-199)
    uint32_t start[kChunk], freq[kChunk], raw[kChunk];
    double rcp[kChunk];
    bool esc[kChunk];
#pragma unroll
    for (int u = 0; u < kChunk; ++u) {                    // state-independent part, loads in flight together
      const int64_t tt = hi - 1 - u;
      start[u] = 0; freq[u] = 0; raw[u] = 0; esc[u] = false; rcp[u] = 0.0;
      if (tt < 0) continue;
      const int64_t i = lane + tt * lanes;
      const int32_t row = __ldg(idx + i);
      if (row < 0 || row >= tb.rows) { bad = 1; continue; }
      const int32_t max_value = __ldg(tb.sizes + row) - 2;
      int32_t v = __ldg(sym + i) - __ldg(tb.offsets + row);
      if (max_value < 0 || max_value + 1 >= tb.cdf_stride) { bad = 1; continue; }
      if (v < 0) { raw[u] = (uint32_t)(-2 * v - 1); v = max_value; }                    // :136-142
      else if (v >= max_value) { raw[u] = (uint32_t)(2 * (v - max_value)); v = max_value; }
      esc[u] = (v == max_value);
      const int32_t* c = tb.cdfs + (int64_t)row * tb.cdf_stride + v;
      const int32_t c0 = __ldg(c), c1 = __ldg(c + 1);
      start[u] = (uint32_t)c0;
      freq[u] = (uint32_t)(c1 - c0);
      if (freq[u] == 0 || freq[u] > 65536u) { bad = 1; freq[u] = 0; continue; }
      rcp[u] = 1.0 / (double)freq[u];
    }
    if (e.p - base < 3 * kChunk + 2) { bad = 2; break; }
#pragma unroll
    for (int u = 0; u < kChunk; ++u) {
      if (freq[u] == 0) continue;
      if (esc[u]) {                                       // bypass digits, :147-175 reversed
        int nb = 0;
        while (((uint64_t)raw[u] >> (nb * kBypassBits)) != 0) ++nb;
        for (int j = nb - 1; j >= 0; --j) e.put_bits((raw[u] >> (j * kBypassBits)) & kBypassMax);
        e.put_bits((uint32_t)nb % kBypassMax);
        for (int q = nb / (int)kBypassMax; q > 0; --q) e.put_bits(kBypassMax);
      }
      e.put(start[u], freq[u], rcp[u]);
    }
  }
  e.p -= 2;                                               // rans64.h:93-100
  e.p[0] = (uint32_t)e.x;
  e.p[1] = (uint32_t)(e.x >> 32);
  lane_words[t] = (uint32_t)(base + lane_cap - e.p);
  if (bad) atomicMax(status, bad);
}

// One warp per lane: exclusive offset = sum of the lane lengths in front of it, then the lane's words move from
// the tail of its scratch slot to the packed stream.  lane_offsets[T] receives the total.
__global__ void __launch_bounds__(256) rans_pack_kernel(const uint32_t* __restrict__ scratch, int64_t lane_cap,
                                                        const uint32_t* __restrict__ lane_words, int64_t T,
                                                        uint32_t* packed, int64_t packed_cap, uint32_t* lane_offsets,
                                                        int* status) {
  pdl_sync();
  const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int l = threadIdx.x & 31;
  if (t >= T) return;
  uint64_t acc = 0;
  for (int64_t j = l; j < t; j += 32) acc += lane_words[j];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  const uint32_t w = lane_words[t];
  if (l == 0) {
    lane_offsets[t] = (uint32_t)acc;
    if (t == T - 1) lane_offsets[T] = (uint32_t)(acc + w);
  }
  if ((int64_t)(acc + w) > packed_cap) { if (l == 0) atomicMax(status, 2); return; }
  const uint32_t* src = scratch + (t + 1) * lane_cap - w;
  for (uint32_t j = l; j < w; j += 32) packed[acc + j] = src[j];
}

struct Dec {
  uint64_t x;
  const uint32_t* p;
  const uint32_t* end;
  __device__ __forceinline__ void refill() {
    if (x < kRansLow) { x = (x << 32) | (p < end ? __ldg(p) : 0u); ++p; }
  }
  __device__ __forceinline__ uint32_t get_bits() {                                      // rans_interface.cpp:88-104
    const uint32_t v = (uint32_t)(x & kBypassMax);
    x >>= kBypassBits;
    refill();
    return v;
  }
};

// largest s in [0, size-2] with cdf[s] <= cf (std::find_if of :244-248).  Two probes around the centre of the
// row (where a trained latent's mass sits) settle most symbols in one memory round trip; the rest bisect.
__device__ __forceinline__ int cdf_search(const int32_t* __restrict__ cdf, int size, uint32_t cf) {
  int lo = 0, hi = size - 1;                 // invariant: cdf[lo] <= cf < cdf[hi]  (cdf[size-1] = 2^16)
  const int m = (size - 2) >> 1;
  if (m > 0 && m + 1 < hi) {
    const uint32_t a = (uint32_t)__ldg(cdf + m), b = (uint32_t)__ldg(cdf + m + 1);
    if (b <= cf) lo = m + 1;
    else if (a <= cf) { lo = m; hi = m + 1; }
    else hi = m;
  }
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((uint32_t)__ldg(cdf + mid) <= cf) lo = mid; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(128) rans_decode_kernel(const uint32_t* __restrict__ packed, int64_t packed_words,
                                                          const uint32_t* __restrict__ lane_offsets,
                                                          const int32_t* __restrict__ indexes, int64_t n, int items,
                                                          int lanes, RansTables tb, int32_t* out, int* status) {
  pdl_sync();
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)items * lanes) return;
  const int item = (int)(t / lanes), lane = (int)(t % lanes);
  const int64_t count = lane < n ? (n - lane + lanes - 1) / lanes : 0;
  const int32_t* idx = indexes + (int64_t)item * n;
  int32_t* dst = out + (int64_t)item * n;
  const uint32_t* w = packed + lane_offsets[t];
  Dec d;
  d.end = packed + packed_words;
  if (w + 2 > d.end) { atomicMax(status, 2); return; }
  d.x = (uint64_t)__ldg(w) | ((uint64_t)__ldg(w + 1) << 32);                            // rans64.h:104-112
  d.p = w + 2;
  int bad = 0;
  for (int64_t lo = 0; lo < count; lo += kChunk) {
    int32_t row[kChunk], size[kChunk], off[kChunk];
#pragma unroll
    for (int u = 0; u < kChunk; ++u) {
      row[u] = -1; size[u] = 0; off[u] = 0;
      if (lo + u >= count) continue;
      const int32_t r = __ldg(idx + lane + (lo + u) * lanes);
      if (r < 0 || r >= tb.rows) { bad = 1; continue; }
      row[u] = r;
      size[u] = __ldg(tb.sizes + r);
      off[u] = __ldg(tb.offsets + r);
      if (size[u] < 2 || size[u] > tb.cdf_stride) { bad = 1; row[u] = -1; }
    }
#pragma unroll
    for (int u = 0; u < kChunk; ++u) {
      if (lo + u >= count) break;
      int32_t value = 0;
      if (row[u] >= 0) {
        const int32_t* cdf = tb.cdfs + (int64_t)row[u] * tb.cdf_stride;
        const int32_t max_value = size[u] - 2;
        const uint32_t cf = (uint32_t)(d.x & 0xFFFFu);
        const int s = cdf_search(cdf, size[u], cf);
        const uint32_t start = (uint32_t)__ldg(cdf + s), freq = (uint32_t)__ldg(cdf + s + 1) - start;
        d.x = (uint64_t)freq * (d.x >> kPrecision) + cf - start;                        // rans64.h:120-137
        d.refill();
        value = s;
        if (s == max_value) {                                                           // :254-275
          uint32_t v = d.get_bits();
          int nb = (int)v;
          while (v == kBypassMax && nb < 64) { v = d.get_bits(); nb += (int)v; }
          uint32_t raw = 0;
          for (int j = 0; j < nb; ++j) {
            const uint32_t dg = d.get_bits();
            if (j < 8) raw |= dg << (j * kBypassBits);
          }
          value = (int32_t)(raw >> 1);
          value = (raw & 1u) ? -value - 1 : value + max_value;
        }
        value += off[u];
      }
      dst[lane + (lo + u) * lanes] = value;
    }
  }
  if (bad) atomicMax(status, bad);
}

// GaussianConditional.build_indexes (entropy_models.py:684-689): T-1 minus the number of table entries (all
// but the last) that are >= max(scale, bound).
__global__ void __launch_bounds__(256) gc_indexes_kernel(const float* __restrict__ scales, int64_t n,
                                                         const float* __restrict__ table, int T, float bound,
                                                         int32_t* __restrict__ out) {
  pdl_sync();
  __shared__ float tab[1024];
  for (int i = threadIdx.x; i < T; i += blockDim.x) tab[i] = table[i];
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float s = fmaxf(scales[i], bound);
    int k = T - 1;
    for (int j = 0; j < T - 1; ++j) k -= (s <= tab[j]) ? 1 : 0;
    out[i] = k;
  }
}

static int check_tables(const RansTables& tb, const char* who) {
  LMSFC_CHECK_ARG(tb.cdfs && tb.sizes && tb.offsets && tb.rows > 0 && tb.cdf_stride >= 2, "%s: bad CDF tables", who);
  return LMSFC_OK;
}

}  // namespace lmsfc
using namespace lmsfc;
#define ST(s) ((cudaStream_t)(s))

extern "C" int64_t lmsfc_rans_lane_cap(int64_t n, int lanes) {
  if (lanes < 1) lanes = 1;
  return 2 * ((n + lanes - 1) / lanes) + 16;      // <= 52 bits per symbol (16 + 9 bypass digits) + flush + slack
}

extern "C" int lmsfc_rans_encode(const int32_t* symbols, const int32_t* indexes, int items, int64_t n, int lanes,
                                 const int32_t* cdfs, int rows, int cdf_stride, const int32_t* sizes,
                                 const int32_t* offsets, uint32_t* scratch, int64_t lane_cap, uint32_t* lane_words,
                                 uint32_t* lane_offsets, uint32_t* packed, int64_t packed_cap, int* status,
                                 void* stream) {
  RansTables tb{cdfs, sizes, offsets, rows, cdf_stride};
  if (int rc = check_tables(tb, "rans_encode")) return rc;
  LMSFC_CHECK_ARG(items >= 0 && n >= 0 && lanes >= 1, "rans_encode: bad sizes");
  LMSFC_CHECK_ARG(scratch && lane_words && lane_offsets && packed && status, "rans_encode: null buffer");
  LMSFC_CHECK_ARG(n == 0 || (symbols && indexes), "rans_encode: null symbols / indexes");
  LMSFC_CHECK_ARG(lane_cap >= lmsfc_rans_lane_cap(n, lanes), "rans_encode: lane_cap below lmsfc_rans_lane_cap()");
  const int64_t T = (int64_t)items * lanes;
  LMSFC_CHECK_ARG(T * lane_cap < (1ll << 32), "rans_encode: more than 2^32 words");
  if (T == 0) return LMSFC_OK;
  LMSFC_LAUNCH(rans_encode_kernel, dim3(ceil_div(T, 128)), dim3(128), 0, ST(stream), symbols, indexes, n, items, lanes,
               tb, scratch, lane_cap, lane_words, status);
  LMSFC_LAUNCH(rans_pack_kernel, dim3(ceil_div(T * 32, 256)), dim3(256), 0, ST(stream), (const uint32_t*)scratch,
               lane_cap, (const uint32_t*)lane_words, T, packed, packed_cap, lane_offsets, status);
  return LMSFC_OK;
}

extern "C" int lmsfc_rans_decode(const uint32_t* packed, int64_t packed_words, const uint32_t* lane_offsets,
                                 const int32_t* indexes, int items, int64_t n, int lanes, const int32_t* cdfs,
                                 int rows, int cdf_stride, const int32_t* sizes, const int32_t* offsets,
                                 int32_t* symbols_out, int* status, void* stream) {
  RansTables tb{cdfs, sizes, offsets, rows, cdf_stride};
  if (int rc = check_tables(tb, "rans_decode")) return rc;
  LMSFC_CHECK_ARG(items >= 0 && n >= 0 && lanes >= 1, "rans_decode: bad sizes");
  const int64_t T = (int64_t)items * lanes;
  if (T == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG(packed && lane_offsets && status && packed_words >= 2 * T, "rans_decode: string shorter than its flush words");
  LMSFC_CHECK_ARG(n == 0 || (indexes && symbols_out), "rans_decode: null indexes / output");
  LMSFC_LAUNCH(rans_decode_kernel, dim3(ceil_div(T, 128)), dim3(128), 0, ST(stream), packed, packed_words, lane_offsets,
               indexes, n, items, lanes, tb, symbols_out, status);
  return LMSFC_OK;
}

extern "C" int lmsfc_gc_build_indexes(const float* scales, int64_t n, const float* scale_table, int table_len,
                                      float scale_bound, int32_t* indexes, void* stream) {
  LMSFC_CHECK_ARG(table_len >= 1 && table_len <= 1024, "gc_build_indexes: scale table of %d entries", table_len);
  if (n == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG(scales && scale_table && indexes && n > 0, "gc_build_indexes: null operand");
  const int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  LMSFC_LAUNCH(gc_indexes_kernel, dim3(grid), dim3(256), 0, ST(stream), scales, n, scale_table, table_len, scale_bound,
               indexes);
  return LMSFC_OK;
}
