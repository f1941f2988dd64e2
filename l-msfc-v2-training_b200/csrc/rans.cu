// Range-ANS entropy coding on the device, bit-exact with the reference's CPU coder
//   compressai/cpp_exts/rans/rans_interface.cpp:108-288 (encode_with_indexes / flush / decode_with_indexes)
//   compressai/third_party/ryg_rans/rans64.h            (64-bit state, 32-bit words, 16-bit probabilities)
//   compressai/entropy_models/entropy_models.py:235-325,684-689 (compress / decompress / build_indexes)
//
// A string is inherently serial (every record depends on the coder state), so parallelism comes from strings:
// lane k of K of an item = symbols[k::K] of that item, coded exactly like the reference codes a whole item.
// K = 1 reproduces the reference's bytes; K > 1 is this library's interleaved container (K strings + a length
// table per item, see coder.py).  One WARP owns one string.  Everything that does not depend on the state is done
// 128 symbols at a time by its 32 threads -- symbol / index / CDF fetches, the reciprocal of each frequency (so the
// encoder's division is a multiply-high on the serial chain) and, in the decoder, the table search itself (32 probes
// per round trip, ballot).  DESIGN.md section 3b has the measurements.
#include "../../include/b200lmsfc.h"
#include "kernels.cuh"

namespace lmsfc {

static constexpr uint32_t kPrecision = 16, kBypassBits = 4, kBypassMax = 15;
static constexpr uint64_t kRansLow = 1ull << 31;

struct RansTables {
  const int32_t* cdfs;     // [rows][cdf_stride]
  const int32_t* sizes;    // [rows]  entries used in each row (entropy_models.py `_cdf_length`)
  const int32_t* offsets;  // [rows]
  int rows, cdf_stride;
};

// Division by an invariant integer (Alverson / Granlund-Montgomery): with sh = ceil(log2 freq) and
// m = ceil(2^(63+sh) / freq), floor(x / freq) = mulhi64(x, m) >> (sh - 1) for every x < 2^63 -- the error term
// x * (m * freq - 2^(63+sh)) / (freq * 2^(63+sh)) stays below 1 / freq because 2^sh >= freq.  The coder's state is
// < 2^47 * freq <= 2^63 when the division happens (rans64.h:80-89), so the quotient is exact.  m needs 2^(63+sh)
// / freq, an 80-bit dividend: two rounds of schoolbook division in 64 bits.  freq = 1: m = 2^64 - 1, sh = 0 gives
// mulhi = x - 1, which the caller compensates in the additive term.
__device__ __forceinline__ uint64_t magic_reciprocal(uint32_t freq, uint32_t& sh) {
  if (freq < 2) { sh = 0; return ~0ull; }
  uint32_t lg = 32 - __clz(freq - 1);                  // ceil(log2 freq), 1..16
  const uint64_t top = 1ull << (lg + 31);              // 2^(63+lg) = top * 2^32
  const uint64_t q1 = top / freq, r1 = top - q1 * freq;
  const uint64_t q0 = ((r1 << 32) + (freq - 1)) / freq;   // ceil of the low round
  sh = lg - 1;
  return (q1 << 32) + q0;
}

// Encoder state (held by thread 0 of the warp that owns the string).
struct Enc {
  uint64_t x;
  uint32_t* p;
  bool writer;
  __device__ __forceinline__ void emit() {
    --p;
    if (writer) *p = (uint32_t)x;
    x >>= 32;
  }
  // rans64.h:75-90: x' = ((x / f) << 16) + x % f + start = x + start + (x / f) * (2^16 - f)
  // lim_hi = freq << 15: the renormalisation bound freq << 47 has a zero low word, so one 32-bit compare decides
  __device__ __forceinline__ void put(uint64_t magic, uint32_t sh, uint32_t add, uint32_t cmpl, uint32_t lim_hi) {
    if ((uint32_t)(x >> 32) >= lim_hi) emit();
    const uint64_t q = __umul64hi(x, magic) >> sh;
    x = x + add + q * cmpl;
  }
  __device__ __forceinline__ void put_bits(uint32_t val) {                              // rans_interface.cpp:68-86
    if (x >= (1ull << 59)) emit();
    x = (x << kBypassBits) | val;
  }
};

// what one symbol contributes, fetched and prepared by one thread: the reciprocal of its frequency (0 marks a
// record that must be skipped: bad index / row), additive term, 2^16 - freq, shift, high word of the
// renormalisation bound, escape payload (raw + 1)
struct __align__(16) Rec { uint64_t magic; uint32_t add, cmpl; uint32_t lim_hi, sh, esc, rare; };   // rare: 1 skip, 2 escape

__device__ __forceinline__ Rec fetch_record(const int32_t* __restrict__ sym, const int32_t* __restrict__ idx, int64_t i,
                                            const RansTables& tb, int& bad) {
  Rec r{0ull, 0u, 0u, 0u, 0u, 0u, 1u};
  const int32_t row = __ldg(idx + i);
  if (row < 0 || row >= tb.rows) { bad = 1; return r; }
  const int32_t max_value = __ldg(tb.sizes + row) - 2;
  int32_t v = __ldg(sym + i) - __ldg(tb.offsets + row);
  if (max_value < 0 || max_value + 1 >= tb.cdf_stride) { bad = 1; return r; }
  uint32_t raw = 0;
  if (v < 0) { raw = (uint32_t)(-2 * v - 1); v = max_value; }                           // :136-142
  else if (v >= max_value) { raw = (uint32_t)(2 * (v - max_value)); v = max_value; }
  const int32_t* c = tb.cdfs + (int64_t)row * tb.cdf_stride + v;
  const uint32_t c0 = (uint32_t)__ldg(c), freq = (uint32_t)__ldg(c + 1) - c0;
  if (freq == 0 || freq > 65536u || c0 > 65535u) { bad = 1; return r; }
  uint32_t sh;
  r.magic = magic_reciprocal(freq, sh);
  r.add = freq < 2 ? c0 + 65535u : c0;                 // mulhi(x, 2^64 - 1) = x - 1: (x - 1) * 65535 + x + 65535 + c0 = (x << 16) + c0
  r.cmpl = 65536u - freq;
  r.sh = sh;
  r.lim_hi = freq << 15;
  r.esc = (v == max_value) ? raw + 1 : 0u;
  r.rare = r.esc ? 2u : 0u;
  return r;
}

// One WARP per string.  Chunks of 128 symbols: every thread fetches and prepares four records (four independent
// load chains in flight) into the warp's slice of shared memory, then thread 0 codes the 128 records in order, last
// symbol first (flush, :189-199) -- a loop without collectives or loads from global memory.
// status: 0 ok, 1 index outside the table, 2 scratch exhausted
static constexpr int kEncChunk = 128;
__global__ void __launch_bounds__(128) rans_encode_kernel(const int32_t* __restrict__ symbols,
                                                          const int32_t* __restrict__ indexes, int64_t n, int items,
                                                          int lanes, RansTables tb, uint32_t* scratch, int64_t lane_cap,
                                                          uint32_t* lane_words, int* status, int ck_log2, int ncp,
                                                          uint32_t* ckpt) {
  pdl_sync();
  __shared__ Rec recs[4][kEncChunk];
  const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int j = threadIdx.x & 31;
  if (t >= (int64_t)items * lanes) return;
  Rec* const mine = recs[threadIdx.x >> 5];
  const int item = (int)(t / lanes), lane = (int)(t % lanes);
  const int64_t count = lane < n ? (n - lane + lanes - 1) / lanes : 0;
  const int32_t* sym = symbols + (int64_t)item * n;
  const int32_t* idx = indexes + (int64_t)item * n;
  uint32_t* const base = scratch + t * lane_cap;
  Enc e{kRansLow, base + lane_cap, true};
  int bad = 0;
  for (int64_t hi = count; hi > 0; hi -= kEncChunk) {
    const int m = hi < kEncChunk ? (int)hi : kEncChunk;
    Rec r[kEncChunk / 32];
#pragma unroll
    for (int k = 0; k < kEncChunk / 32; ++k) {            // slot u of the chunk = symbol hi-1-u
      const int u = j + 32 * k;
      r[k] = u < m ? fetch_record(sym, idx, lane + (hi - 1 - u) * lanes, tb, bad) : Rec{0ull, 0u, 0u, 0u, 0u, 0u, 1u};
    }
    __syncwarp();                                         // thread 0 is done with the previous chunk
#pragma unroll
    for (int k = 0; k < kEncChunk / 32; ++k) mine[j + 32 * k] = r[k];
    __syncwarp();
    if (j == 0 && e.p - base < 2 * m + 2) bad = 2;         // cannot happen with lane_cap >= lmsfc_rans_lane_cap()
    if (j == 0 && bad != 2) {
      // decoder checkpoints (lmsfc_rans_encode_indexed): the coder state right after symbol i = c * 2^ck_log2 has been
      // coded IS the decoder's state right before it decodes symbol i, and the words emitted so far are the ones the
      // decoder has not read yet.  Chunks hold 128 symbols and the interval is a multiple of 128: at most one per chunk.
      int ck_u = -1;
      if (ckpt) {
        const int64_t i_lo = hi - m;                      // symbol of slot u: hi - 1 - u, slots cover [i_lo, hi)
        const int64_t i_ck = ((hi - 1) >> ck_log2) << ck_log2;
        if (i_ck >= i_lo && i_ck > 0) ck_u = (int)(hi - 1 - i_ck);
      }
#pragma unroll 2
      for (int u = 0; u < m; ++u) {
        const Rec c = mine[u];
        if (c.rare) {                                     // one branch for both uncommon cases
          if (c.rare & 1u) {
            if (u == ck_u) {
              uint32_t* ck = ckpt + ((size_t)t * ncp + (size_t)(((hi - 1 - u) >> ck_log2) - 1)) * 3;
              ck[0] = (uint32_t)e.x; ck[1] = (uint32_t)(e.x >> 32); ck[2] = (uint32_t)(base + lane_cap - e.p);
            }
            continue;
          }
          // bypass digits, :147-175 reversed
          const uint32_t raw = c.esc - 1u;
          int nb = 0;
          while (((uint64_t)raw >> (nb * kBypassBits)) != 0) ++nb;
          for (int d = nb - 1; d >= 0; --d) e.put_bits((raw >> (d * kBypassBits)) & kBypassMax);
          e.put_bits((uint32_t)nb % kBypassMax);
          for (int q = nb / (int)kBypassMax; q > 0; --q) e.put_bits(kBypassMax);
        }
        e.put(c.magic, c.sh, c.add, c.cmpl, c.lim_hi);
        if (u == ck_u) {
          uint32_t* ck = ckpt + ((size_t)t * ncp + (size_t)(((hi - 1 - u) >> ck_log2) - 1)) * 3;
          ck[0] = (uint32_t)e.x; ck[1] = (uint32_t)(e.x >> 32); ck[2] = (uint32_t)(base + lane_cap - e.p);
        }
      }
    }
  }
  if (j == 0) {
    e.p -= 2;                                             // rans64.h:93-100
    e.p[0] = (uint32_t)e.x;
    e.p[1] = (uint32_t)(e.x >> 32);
    lane_words[t] = (uint32_t)(base + lane_cap - e.p);
  }
  __syncwarp();
  bad = __reduce_max_sync(0xffffffffu, bad);
  if (bad && j == 0) atomicMax(status, bad);
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

// Decoder state, replicated in the warp like the encoder's.  The next stream word is kept loaded one refill ahead.
struct Dec {
  uint64_t x;
  const uint32_t* p;
  const uint32_t* end;
  uint32_t ahead;                          // *p
  __device__ __forceinline__ void refill() {
    if (x < kRansLow) {
      x = (x << 32) | ahead;
      ++p;
      ahead = p < end ? __ldg(p) : 0u;
    }
  }
  __device__ __forceinline__ uint32_t get_bits() {                                      // rans_interface.cpp:88-104
    const uint32_t v = (uint32_t)(x & kBypassMax);
    x >>= kBypassBits;
    refill();
    return v;
  }
};

// Largest s in [0, size-2] with cdf[s] <= cf (std::find_if of :244-248), found by the whole warp: thread j probes
// one entry, a ballot counts the entries <= cf.  The first probe set is the 32 entries around the centre of the row,
// where a latent's mass sits (rows of <= 32 entries are covered entirely): most symbols cost ONE memory round trip,
// and start / next come out of the probed registers.  Otherwise 32-ary steps over what is left.
static constexpr uint32_t kAbove = 0x7fffffffu;
__device__ __forceinline__ int window_base(int size) { return size > 32 ? ((size - 2) >> 1) - 15 : 0; }
__device__ __forceinline__ int cdf_search_warp(const int32_t* __restrict__ cdf, int size, uint32_t cf, int j,
                                               uint32_t& start, uint32_t& freq) {
  const int w0 = window_base(size);
  // thread j probes entry w0+j and its successor: the thread that holds the answer hands over start and freq
  // (freq - 1 fits 16 bits) in ONE shuffle
  const uint32_t val = (w0 + j < size) ? (uint32_t)__ldg(cdf + w0 + j) : kAbove;
  const uint32_t nxt = (w0 + j + 1 < size) ? (uint32_t)__ldg(cdf + w0 + j + 1) : kAbove;
  const uint32_t pack = (val & 0xFFFFu) | ((nxt - val - 1u) << 16);
  int cnt = __popc(__ballot_sync(0xffffffffu, val <= cf));
  int lo, hi;
  if (cnt > 0 && cnt < 32) {
    const uint32_t got = __shfl_sync(0xffffffffu, pack, cnt - 1);
    start = got & 0xFFFFu;
    freq = (got >> 16) + 1u;
    return w0 + cnt - 1;
  }
  uint32_t v2;
  if (cnt == 0) { lo = 0; hi = w0; } else { lo = w0 + 31; hi = size - 1; }
  while (hi - lo > 1) {                    // invariant cdf[lo] <= cf < cdf[hi]
    const int step = (hi - lo + 31) >> 5;
    const int pidx = lo + (j + 1) * step;
    v2 = pidx < hi ? (uint32_t)__ldg(cdf + pidx) : kAbove;
    cnt = __popc(__ballot_sync(0xffffffffu, v2 <= cf));
    lo += cnt * step;
    hi = min(hi, lo + step);
  }
  start = (uint32_t)__ldg(cdf + lo);
  freq = (uint32_t)__ldg(cdf + lo + 1) - start;
  return lo;
}

// One WARP per string.  Chunks of 128 symbols: the threads fetch (row, size, offset) of four symbols each into the
// warp's slice of shared memory, the warp decodes them in order (state replicated, search by ballot), and the
// threads store four decoded values each.
static constexpr int kDecChunk = 128;
__global__ void __launch_bounds__(128) rans_decode_kernel(const uint32_t* __restrict__ packed, int64_t packed_words,
                                                          const uint32_t* __restrict__ lane_offsets,
                                                          const int32_t* __restrict__ indexes, int64_t n, int items,
                                                          int lanes, RansTables tb, int32_t* out, int* status,
                                                          int ck_log2, int ncp, const uint32_t* __restrict__ ckpt) {
  pdl_sync();
  __shared__ int4 slots[4][kDecChunk];                    // row, size, offset, decoded value
  // one warp per (string, segment): segment c of a string starts at symbol c * 2^ck_log2 from checkpoint c - 1
  // ({state low, state high, word position in `packed`}, lmsfc_rans_decode_indexed); without checkpoints nseg = 1
  const int nseg = ncp + 1;
  const int64_t wg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t t = wg / nseg;
  const int seg = (int)(wg - t * nseg);
  const int j = threadIdx.x & 31;
  if (t >= (int64_t)items * lanes) return;
  int4* const mine = slots[threadIdx.x >> 5];
  const int item = (int)(t / lanes), lane = (int)(t % lanes);
  const int64_t count_all = lane < n ? (n - lane + lanes - 1) / lanes : 0;
  const int64_t first = (int64_t)seg << ck_log2;
  if (seg > 0 && first >= count_all) return;
  const int64_t count = (ckpt && first + (1ll << ck_log2) < count_all) ? first + (1ll << ck_log2) : count_all;
  const bool has_next = count < count_all;                // the state this segment ends in is the next checkpoint
  const int32_t* idx = indexes + (int64_t)item * n;
  int32_t* dst = out + (int64_t)item * n;
  const uint32_t* w = packed + lane_offsets[t];
  Dec d;
  d.end = packed + packed_words;
  if (seg == 0) {
    if (w + 2 > d.end) { if (j == 0) atomicMax(status, 2); return; }
    d.x = (uint64_t)__ldg(w) | ((uint64_t)__ldg(w + 1) << 32);                          // rans64.h:104-112
    d.p = w + 2;
  } else {
    const uint32_t* ck = ckpt + ((size_t)t * ncp + (size_t)(seg - 1)) * 3;
    d.x = (uint64_t)__ldg(ck) | ((uint64_t)__ldg(ck + 1) << 32);
    const uint32_t pos = __ldg(ck + 2);
    if ((int64_t)pos > packed_words || d.x < kRansLow) { if (j == 0) atomicMax(status, 3); return; }
    d.p = packed + pos;
  }
  d.ahead = d.p < d.end ? __ldg(d.p) : 0u;
  int bad = 0;
  for (int64_t lo = first; lo < count; lo += kDecChunk) {
    const int m = count - lo < kDecChunk ? (int)(count - lo) : kDecChunk;
#pragma unroll
    for (int k = 0; k < kDecChunk / 32; ++k) {
      const int u = j + 32 * k;
      int4 sl = make_int4(-1, 0, 0, 0);
      if (u < m) {
        const int32_t r = __ldg(idx + lane + (lo + u) * lanes);
        if (r < 0 || r >= tb.rows) bad = 1;
        else {
          sl.y = __ldg(tb.sizes + r);
          sl.z = __ldg(tb.offsets + r);
          if (sl.y < 2 || sl.y > tb.cdf_stride) bad = 1; else sl.x = r;
        }
      }
      mine[u] = sl;
    }
    __syncwarp();
#pragma unroll 2
    for (int u = 0; u < m; ++u) {
      const int4 sl = mine[u];
      if (sl.x < 0) continue;
      const int32_t* cdf = tb.cdfs + (int64_t)sl.x * tb.cdf_stride;
      const int32_t max_value = sl.y - 2;
      const uint32_t cf = (uint32_t)(d.x & 0xFFFFu);
      uint32_t start, freq;
      const int s = cdf_search_warp(cdf, sl.y, cf, j, start, freq);
      d.x = (uint64_t)freq * (d.x >> kPrecision) + cf - start;                          // rans64.h:120-137
      d.refill();
      int32_t value = s;
      if (s == max_value) {                                                             // :254-275
        uint32_t v = d.get_bits();
        int nb = (int)v;
        while (v == kBypassMax && nb < 64) { v = d.get_bits(); nb += (int)v; }
        uint32_t raw = 0;
        for (int g = 0; g < nb; ++g) {
          const uint32_t dg = d.get_bits();
          if (g < 8) raw |= dg << (g * kBypassBits);
        }
        value = (int32_t)(raw >> 1);
        value = (raw & 1u) ? -value - 1 : value + max_value;
      }
      mine[u].w = value + sl.z;                          // every thread holds the same value
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kDecChunk / 32; ++k) {
      const int u = j + 32 * k;
      if (u < m) dst[lane + (lo + u) * lanes] = mine[u].w;
    }
    __syncwarp();
  }
  if (has_next) {                                         // integrity: a wrong side table cannot pass silently
    const uint32_t* ck = ckpt + ((size_t)t * ncp + (size_t)seg) * 3;
    const uint64_t want = (uint64_t)__ldg(ck) | ((uint64_t)__ldg(ck + 1) << 32);
    if (d.x != want || (uint32_t)(d.p - packed) != __ldg(ck + 2)) bad = bad > 3 ? bad : 3;
  }
  bad = __reduce_max_sync(0xffffffffu, bad);
  if (bad && j == 0) atomicMax(status, bad);
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

static int check_ckpt(int interval, int64_t n, int lanes, int ncp, const char* who) {
  LMSFC_CHECK_ARG(interval >= 128 && (interval & (interval - 1)) == 0, "%s: checkpoint interval %d is not a power of two >= 128",
                  who, interval);
  LMSFC_CHECK_ARG(ncp == lmsfc_rans_checkpoints(n, lanes < 1 ? 1 : lanes, interval), "%s: ncp=%d, lmsfc_rans_checkpoints() says %d", who, ncp,
                  lmsfc_rans_checkpoints(n, lanes < 1 ? 1 : lanes, interval));
  return LMSFC_OK;
}
static int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

extern "C" int lmsfc_rans_checkpoints(int64_t n, int lanes, int interval) {
  if (lanes < 1 || interval < 1 || n <= 0) return 0;
  const int64_t count = (n + lanes - 1) / lanes;          // the longest lane
  return (int)((count + interval - 1) / interval - 1);
}

extern "C" int lmsfc_rans_encode_indexed(const int32_t* symbols, const int32_t* indexes, int items, int64_t n, int lanes,
                                         const int32_t* cdfs, int rows, int cdf_stride, const int32_t* sizes,
                                         const int32_t* offsets, uint32_t* scratch, int64_t lane_cap, uint32_t* lane_words,
                                         uint32_t* lane_offsets, uint32_t* packed, int64_t packed_cap, int* status,
                                         int ckpt_interval, int ncp, uint32_t* ckpt, void* stream) {
  RansTables tb{cdfs, sizes, offsets, rows, cdf_stride};
  if (int rc = check_tables(tb, "rans_encode")) return rc;
  LMSFC_CHECK_ARG(items >= 0 && n >= 0 && lanes >= 1, "rans_encode: bad sizes");
  LMSFC_CHECK_ARG(scratch && lane_words && lane_offsets && packed && status, "rans_encode: null buffer");
  LMSFC_CHECK_ARG(n == 0 || (symbols && indexes), "rans_encode: null symbols / indexes");
  LMSFC_CHECK_ARG(lane_cap >= lmsfc_rans_lane_cap(n, lanes), "rans_encode: lane_cap below lmsfc_rans_lane_cap()");
  int lg = 0;
  if (ckpt_interval) {
    if (int rc = check_ckpt(ckpt_interval, n, lanes, ncp, "rans_encode_indexed")) return rc;
    LMSFC_CHECK_ARG(ncp == 0 || ckpt, "rans_encode_indexed: null checkpoint table");
    lg = ilog2(ckpt_interval);
  } else {
    ncp = 0; ckpt = nullptr;
  }
  const int64_t T = (int64_t)items * lanes;
  LMSFC_CHECK_ARG(T * lane_cap < (1ll << 32), "rans_encode: more than 2^32 words");
  if (T == 0) return LMSFC_OK;
  LMSFC_LAUNCH(rans_encode_kernel, dim3(ceil_div(T * 32, 128)), dim3(128), 0, ST(stream), symbols, indexes, n, items, lanes,
               tb, scratch, lane_cap, lane_words, status, lg, ncp, ncp ? ckpt : (uint32_t*)nullptr);
  LMSFC_LAUNCH(rans_pack_kernel, dim3(ceil_div(T * 32, 256)), dim3(256), 0, ST(stream), (const uint32_t*)scratch,
               lane_cap, (const uint32_t*)lane_words, T, packed, packed_cap, lane_offsets, status);
  return LMSFC_OK;
}

extern "C" int lmsfc_rans_decode_indexed(const uint32_t* packed, int64_t packed_words, const uint32_t* lane_offsets,
                                         const int32_t* indexes, int items, int64_t n, int lanes, const int32_t* cdfs,
                                         int rows, int cdf_stride, const int32_t* sizes, const int32_t* offsets,
                                         int32_t* symbols_out, int* status, int ckpt_interval, int ncp,
                                         const uint32_t* ckpt, void* stream) {
  RansTables tb{cdfs, sizes, offsets, rows, cdf_stride};
  if (int rc = check_tables(tb, "rans_decode")) return rc;
  LMSFC_CHECK_ARG(items >= 0 && n >= 0 && lanes >= 1, "rans_decode: bad sizes");
  int lg = 0;
  if (ckpt_interval) {
    if (int rc = check_ckpt(ckpt_interval, n, lanes, ncp, "rans_decode_indexed")) return rc;
    LMSFC_CHECK_ARG(ncp == 0 || ckpt, "rans_decode_indexed: null checkpoint table");
    lg = ilog2(ckpt_interval);
  } else {
    ncp = 0; ckpt = nullptr;
  }
  const int64_t T = (int64_t)items * lanes;
  if (T == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG(packed && lane_offsets && status && packed_words >= 2 * T, "rans_decode: string shorter than its flush words");
  LMSFC_CHECK_ARG(n == 0 || (indexes && symbols_out), "rans_decode: null indexes / output");
  LMSFC_LAUNCH(rans_decode_kernel, dim3(ceil_div(T * (ncp + 1) * 32, 128)), dim3(128), 0, ST(stream), packed, packed_words,
               lane_offsets, indexes, n, items, lanes, tb, symbols_out, status, lg, ncp, ncp ? ckpt : (const uint32_t*)nullptr);
  return LMSFC_OK;
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
  LMSFC_LAUNCH(rans_encode_kernel, dim3(ceil_div(T * 32, 128)), dim3(128), 0, ST(stream), symbols, indexes, n, items, lanes,
               tb, scratch, lane_cap, lane_words, status, 0, 0, (uint32_t*)nullptr);
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
  LMSFC_LAUNCH(rans_decode_kernel, dim3(ceil_div(T * 32, 128)), dim3(128), 0, ST(stream), packed, packed_words, lane_offsets,
               indexes, n, items, lanes, tb, symbols_out, status, 0, 0, (const uint32_t*)nullptr);
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
