/* Oracle (TEST INFRASTRUCTURE, see oracle/__init__.py): plain-C restatement of
 * the reference's range-ANS coder for one string,
 *   /root/reference/compressai/compressai/cpp_exts/rans/rans_interface.cpp
 *     :108-178 (BufferedRansEncoder::encode_with_indexes: symbol -> (start,range) records,
 *               out-of-table values through the max_value sentinel + 4-bit bypass digits),
 *     :180-207 (flush: records coded last-to-first, 32-bit words written back-to-front),
 *     :222-288 (RansDecoder::decode_with_indexes),
 *     :68-104  (Rans64EncPutBits / Rans64DecGetBits),
 *   /root/reference/compressai/third_party/ryg_rans/rans64.h (Rans64EncPut :75-90,
 *     Rans64EncFlush :93-100, Rans64DecInit :104-112, Rans64DecGet/Advance :115-137).
 * Parity: pinned by tests/golden/rans.npz (strings produced by the reference's own compiled
 * `ans` module, tests/golden/make_golden_rans.py) and, in the authoring container, against
 * oracle/_ref/ans*.so directly.
 *
 * A "lane" view (first, step, count) selects symbols first, first+step, ...: with step 1 this
 * is the reference call on the whole item; the product's interleaved container codes lane k of
 * K as the reference string of symbols[k::K].
 */
#include <stdint.h>
#include <stdlib.h>

#define PRECISION 16u
#define BYPASS_BITS 4u
#define BYPASS_MAX 15u
#define RANS_LOW (1ull << 31)

typedef struct { uint64_t x; uint32_t *p; } enc_t;

static void put(enc_t *e, uint32_t start, uint32_t freq, uint32_t bits) {
  /* rans64.h:75-90 (and rans_interface.cpp:68-86 with freq = 1, start = val) */
  uint64_t x = e->x;
  const uint64_t lim = ((RANS_LOW >> bits) << 32) * freq;
  if (x >= lim) { *--e->p = (uint32_t)x; x >>= 32; }
  e->x = ((x / freq) << bits) + (x % freq) + start;
}

/* how many records one symbol expands to, and the raw value, rans_interface.cpp:128-175 */
static int classify(int32_t sym, int32_t offset, int32_t max_value, int32_t *value, uint32_t *raw) {
  int32_t v = sym - offset;
  *raw = 0;
  if (v < 0) { *raw = (uint32_t)(-2 * v - 1); v = max_value; }
  else if (v >= max_value) { *raw = (uint32_t)(2 * (v - max_value)); v = max_value; }
  *value = v;
  return v == max_value;
}

/* Encodes lane (first, step, count).  `words` is a scratch of `cap` uint32; the string is the
 * tail words[cap - n .. cap) and n (words) is returned; -1 when cap is too small. */
long lmsfc_oracle_rans_encode(const int32_t *symbols, const int32_t *indexes, long first, long step,
                              long count, const int32_t *cdfs, int cdf_stride, const int32_t *sizes,
                              const int32_t *offsets, uint32_t *words, long cap) {
  enc_t e = {RANS_LOW, words + cap};
  for (long t = count - 1; t >= 0; --t) {           /* flush pops records from the back :189-199 */
    const long i = first + t * step;
    const int32_t row = indexes[i];
    const int32_t *cdf = cdfs + (long)row * cdf_stride;
    const int32_t max_value = sizes[row] - 2;
    int32_t v; uint32_t raw;
    if ((e.p - words) < 24) return -1;
    if (classify(symbols[i], offsets[row], max_value, &v, &raw)) {
      int n = 0;
      /* :150-153.  The reference shifts a uint32 by up to 32 (undefined for raw >= 2^28, i.e. values
       * ~2^27 outside the table); the 64-bit shift here defines that corner: n <= 8. */
      while (((uint64_t)raw >> (n * BYPASS_BITS)) != 0) ++n;
      for (int j = n - 1; j >= 0; --j)               /* raw digits were pushed low-to-high :166-172 */
        put(&e, (raw >> (j * BYPASS_BITS)) & BYPASS_MAX, 1, BYPASS_BITS);
      /* count prefix :156-163: q records of 15 then the remainder; reversed here */
      put(&e, (uint32_t)(n % (int)BYPASS_MAX), 1, BYPASS_BITS);
      for (int q = n / (int)BYPASS_MAX; q > 0; --q) put(&e, BYPASS_MAX, 1, BYPASS_BITS);
    }
    put(&e, (uint32_t)cdf[v], (uint32_t)(cdf[v + 1] - cdf[v]), PRECISION);
  }
  e.p -= 2;                                          /* rans64.h:93-100 */
  e.p[0] = (uint32_t)e.x;
  e.p[1] = (uint32_t)(e.x >> 32);
  return (long)(words + cap - e.p);
}

typedef struct { uint64_t x; const uint32_t *p; } dec_t;

static uint32_t get_bits(dec_t *d) {                 /* rans_interface.cpp:88-104 */
  const uint32_t val = (uint32_t)(d->x & BYPASS_MAX);
  d->x >>= BYPASS_BITS;
  if (d->x < RANS_LOW) d->x = (d->x << 32) | *d->p++;
  return val;
}

/* Decodes lane (first, step, count) of `indexes` from `words` into out[first + t*step]. */
void lmsfc_oracle_rans_decode(const uint32_t *words, const int32_t *indexes, long first, long step,
                              long count, const int32_t *cdfs, int cdf_stride, const int32_t *sizes,
                              const int32_t *offsets, int32_t *out) {
  dec_t d;
  d.x = (uint64_t)words[0] | ((uint64_t)words[1] << 32);   /* rans64.h:104-112 */
  d.p = words + 2;
  for (long t = 0; t < count; ++t) {
    const long i = first + t * step;
    const int32_t row = indexes[i];
    const int32_t *cdf = cdfs + (long)row * cdf_stride;
    const int32_t max_value = sizes[row] - 2;
    const uint32_t cf = (uint32_t)(d.x & 0xFFFFu);
    int32_t s = 0;                                   /* first entry > cf, minus one :244-248 */
    while (s + 1 < sizes[row] && (uint32_t)cdf[s + 1] <= cf) ++s;
    const uint32_t start = (uint32_t)cdf[s], freq = (uint32_t)(cdf[s + 1] - cdf[s]);
    d.x = freq * (d.x >> PRECISION) + (d.x & 0xFFFFu) - start;   /* rans64.h:120-137 */
    if (d.x < RANS_LOW) d.x = (d.x << 32) | *d.p++;
    int32_t value = s;
    if (value == max_value) {                        /* :254-275 */
      uint32_t val = get_bits(&d);
      int32_t n = (int32_t)val;
      while (val == BYPASS_MAX) { val = get_bits(&d); n += (int32_t)val; }
      uint32_t raw = 0;
      for (int j = 0; j < n; ++j) raw |= get_bits(&d) << (j * BYPASS_BITS);
      value = (int32_t)(raw >> 1);
      value = (raw & 1u) ? -value - 1 : value + max_value;
    }
    out[i] = value + offsets[row];
  }
}
