/* Oracle (TEST INFRASTRUCTURE, see oracle/__init__.py): plain-C restatement of
 * the reference's integer pmf -> quantised CDF routine,
 * /root/reference/compressai/compressai/cpp_exts/ops/ops.cpp:40-109.
 * Parity: pinned by the reference KAT tests/test_ops.py:104-118 and by
 * tests/golden/cdf_*.npz generated from the built reference.
 *
 * Return: 0 ok, 1 negative/non-finite pmf entry (ops.cpp:46-52),
 *         2 all-zero pmf (ops.cpp:61-64).  cdf must hold n+1 uint32.
 */
#include <math.h>
#include <stdint.h>

int lmsfc_oracle_pmf_to_quantized_cdf(const float *pmf, int n, int precision,
                                      uint32_t *cdf) {
  for (int i = 0; i < n; ++i)
    if (pmf[i] < 0 || !isfinite(pmf[i])) return 1;
  cdf[0] = 0;
  /* ops.cpp:57-58: float * int -> float, std::round(float), -> uint32 */
  for (int i = 0; i < n; ++i)
    cdf[i + 1] = (uint32_t)roundf(pmf[i] * (float)(1 << precision));
  /* ops.cpp:60: std::accumulate with an int init -> int accumulator */
  int acc = 0;
  for (int i = 0; i <= n; ++i) acc = (int)((uint32_t)acc + cdf[i]);
  const uint32_t total = (uint32_t)acc;
  if (total == 0) return 2;
  /* ops.cpp:66-69 */
  for (int i = 0; i <= n; ++i)
    cdf[i] = (uint32_t)((((uint64_t)(1 << precision)) * cdf[i]) / total);
  /* ops.cpp:71-72 */
  for (int i = 1; i <= n; ++i) cdf[i] += cdf[i - 1];
  cdf[n] = 1u << precision;
  /* ops.cpp:74-100: steal from the smallest freq>1 symbol */
  for (int i = 0; i < n; ++i) {
    if (cdf[i] == cdf[i + 1]) {
      uint32_t best_freq = ~0u;
      int best_steal = -1;
      for (int j = 0; j < n; ++j) {
        uint32_t freq = cdf[j + 1] - cdf[j];
        if (freq > 1 && freq < best_freq) {
          best_freq = freq;
          best_steal = j;
        }
      }
      if (best_steal < 0) return 3; /* reference asserts */
      if (best_steal < i) {
        for (int j = best_steal + 1; j <= i; ++j) cdf[j]--;
      } else {
        for (int j = i + 1; j <= best_steal; ++j) cdf[j]++;
      }
    }
  }
  return 0;
}
