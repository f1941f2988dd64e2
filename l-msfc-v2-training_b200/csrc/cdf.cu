// Integer pmf -> quantised CDF tables, bit-exact with the reference's C++ routine
//   compressai/cpp_exts/ops/ops.cpp:40-109   (pmf_to_quantized_cdf)
//   compressai/entropy_models/entropy_models.py:204-212 (_pmf_to_cdf: one row per channel/scale)
// One thread block per row: the argmin scan and the range shift of the "steal" loop run
// block-parallel; the loop over zero-width bins keeps the reference's order, so the result is
// identical.  The single-row entry point with HOST pointers (the signature of the reference's
// pybind11 function) stages the row through the same device kernel: there is no CPU algorithm in
// this library.
#include <math.h>

#include "kernels.cuh"

namespace lmsfc {

// status[row]: 0 ok, 1 bad element, 2 all-zero, 3 nothing to steal
__global__ void __launch_bounds__(256) cdf_rows_kernel(const float* pmf, int pmf_stride, const float* tail,
                                                       const int* pmf_length, int precision, int32_t* cdf_out,
                                                       int cdf_stride, int* status) {
  pdl_sync();
  extern __shared__ uint32_t sm[];       // n + 1 entries
  __shared__ unsigned long long best_key;
  __shared__ int flag;
  __shared__ uint32_t wsum[8];
  const int row = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
  const int len = pmf_length[row];
  const int n = len + 1;                 // clipped pmf ++ tail mass (entropy_models.py:208-211)
  uint32_t* c = sm;
  int32_t* out = cdf_out + (size_t)row * cdf_stride;
  for (int i = tid; i < cdf_stride; i += nt) out[i] = 0;
  if (tid == 0) flag = 0;
  __syncthreads();
  const float scale = (float)(1 << precision);
  for (int i = tid; i < n; i += nt) {
    const float p = (i < len) ? pmf[(size_t)row * pmf_stride + i] : tail[row];
    if (p < 0.f || !isfinite(p)) flag = 1;
    c[i + 1] = (uint32_t)roundf(p * scale);
  }
  if (tid == 0) c[0] = 0;
  __syncthreads();
  if (flag) { if (tid == 0) status[row] = 1; return; }
  // total (mod 2^32, like the int accumulator)
  uint32_t part = 0;
  for (int i = tid; i <= n; i += nt) part += c[i];
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if ((tid & 31) == 0) wsum[tid >> 5] = part;
  __syncthreads();
  uint32_t total = 0;
  for (int w = 0; w < (nt + 31) / 32; ++w) total += wsum[w];
  if (total == 0) { if (tid == 0) status[row] = 2; return; }
  for (int i = tid; i <= n; i += nt) c[i] = (uint32_t)((((uint64_t)1 << precision) * c[i]) / total);
  __syncthreads();
  if (tid == 0) {                         // n <= ~3.1k: a serial prefix sum is negligible
    for (int i = 1; i <= n; ++i) c[i] += c[i - 1];
    c[n] = 1u << precision;
  }
  __syncthreads();
  for (int i = 0; i < n; ++i) {
    if (c[i] != c[i + 1]) continue;       // uniform branch (shared data, synced below)
    if (tid == 0) best_key = ~0ull;
    __syncthreads();
    unsigned long long key = ~0ull;      // (freq << 32 | j): smallest freq, first index on ties
    for (int j = tid; j < n; j += nt) {
      const uint32_t f = c[j + 1] - c[j];
      if (f > 1) {
        const unsigned long long k = ((unsigned long long)f << 32) | (uint32_t)j;
        if (k < key) key = k;
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
      if (other < key) key = other;
    }
    if ((tid & 31) == 0 && key != ~0ull) atomicMin(&best_key, key);
    __syncthreads();
    const unsigned long long bk = best_key;
    if (bk == ~0ull) { if (tid == 0) status[row] = 3; return; }
    const int best = (int)(uint32_t)bk;
    if (best < i) { for (int j = best + 1 + tid; j <= i; j += nt) c[j]--; }
    else          { for (int j = i + 1 + tid; j <= best; j += nt) c[j]++; }
    __syncthreads();
  }
  for (int i = tid; i <= n; i += nt) out[i] = (int32_t)c[i];
  if (tid == 0) status[row] = 0;
}

int pmf_rows_to_cdf_device(const float* pmf, int pmf_stride, const float* tail, const int* pmf_length, int rows,
                           int precision, int32_t* cdf, int cdf_stride, int* status, cudaStream_t s) {
  if (rows == 0) return LMSFC_OK;
  LMSFC_CHECK_ARG(cdf_stride >= 2 && precision > 0 && precision <= 30, "cdf: bad stride/precision");
  const size_t smem = (size_t)(cdf_stride + 1) * sizeof(uint32_t);
  LMSFC_CHECK_ARG(smem <= 200 * 1024, "cdf: row of %d entries does not fit shared memory", cdf_stride);
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) {
    LMSFC_CUDA(cudaFuncSetAttribute(cdf_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = smem;
  }
  LMSFC_LAUNCH(cdf_rows_kernel, dim3(rows), dim3(256), smem, s, pmf, pmf_stride, tail, pmf_length, precision, cdf, cdf_stride, status);
  return LMSFC_OK;
}

// Host-pointer signature of the reference's `_CXX.pmf_to_quantized_cdf(list, precision)`: the row is
// copied to the device, converted by cdf_rows_kernel and copied back (synchronous, like the original).
int pmf_to_quantized_cdf_host(const float* pmf, int n, int precision, uint32_t* cdf) {
  if (n <= 0) {
    set_error("Invalid `pmf`: at least one element must have a non-zero probability.");
    return LMSFC_ERR_DOMAIN;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("pmf_to_quantized_cdf: no CUDA device (this library has no CPU fallback)");
    return LMSFC_ERR_CUDA;
  }
  const int len = n - 1;                       // kernel row = pmf[0:len] ++ tail
  uint8_t* buf = nullptr;
  const size_t off_cdf = (size_t)n * sizeof(float), off_len = off_cdf + (size_t)(n + 1) * sizeof(int32_t);
  LMSFC_CUDA(cudaMalloc(&buf, off_len + 2 * sizeof(int)));
  float* d_pmf = (float*)buf;
  int32_t* d_cdf = (int32_t*)(buf + off_cdf);
  int* d_len = (int*)(buf + off_len);
  int host_meta[2] = {len, 0};
  int rc = LMSFC_OK;
  cudaError_t e = cudaMemcpy(d_pmf, pmf, (size_t)n * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_len, host_meta, sizeof(host_meta), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    rc = pmf_rows_to_cdf_device(d_pmf, n, d_pmf + len, d_len, 1, precision, d_cdf, n + 1, d_len + 1, 0);
    if (rc == LMSFC_OK) {
      e = cudaMemcpy(cdf, d_cdf, (size_t)(n + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost);
      if (e == cudaSuccess) e = cudaMemcpy(host_meta, d_len, sizeof(host_meta), cudaMemcpyDeviceToHost);
    }
  }
  cudaFree(buf);
  if (rc) return rc;
  if (e != cudaSuccess) {
    set_error("pmf_to_quantized_cdf: %s", cudaGetErrorString(e));
    return LMSFC_ERR_CUDA;
  }
  switch (host_meta[1]) {
    case 0: return LMSFC_OK;
    case 1: set_error("Invalid `pmf`, non-finite or negative element found"); break;
    case 2: set_error("Invalid `pmf`: at least one element must have a non-zero probability."); break;
    default: set_error("pmf_to_quantized_cdf: no symbol to steal from"); break;
  }
  return LMSFC_ERR_DOMAIN;
}

}  // namespace lmsfc
