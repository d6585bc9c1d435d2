// rand.cu -- counter-based synthetic data generator (see include/sysml_b200.h, "Synthetic data").
// cell i of the global array = lo + (hi-lo) * u,  u = (splitmix64(splitmix64(seed) ^ (offset+i)) >> 11) * 2^-53.
// Every step is exact integer arithmetic or a single correctly rounded FP64 operation WITHOUT
// contraction, so the CPU twin (oracle/sysml_oracle.cpp: orc_rand_uniform) is bit-identical.
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace {
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

__global__ void rand_uniform_kernel(double* __restrict__ out, i64 len, uint64_t s, i64 offset, double lo, double span) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) {
    const uint64_t h = splitmix64(s ^ (uint64_t)(offset + i));
    const double u = __dmul_rn((double)(h >> 11), 0x1.0p-53);
    out[i] = __dadd_rn(lo, __dmul_rn(span, u));
  }
}

// ---- synthetic CSR (BASELINE config C3: every cell non-zero with a small probability p, so the number of entries of a
// row is Binomial(cols, p) ~ Poisson(cols*p)).  Row r of the GLOBAL matrix depends on (seed, r) only:
//   hr    = splitmix64(s ^ r)
//   count = smallest k with PoissonCDF(k; mean) > u0, u0 from splitmix64(hr), capped at min(cols, 4096);
//           the CDF is accumulated as p_0 = exp(-mean) (computed by the caller, passed in), p_k = p_{k-1} * mean / k
//           with separately rounded operations, so the CPU twin (oracle: orc_rand_csr_*) gets the same counts
//   entry j (0 <= j < count): column = j*W + (splitmix64(hr + 2j + 1) mod W), W = cols / count  (sorted, distinct),
//                             value  = -1 + 2*u, u from splitmix64(hr + 2j + 2), never exactly 0
__host__ __device__ __forceinline__ int csr_row_count(uint64_t hr, double mean, double p0, i64 cols) {
  const uint64_t h = splitmix64(hr);
  const double u0 = (double)(h >> 11) * 0x1.0p-53;
  double pk = p0, cdf = p0;
  int k = 0;
  const int cap = (int)(cols < 4096 ? cols : 4096);
  while (cdf <= u0 && k < cap) {
    k++;
#ifdef __CUDA_ARCH__
    pk = __ddiv_rn(__dmul_rn(pk, mean), (double)k);
    cdf = __dadd_rn(cdf, pk);
#else
    pk = (pk * mean) / (double)k;
    cdf = cdf + pk;
#endif
    if (pk < 1e-300) break;
  }
  return k;
}

__global__ void rand_csr_count_kernel(int* __restrict__ cnt, i64 rows, i64 row_offset, uint64_t s, double mean, double p0, i64 cols) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r <= rows; r += stride)
    cnt[r] = (r < rows) ? csr_row_count(splitmix64(s ^ (uint64_t)(row_offset + r)), mean, p0, cols) : 0;
}

// one thread per row (rows are short); consecutive threads write consecutive rows' entries
__global__ void rand_csr_fill_kernel(const int* __restrict__ rp, int* __restrict__ ci, double* __restrict__ va, i64 rows, i64 row_offset,
                                     uint64_t s, i64 cols) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    const uint64_t hr = splitmix64(s ^ (uint64_t)(row_offset + r));
    const int lo = rp[r], c = rp[r + 1] - lo;
    if (c <= 0) continue;
    const uint64_t W = (uint64_t)(cols / c);
    for (int j = 0; j < c; j++) {
      const uint64_t hc = splitmix64(hr + 2ull * (uint64_t)j + 1ull), hv = splitmix64(hr + 2ull * (uint64_t)j + 2ull);
      ci[lo + j] = (int)((uint64_t)j * W + hc % W);
      const double u = __dmul_rn((double)((hv >> 11) | 1ull), 0x1.0p-53);  // odd numerator: never 0, so never a stored zero
      va[lo + j] = __dadd_rn(-1.0, __dmul_rn(2.0, u));
    }
  }
}
}  // namespace

extern "C" int sysml_b200_rand_csr(sysml_b200_ctx* ctx, int64_t rows, int64_t cols, double mean_per_row, uint64_t seed, int64_t row_offset,
                                   int32_t* rowptr, int32_t* colind, double* val, int64_t* h_nnz) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 1 || !(mean_per_row >= 0.0) || mean_per_row > 700.0 || !rowptr)
    return sysml_set_error(ctx, SYSML_E_INVALID, "rand_csr: bad arguments (rows >= 0, cols >= 1, 0 <= mean <= 700, rowptr non-null)");
  if (rows > 0x7ffffffeLL) return sysml_set_error(ctx, SYSML_E_LIMIT, "rand_csr: more than 2^31-2 rows");
  const uint64_t s = splitmix64(seed ^ 0xC5A1ull);
  const double p0 = exp(-mean_per_row);
  i64 blocks = (rows + 256) / 256;
  const i64 maxb = (i64)ctx->sm_count * 16;
  if (blocks > maxb) blocks = maxb;
  if (!colind) {
    // phase 1: rowptr (exclusive scan of the per-row counts) and the total
    size_t scan_bytes = 0;
    SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const int*)nullptr, (int*)nullptr, (int)(rows + 1), ctx->stream));
    void* ws = nullptr;
    const size_t cnt_bytes = (((size_t)(rows + 1) * 4) + 255) & ~(size_t)255;
    int rc = sysml_workspace(ctx, cnt_bytes + scan_bytes + 256, &ws);
    if (rc) return rc;
    int* cnt = (int*)ws;
    rand_csr_count_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(cnt, rows, row_offset, s, mean_per_row, p0, cols);
    SYSML_LAUNCH_CHECK(ctx, "rand_csr_count_kernel");
    // the scan is 32-bit: totals beyond 2^31-1 wrap to a negative last element and are refused below
    SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum((char*)ws + cnt_bytes, scan_bytes, cnt, rowptr, (int)(rows + 1), ctx->stream));
    ctx->launches++;
    int last = 0;
    SYSML_CUDA(ctx, cudaMemcpyAsync(&last, rowptr + rows, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (last < 0 || (double)rows * mean_per_row > 2.1e9)
      return sysml_set_error(ctx, SYSML_E_LIMIT, "rand_csr: more than 2^31-1 entries (int32 row pointers, CSRPointer.java:76-100)");
    if (h_nnz) *h_nnz = last;
    return SYSML_OK;
  }
  if (!val) return sysml_set_error(ctx, SYSML_E_INVALID, "rand_csr: val is null");
  if (rows == 0) return SYSML_OK;
  rand_csr_fill_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(rowptr, colind, val, rows, row_offset, s, cols);
  SYSML_LAUNCH_CHECK(ctx, "rand_csr_fill_kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_rand_uniform(sysml_b200_ctx* ctx, double* out, int64_t len, uint64_t seed, int64_t offset,
                                       double lo, double hi) {
  SYSML_ENTER(ctx);
  if (len < 0 || (len > 0 && !out)) return sysml_set_error(ctx, SYSML_E_INVALID, "rand_uniform: bad arguments");
  if (len == 0) return SYSML_OK;
  i64 blocks = (len + 255) / 256;
  const i64 maxb = (i64)ctx->sm_count * 16;
  if (blocks > maxb) blocks = maxb;
  rand_uniform_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(out, len, splitmix64(seed), offset, lo, hi - lo);
  SYSML_LAUNCH_CHECK(ctx, "rand_uniform_kernel");
  return SYSML_OK;
}
