// sparse_out.cu -- operations whose RESULT is CSR: transpose, C = alpha*A + beta*B, and sparse x sparse matmult.
//
// Reference: CSRPointer.transpose (csr2csc, CSRPointer.java:463-481), LibMatrixCUDA.dgeam's sparse branch
// (cusparseDcsrgeam, LibMatrixCUDA.java:1593-1701 via CSRPointer.allocateForDgeam :227-299) and
// LibMatrixCuMatMult.sparseSparseMatMult (cusparseDcsrgemm, LibMatrixCuMatMult.java:150-195 via
// CSRPointer.allocateForMatrixMultiply :301-399: a symbolic nnz pass, then allocation, then the numeric pass).
// CPU twins: LibMatrixMult.matrixMultSparseSparse (LibMatrixMult.java:1397-1500: per output row, contributions
// added in ascending k), LibMatrixBincell's sparse-sparse merge (LibMatrixBincell.java:250-339) and
// LibMatrixReorg.transpose.  Like the CPU sparse blocks -- and unlike cuSPARSE -- results hold no explicit zeros
// (SparseRow.append skips them), so nnz equals MatrixBlock.recomputeNonZeros of the dense result.
//
// Results whose size is only known after a symbolic pass are allocated by the library (cudaMalloc; the caller
// releases them with sysml_b200_free) -- the same two-step protocol as CSRPointer.allocateFor*.
//
// sparse x sparse is expand - sort - compress: every product a(i,k)*b(k,j) is written with the 64-bit key
// (i << 32 | j) in (i, then k ascending) order, a STABLE radix sort (cub::DeviceRadixSort) groups equal keys
// without reordering them, and one thread per distinct key adds its run left to right -- the reference's
// summation order, so the result is bit-identical to the CPU algorithm (products and sums are separately rounded).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace {

typedef unsigned long long u64;

inline unsigned grid_for(sysml_b200_ctx* ctx, i64 items, int threads) {
  i64 b = (items + threads - 1) / threads;
  const i64 cap = (i64)ctx->sm_count * 32;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

int fetch_int(sysml_b200_ctx* ctx, const int* d, int* h) {
  SYSML_CUDA(ctx, cudaMemcpyAsync(h, d, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SYSML_OK;
}
int fetch_u64(sysml_b200_ctx* ctx, const u64* d, u64* h) {
  SYSML_CUDA(ctx, cudaMemcpyAsync(h, d, sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SYSML_OK;
}

// carve a 256-byte aligned piece out of a workspace block
template <class T>
T* carve(char*& p, size_t count) {
  T* r = reinterpret_cast<T*>(p);
  p += ((count * sizeof(T) + 255) / 256) * 256;
  return r;
}
inline size_t padded(size_t bytes) { return ((bytes + 255) / 256) * 256; }

// ---------------- transpose ----------------
__global__ void tr_keys_kernel(const int* __restrict__ rp, const int* __restrict__ ci, i64 rows, u64* __restrict__ keys,
                               int* __restrict__ colcnt) {
  // one warp per row: key = (col << 32) | row; column histogram for the new row pointers
  const int lane = threadIdx.x & 31;
  const i64 w = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = w; r < rows; r += nw) {
    const int lo = rp[r], hi = rp[r + 1];
    for (int k = lo + lane; k < hi; k += 32) {
      const int c = ci[k];
      keys[k] = ((u64)(unsigned)c << 32) | (u64)(unsigned)r;
      atomicAdd(colcnt + c, 1);
    }
  }
}
__global__ void low32_kernel(const u64* __restrict__ keys, i64 n, int* __restrict__ out) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (int)(keys[i] & 0xffffffffull);
}

// ---------------- C = alpha*A + beta*B (sorted merge per row) ----------------
template <bool FILL>
__global__ void geam_kernel(const int* __restrict__ arp, const int* __restrict__ aci, const double* __restrict__ ava,
                            const int* __restrict__ brp, const int* __restrict__ bci, const double* __restrict__ bva, double alpha,
                            double beta, i64 rows, int* __restrict__ cnt_or_rp, int* __restrict__ cci, double* __restrict__ cva) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    int i = arp[r], j = brp[r];
    const int ie = arp[r + 1], je = brp[r + 1];
    int n = 0;
    const int base = FILL ? cnt_or_rp[r] : 0;
    while (i < ie || j < je) {
      const int ca = (i < ie) ? aci[i] : 0x7fffffff, cb = (j < je) ? bci[j] : 0x7fffffff;
      double v;
      int c;
      if (ca == cb) {
        v = __dadd_rn(__dmul_rn(alpha, ava[i]), __dmul_rn(beta, bva[j]));
        c = ca;
        i++;
        j++;
      } else if (ca < cb) {
        v = __dmul_rn(alpha, ava[i]);
        c = ca;
        i++;
      } else {
        v = __dmul_rn(beta, bva[j]);
        c = cb;
        j++;
      }
      if (v != 0.0) {  // sparse rows hold no explicit zeros (SparseRow.append)
        if (FILL) {
          cci[base + n] = c;
          cva[base + n] = v;
        }
        n++;
      }
    }
    if (!FILL) cnt_or_rp[r] = n;
  }
}

// ---------------- sparse x sparse: expand / compress ----------------
__global__ void spgemm_count_kernel(const int* __restrict__ arp, const int* __restrict__ aci, const int* __restrict__ brp, i64 rows,
                                    u64* __restrict__ cnt) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    u64 n = 0;
    for (int k = arp[r]; k < arp[r + 1]; k++) {
      const int kk = aci[k];
      n += (u64)(brp[kk + 1] - brp[kk]);
    }
    cnt[r] = n;
  }
}
__global__ void spgemm_expand_kernel(const int* __restrict__ arp, const int* __restrict__ aci, const double* __restrict__ ava,
                                     const int* __restrict__ brp, const int* __restrict__ bci, const double* __restrict__ bva, i64 rows,
                                     const u64* __restrict__ off, u64* __restrict__ keys, double* __restrict__ vals) {
  // one warp per row of A; its products land at off[r].. in (k ascending, then j ascending) order
  const int lane = threadIdx.x & 31;
  const i64 w = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = w; r < rows; r += nw) {
    u64 pos = off[r];
    for (int k = arp[r]; k < arp[r + 1]; k++) {
      const int kk = aci[k];
      const double a = ava[k];
      const int lo = brp[kk], hi = brp[kk + 1];
      for (int j = lo + lane; j < hi; j += 32) {
        keys[pos + (u64)(j - lo)] = ((u64)r << 32) | (u64)(unsigned)bci[j];
        vals[pos + (u64)(j - lo)] = __dmul_rn(a, bva[j]);
      }
      pos += (u64)(hi - lo);
    }
  }
}
// flag[i] = 1 when position i starts a run of equal keys whose left-to-right sum is non-zero; sums[i] = that sum
__global__ void spgemm_runs_kernel(const u64* __restrict__ keys, const double* __restrict__ vals, u64 n, int* __restrict__ flag,
                                   double* __restrict__ sums) {
  const u64 stride = (u64)gridDim.x * blockDim.x;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const u64 key = keys[i];
    int f = 0;
    if (i == 0 || keys[i - 1] != key) {
      double s = vals[i];
      for (u64 q = i + 1; q < n && keys[q] == key; q++) s = __dadd_rn(s, vals[q]);
      sums[i] = s;
      f = (s != 0.0) ? 1 : 0;
    }
    flag[i] = f;
  }
}
__global__ void spgemm_emit_kernel(const u64* __restrict__ keys, const double* __restrict__ sums, const int* __restrict__ flag,
                                   const int* __restrict__ pos, u64 n, int* __restrict__ cci, double* __restrict__ cva,
                                   int* __restrict__ rowcnt) {
  const u64 stride = (u64)gridDim.x * blockDim.x;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (!flag[i]) continue;
    const u64 key = keys[i];
    const int p = pos[i];
    cci[p] = (int)(key & 0xffffffffull);
    cva[p] = sums[i];
    atomicAdd(rowcnt + (int)(key >> 32), 1);
  }
}

// ---------------- ultra-sparse left operand x dense: CSR result ----------------
// LibMatrixMult.matrixMultUltraSparse, left branch (LibMatrixMult.java:1510-1566): only the rows of A that hold cells produce
// anything; a row with one cell selects (and scales) a row of B, a longer row adds its products in ascending k, skipping
// products that are zero; results that are zero are not stored.  A thread looks at one row of A; the rows that hold cells are
// then taken one after the other by the whole warp, lanes over the columns of B.
template <bool FILL>
__global__ void __launch_bounds__(256) us_left_kernel(const int* __restrict__ arp, const int* __restrict__ aci, const double* __restrict__ ava,
                                                       const double* __restrict__ B, i64 rows, i64 n, int* __restrict__ cnt, const int* __restrict__ crp,
                                                       int* __restrict__ cci, double* __restrict__ cva) {
  const int lane = threadIdx.x & 31;
  const i64 nthreads = (i64)gridDim.x * blockDim.x;
  for (i64 base = ((i64)blockIdx.x * blockDim.x + threadIdx.x) - lane; base < rows; base += nthreads) {
    const i64 r = base + lane;
    int lo = 0, hi = 0;
    if (r < rows) {
      lo = arp[r];
      hi = arp[r + 1];
      if (!FILL && hi <= lo) cnt[r] = 0;  // (rows with cells get their count from lane 0 below)
    }
    unsigned todo = __ballot_sync(0xffffffffu, hi > lo);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const int klo = __shfl_sync(0xffffffffu, lo, src), khi = __shfl_sync(0xffffffffu, hi, src);
      const i64 row = base + src;
      int written = 0;
      const int obase = FILL ? crp[row] : 0;
      for (i64 j0 = 0; j0 < n; j0 += 32) {
        const i64 j = j0 + lane;
        double c = 0.0;
        if (j < n) {
          for (int k = klo; k < khi; k++) {
            const double p = __dmul_rn(__ldg(ava + k), __ldg(B + (i64)__ldg(aci + k) * n + j));
            if (p != 0.0) c = __dadd_rn(c, p);
          }
        }
        const bool keep = (j < n) && (c != 0.0);
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (FILL && keep) {
          const int pos = obase + written + __popc(m & ((1u << lane) - 1u));
          cci[pos] = (int)j;
          cva[pos] = c;
        }
        written += __popc(m);
      }
      if (!FILL && lane == 0) cnt[row] = written;
    }
  }
}

int exclusive_scan_int(sysml_b200_ctx* ctx, const int* in, int* out, i64 n, void* tmp, size_t tmp_bytes) {
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, in, out, (int)n, ctx->stream));
  return SYSML_OK;
}

}  // namespace

extern "C" {

int sysml_b200_csr_transpose(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, int64_t rows,
                             int64_t cols, int32_t* t_rowptr, int32_t* t_colind, double* t_val) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0 || rows > 0x7fffffffLL || cols > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_DIM, "csr_transpose: bad dimensions");
  if (!t_rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_transpose: null output");
  SYSML_CUDA(ctx, cudaMemsetAsync(t_rowptr, 0, (size_t)(cols + 1) * sizeof(int), ctx->stream));
  if (rows == 0 || cols == 0) return SYSML_OK;
  if (!rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_transpose: null input");
  int nnz = 0;
  int rc = fetch_int(ctx, rowptr + rows, &nnz);
  if (rc) return rc;
  if (nnz == 0) return SYSML_OK;
  if (!colind || !val || !t_colind || !t_val) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_transpose: null pointer");
  size_t sort_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const u64*)nullptr, (u64*)nullptr, (const double*)nullptr,
                                                  (double*)nullptr, nnz, 0, 64, ctx->stream));
  size_t scan_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const int*)nullptr, (int*)nullptr, (int)(cols + 1), ctx->stream));
  const size_t tmp_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
  void* ws = nullptr;
  rc = sysml_workspace(ctx, 2 * padded((size_t)nnz * 8) + padded((size_t)(cols + 1) * 4) + padded(tmp_bytes) + 1024, &ws);
  if (rc) return rc;
  char* p = (char*)ws;
  u64* keys = carve<u64>(p, nnz);
  u64* keys2 = carve<u64>(p, nnz);
  int* colcnt = carve<int>(p, cols + 1);
  void* tmp = p;
  SYSML_CUDA(ctx, cudaMemsetAsync(colcnt, 0, (size_t)(cols + 1) * sizeof(int), ctx->stream));
  tr_keys_kernel<<<grid_for(ctx, rows * 32, 256), 256, 0, ctx->stream>>>(rowptr, colind, rows, keys, colcnt);
  SYSML_LAUNCH_CHECK(ctx, "tr_keys_kernel");
  int bits = 33;  // the column sits above bit 32
  while (bits < 64 && (1ll << (bits - 32)) < cols) bits++;
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, sort_bytes, keys, keys2, val, t_val, nnz, 0, bits, ctx->stream));
  low32_kernel<<<grid_for(ctx, nnz, 256), 256, 0, ctx->stream>>>(keys2, nnz, t_colind);
  SYSML_LAUNCH_CHECK(ctx, "low32_kernel");
  rc = exclusive_scan_int(ctx, colcnt, t_rowptr, cols + 1, tmp, scan_bytes);
  return rc;
}

int sysml_b200_csr_add(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val, double alpha,
                       const int32_t* b_rowptr, const int32_t* b_colind, const double* b_val, double beta, int64_t rows, int64_t cols,
                       int32_t* c_rowptr, int32_t** c_colind, double** c_val, int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0 || rows > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_DIM, "csr_add: bad dimensions");
  if (!c_rowptr || !c_colind || !c_val || !c_nnz) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_add: null output");
  *c_colind = nullptr;
  *c_val = nullptr;
  *c_nnz = 0;
  SYSML_CUDA(ctx, cudaMemsetAsync(c_rowptr, 0, (size_t)(rows + 1) * sizeof(int), ctx->stream));
  if (rows == 0) return SYSML_OK;
  if (!a_rowptr || !b_rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_add: null input");
  size_t scan_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const int*)nullptr, (int*)nullptr, (int)(rows + 1), ctx->stream));
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, padded((size_t)(rows + 1) * 4) + padded(scan_bytes) + 512, &ws);
  if (rc) return rc;
  char* p = (char*)ws;
  int* cnt = carve<int>(p, rows + 1);
  void* tmp = p;
  SYSML_CUDA(ctx, cudaMemsetAsync(cnt, 0, (size_t)(rows + 1) * sizeof(int), ctx->stream));
  geam_kernel<false><<<grid_for(ctx, rows, 128), 128, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, b_rowptr, b_colind, b_val, alpha, beta,
                                                                        rows, cnt, nullptr, nullptr);
  SYSML_LAUNCH_CHECK(ctx, "geam_kernel<count>");
  rc = exclusive_scan_int(ctx, cnt, c_rowptr, rows + 1, tmp, scan_bytes);
  if (rc) return rc;
  int nnz = 0;
  rc = fetch_int(ctx, c_rowptr + rows, &nnz);
  if (rc) return rc;
  *c_nnz = nnz;
  if (nnz == 0) return SYSML_OK;
  cudaError_t e = cudaMalloc((void**)c_colind, (size_t)nnz * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc((void**)c_val, (size_t)nnz * sizeof(double));
  if (e != cudaSuccess) {
    if (*c_colind) cudaFree(*c_colind);
    *c_colind = nullptr;
    *c_val = nullptr;
    return sysml_set_error(ctx, SYSML_E_OOM, "csr_add: cudaMalloc of %d result entries: %s", nnz, cudaGetErrorString(e));
  }
  geam_kernel<true><<<grid_for(ctx, rows, 128), 128, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, b_rowptr, b_colind, b_val, alpha, beta,
                                                                       rows, c_rowptr, *c_colind, *c_val);
  SYSML_LAUNCH_CHECK(ctx, "geam_kernel<fill>");
  return SYSML_OK;
}

int sysml_b200_matmult_ultra_sparse_left(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val, const double* B,
                                         int64_t a_rows, int64_t a_cols, int64_t b_cols, int32_t* c_rowptr, int32_t* c_colind, double* c_val,
                                         int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  if (a_rows < 0 || a_cols < 0 || b_cols < 0 || a_rows >= 0x7fffffffLL || b_cols > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_ultra_sparse_left: bad dimensions");
  if (!c_rowptr || !c_nnz) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_ultra_sparse_left: null c_rowptr / c_nnz");
  if ((c_colind == nullptr) != (c_val == nullptr)) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_ultra_sparse_left: c_colind and c_val must both be NULL (symbolic pass) or both be set");
  const bool fill = c_colind != nullptr;
  if (a_rows == 0 || b_cols == 0 || a_cols == 0) {
    if (!fill) {
      SYSML_CUDA(ctx, cudaMemsetAsync(c_rowptr, 0, (size_t)(a_rows + 1) * sizeof(int), ctx->stream));
      *c_nnz = 0;
    }
    return SYSML_OK;
  }
  if (!a_rowptr || !B) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_ultra_sparse_left: null input");
  const unsigned g = grid_for(ctx, a_rows, 256);
  if (fill) {
    us_left_kernel<true><<<g, 256, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, B, a_rows, b_cols, nullptr, c_rowptr, c_colind, c_val);
    SYSML_LAUNCH_CHECK(ctx, "us_left_kernel<fill>");
    return SYSML_OK;
  }
  size_t scan_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const int*)nullptr, (int*)nullptr, (int)(a_rows + 1), ctx->stream));
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, padded((size_t)(a_rows + 1) * 4) + padded(scan_bytes) + 512, &ws);
  if (rc) return rc;
  char* p = (char*)ws;
  int* cnt = carve<int>(p, a_rows + 1);
  void* tmp = p;
  SYSML_CUDA(ctx, cudaMemsetAsync(cnt + a_rows, 0, sizeof(int), ctx->stream));
  us_left_kernel<false><<<g, 256, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, B, a_rows, b_cols, cnt, nullptr, nullptr, nullptr);
  SYSML_LAUNCH_CHECK(ctx, "us_left_kernel<count>");
  rc = exclusive_scan_int(ctx, cnt, c_rowptr, a_rows + 1, tmp, scan_bytes);
  if (rc) return rc;
  int nnz = 0;
  rc = fetch_int(ctx, c_rowptr + a_rows, &nnz);
  if (rc) return rc;
  *c_nnz = nnz;
  return SYSML_OK;
}

int sysml_b200_matmult_csr_csr_sparse(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val,
                                      const int32_t* b_rowptr, const int32_t* b_colind, const double* b_val, int64_t a_rows,
                                      int64_t a_cols, int64_t b_rows, int64_t b_cols, int32_t* c_rowptr, int32_t** c_colind,
                                      double** c_val, int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  if (a_rows < 0 || a_cols < 0 || b_rows < 0 || b_cols < 0 || a_rows > 0x7fffffffLL || b_cols > 0x7fffffffLL)
    return sysml_set_error(ctx, SYSML_E_DIM, "matmult_csr_csr_sparse: bad dimensions");
  if (a_cols != b_rows) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_csr_csr_sparse: inner dimensions %lld vs %lld", (long long)a_cols, (long long)b_rows);
  if (!c_rowptr || !c_colind || !c_val || !c_nnz) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_csr_csr_sparse: null output");
  *c_colind = nullptr;
  *c_val = nullptr;
  *c_nnz = 0;
  SYSML_CUDA(ctx, cudaMemsetAsync(c_rowptr, 0, (size_t)(a_rows + 1) * sizeof(int), ctx->stream));
  if (a_rows == 0 || b_cols == 0 || a_cols == 0) return SYSML_OK;
  if (!a_rowptr || !b_rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_csr_csr_sparse: null input");
  // ---- symbolic: how many products per row ----
  size_t scan64_bytes = 0, scan_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan64_bytes, (const u64*)nullptr, (u64*)nullptr, (int)(a_rows + 1), ctx->stream));
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const int*)nullptr, (int*)nullptr, (int)(a_rows + 1), ctx->stream));
  void* ws0 = nullptr;
  int rc = sysml_workspace(ctx, 2 * padded((size_t)(a_rows + 1) * 8) + padded(scan64_bytes) + 1024, &ws0);
  if (rc) return rc;
  u64 total = 0;
  {
    char* p = (char*)ws0;
    u64* cnt = carve<u64>(p, a_rows + 1);
    u64* off = carve<u64>(p, a_rows + 1);
    void* tmp = p;
    SYSML_CUDA(ctx, cudaMemsetAsync(cnt, 0, (size_t)(a_rows + 1) * sizeof(u64), ctx->stream));
    spgemm_count_kernel<<<grid_for(ctx, a_rows, 128), 128, 0, ctx->stream>>>(a_rowptr, a_colind, b_rowptr, a_rows, cnt);
    SYSML_LAUNCH_CHECK(ctx, "spgemm_count_kernel");
    SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, scan64_bytes, cnt, off, (int)(a_rows + 1), ctx->stream));
    rc = fetch_u64(ctx, off + a_rows, &total);
    if (rc) return rc;
  }
  if (total == 0) return SYSML_OK;
  if (total > (1ull << 28))
    return sysml_set_error(ctx, SYSML_E_LIMIT, "matmult_csr_csr_sparse: %llu intermediate products exceed the limit of 2^28 (use the dense-output entry point)", total);
  // ---- numeric: expand, stable sort by (row, col), add each run left to right ----
  size_t sort_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const u64*)nullptr, (u64*)nullptr, (const double*)nullptr,
                                                  (double*)nullptr, (int)total, 0, 64, ctx->stream));
  size_t scan_tot_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_tot_bytes, (const int*)nullptr, (int*)nullptr, (int)total, ctx->stream));
  size_t tmp_bytes = sort_bytes > scan_tot_bytes ? sort_bytes : scan_tot_bytes;
  if (scan_bytes > tmp_bytes) tmp_bytes = scan_bytes;
  if (scan64_bytes > tmp_bytes) tmp_bytes = scan64_bytes;
  const size_t head = 2 * padded((size_t)(a_rows + 1) * 8) + padded((size_t)(a_rows + 1) * 4);
  void* ws = nullptr;
  rc = sysml_workspace(ctx, head + 4 * padded(total * 8) + 2 * padded(total * 4) + padded(tmp_bytes) + 2048, &ws);
  if (rc) return rc;
  char* p = (char*)ws;
  u64* cnt = carve<u64>(p, a_rows + 1);
  u64* off = carve<u64>(p, a_rows + 1);
  int* rowcnt = carve<int>(p, a_rows + 1);
  u64* keys = carve<u64>(p, total);
  u64* keys2 = carve<u64>(p, total);
  double* vals = carve<double>(p, total);
  double* vals2 = carve<double>(p, total);
  int* flag = carve<int>(p, total);
  int* pos = carve<int>(p, total);
  void* tmp = p;
  // the workspace may have moved when it grew: recompute the offsets
  SYSML_CUDA(ctx, cudaMemsetAsync(cnt, 0, (size_t)(a_rows + 1) * sizeof(u64), ctx->stream));
  spgemm_count_kernel<<<grid_for(ctx, a_rows, 128), 128, 0, ctx->stream>>>(a_rowptr, a_colind, b_rowptr, a_rows, cnt);
  SYSML_LAUNCH_CHECK(ctx, "spgemm_count_kernel");
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, scan64_bytes, cnt, off, (int)(a_rows + 1), ctx->stream));
  spgemm_expand_kernel<<<grid_for(ctx, a_rows * 32, 256), 256, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, b_rowptr, b_colind, b_val, a_rows,
                                                                                 off, keys, vals);
  SYSML_LAUNCH_CHECK(ctx, "spgemm_expand_kernel");
  int bits = 33;
  while (bits < 64 && (1ll << (bits - 32)) < a_rows) bits++;
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, sort_bytes, keys, keys2, vals, vals2, (int)total, 0, bits, ctx->stream));
  spgemm_runs_kernel<<<grid_for(ctx, (i64)total, 256), 256, 0, ctx->stream>>>(keys2, vals2, total, flag, vals);
  SYSML_LAUNCH_CHECK(ctx, "spgemm_runs_kernel");
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, scan_tot_bytes, flag, pos, (int)total, ctx->stream));
  int last_pos = 0, last_flag = 0;
  rc = fetch_int(ctx, pos + (total - 1), &last_pos);
  if (rc) return rc;
  rc = fetch_int(ctx, flag + (total - 1), &last_flag);
  if (rc) return rc;
  const int nnz = last_pos + last_flag;
  *c_nnz = nnz;
  if (nnz == 0) return SYSML_OK;
  cudaError_t e = cudaMalloc((void**)c_colind, (size_t)nnz * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc((void**)c_val, (size_t)nnz * sizeof(double));
  if (e != cudaSuccess) {
    if (*c_colind) cudaFree(*c_colind);
    *c_colind = nullptr;
    *c_val = nullptr;
    return sysml_set_error(ctx, SYSML_E_OOM, "matmult_csr_csr_sparse: cudaMalloc of %d result entries: %s", nnz, cudaGetErrorString(e));
  }
  SYSML_CUDA(ctx, cudaMemsetAsync(rowcnt, 0, (size_t)(a_rows + 1) * sizeof(int), ctx->stream));
  spgemm_emit_kernel<<<grid_for(ctx, (i64)total, 256), 256, 0, ctx->stream>>>(keys2, vals, flag, pos, total, *c_colind, *c_val, rowcnt);
  SYSML_LAUNCH_CHECK(ctx, "spgemm_emit_kernel");
  return exclusive_scan_int(ctx, rowcnt, c_rowptr, a_rows + 1, tmp, scan_bytes);
}

}  // extern "C"
