// pcsc.cu -- a second, column-ordered copy of a constant CSR matrix ("panel CSC") and the products that want it:
//
//     out = t(X) %*% y                       (matrixMultSparseDense VM after the left-transpose rewrite,
//                                             LibMatrixMult.java:1262-1286; AggBinaryOp.java:689-713)
//     out = t(X) %*% (w * (X %*% v)) ...     (matrixMultChainSparse, LibMatrixMult.java:1655-1684)
//
// Why.  Row-ordered CSR makes t(X) %*% y a scatter: c[colind[k]] += y[i] * val[k].  On the GPU that is an fp64 atomic
// per non-zero -- shared-memory fp64 atomics are compare-and-swap loops, L2 reductions are a fraction of the HBM
// rate -- and the order of the additions, hence the low bits of the result, changes from run to run.  X is constant
// across the iterations of L2SVM / GLM / LinearRegCG, so it pays to keep a copy in which the scatter is a gather.
// A plain CSC copy is not that copy: its gathers go to y, an n-vector (400 MB at C3) that no cache holds.
//
// Layout.  Rows are cut into panels of R rows (R <= 65536, 8192 by default).  Inside a panel the entries are sorted
// by (column, row) and stored as three parallel arrays -- value f64, row-within-panel u16, column u16 -- 12 bytes per
// non-zero, the same as CSR.  Panels follow each other in row order, so the arrays have exactly nnz entries and a
// panel is one contiguous range.  Each panel range is further cut into 32 column slices (equal column counts);
// slice_ptr[p*32 + s] is where slice s of panel p starts (slice_ptr[npanels*32] = nnz).
//
// Kernel.  One persistent CTA per SM, 32 warps.  Per panel: the R values of y (times / minus w for the chain types)
// are parked in shared memory; warp s streams slice s of the panel with coalesced loads, four consecutive entries
// per lane, multiplies by the gathered y, and reduces equal-column runs with a segmented warp scan; finished runs are
// added to the CTA's private m-vector in shared memory with plain read-modify-writes -- a column belongs to exactly
// one slice, hence to one warp, so there is no atomic anywhere and the order of every addition is fixed.  The per-CTA
// vectors are summed by the fixed-order compensated reduction the dense kernels use.  Result: bit-identical run to run.
//
// Algorithmic bytes per product: 12*nnz + 8*n (y) [+ 8*n (w)] + 8*m.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

int sysml_reduce_partial_rows(sysml_b200_ctx* ctx, const double* partial, int P, int m, double* out);
extern "C" int sysml_b200_matmult_csr_dense(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                            const double* B, double* C, int64_t a_rows, int64_t a_cols, int64_t b_rows, int64_t b_cols,
                                            int trans_a);

namespace {

typedef unsigned long long u64;
constexpr int NSLICE = 32;

__global__ void pcsc_keys_kernel(const int* __restrict__ rp, const int* __restrict__ ci, i64 rows, int R, u64* __restrict__ keys) {
  // one thread per row (rows are short): key = panel << 32 | column << 16 | row-within-panel
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    const u64 hi = ((u64)(r / R) << 32) | (u64)(r % R);
    for (int k = rp[r]; k < rp[r + 1]; k++) keys[k] = hi | ((u64)(unsigned)ci[k] << 16);
  }
}

// unpack the sorted keys; entry e starts every (panel, slice) group between its predecessor's group and its own
__global__ void pcsc_unpack_kernel(const u64* __restrict__ keys, i64 nnz, int slice_cols, i64 ngroups, unsigned short* __restrict__ lrow,
                                   unsigned short* __restrict__ col, int* __restrict__ slice_ptr) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 e = (i64)blockIdx.x * blockDim.x + threadIdx.x; e <= nnz; e += stride) {
    i64 g, gprev;
    if (e < nnz) {
      const u64 k = keys[e];
      const unsigned c = (unsigned)(k >> 16) & 0xffffu;
      lrow[e] = (unsigned short)(k & 0xffffu);
      col[e] = (unsigned short)c;
      g = (i64)(k >> 32) * NSLICE + c / slice_cols;
    } else {
      g = ngroups;  // one past the last group: closes every trailing empty group and writes the total
    }
    if (e == 0) {
      gprev = -1;
    } else {
      const u64 kp = keys[e - 1];
      gprev = (i64)(kp >> 32) * NSLICE + ((unsigned)(kp >> 16) & 0xffffu) / slice_cols;
    }
    for (i64 q = gprev + 1; q <= g; q++) slice_ptr[q] = (int)e;
  }
}

struct PcscArgs {
  const double* val;
  const unsigned short* lrow;
  const unsigned short* col;
  const int* slice_ptr;
  const double* y;
  const double* w;  // nullable
  int wop;          // 0: y ; 1: y * w ; 2: y - w
  double* partial;  // [gridDim.x][m]
  i64 n;
  int m, R;
  i64 npanels;
};

__global__ void __launch_bounds__(1024, 1) pcsc_vm_kernel(const PcscArgs a) {
  extern __shared__ __align__(16) double smem[];
  double* c_s = smem;        // m
  double* y_s = smem + a.m;  // R
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < a.m; j += 1024) c_s[j] = 0.0;
  for (i64 p = blockIdx.x; p < a.npanels; p += gridDim.x) {
    __syncthreads();  // previous panel's gathers are done (and c_s is zeroed on the first trip)
    const i64 r0 = p * (i64)a.R;
    for (int i = tid; i < a.R; i += 1024) {
      double yv = 0.0;
      if (r0 + i < a.n) {
        yv = ldg_stream(a.y + r0 + i);
        if (a.wop == 1) yv = __dmul_rn(yv, ldg_stream(a.w + r0 + i));
        else if (a.wop == 2) yv = __dsub_rn(yv, ldg_stream(a.w + r0 + i));
      }
      y_s[i] = yv;
    }
    __syncthreads();
    const int* sp = a.slice_ptr + p * NSLICE + warp;
    const i64 lo = __ldg(sp), hi = __ldg(sp + 1);
    // The warp walks its slice in steps of 128 entries, lane l holding the four consecutive entries 4l..4l+3 of the step
    // (two 16-byte and two 8-byte loads per lane; the walk starts at a multiple of four so that they are aligned).
    // Entries before `lo` belong to the previous slice: they get the key of the lane's first own entry and a zero
    // product; entries past `hi` get a sentinel key that forms its own run and is never written.  The loads of UNR steps
    // are issued before the first of them is reduced: one step in flight per warp does not cover the HBM latency.
    constexpr int UNR = 2;
    constexpr unsigned SENT = 0xffffffffu;
    for (i64 base0 = lo & ~(i64)3; base0 < hi; base0 += 128 * UNR) {
      double pv[UNR][4];
      unsigned kk[UNR][4], rw[UNR][4];
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const i64 e0 = base0 + 128 * u + 4 * lane;
        if (e0 + 3 < hi) {
          const double2 x01 = ldg_stream2(reinterpret_cast<const double2*>(a.val + e0));
          const double2 x23 = ldg_stream2(reinterpret_cast<const double2*>(a.val + e0 + 2));
          const uint2 c4 = __ldg(reinterpret_cast<const uint2*>(a.col + e0));
          const uint2 r4 = __ldg(reinterpret_cast<const uint2*>(a.lrow + e0));
          pv[u][0] = x01.x; pv[u][1] = x01.y; pv[u][2] = x23.x; pv[u][3] = x23.y;
          kk[u][0] = c4.x & 0xffffu; kk[u][1] = c4.x >> 16; kk[u][2] = c4.y & 0xffffu; kk[u][3] = c4.y >> 16;
          rw[u][0] = r4.x & 0xffffu; rw[u][1] = r4.x >> 16; rw[u][2] = r4.y & 0xffffu; rw[u][3] = r4.y >> 16;
        } else {
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const bool in = e0 + j < hi;
            pv[u][j] = in ? ldg_stream(a.val + e0 + j) : 0.0;
            kk[u][j] = in ? (unsigned)a.col[e0 + j] : SENT;
            rw[u][j] = in ? (unsigned)a.lrow[e0 + j] : 0u;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const i64 e0 = base0 + 128 * u + 4 * lane;
        if (base0 + 128 * u >= hi) break;  // warp-uniform
        double p[4];
        unsigned k[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          k[j] = kk[u][j];
          p[j] = (k[j] != SENT && e0 + j >= lo) ? __dmul_rn(pv[u][j], y_s[rw[u][j]]) : 0.0;
        }
        if (e0 < lo) {  // only lane 0 of a slice's first step: 1..3 foreign entries, then at least one own entry
          const unsigned own = (e0 + 1 >= lo) ? k[1] : (e0 + 2 >= lo) ? k[2] : k[3];
#pragma unroll
          for (int j = 0; j < 3; j++)
            if (e0 + j < lo) k[j] = own;
        }
        // runs inside the lane: the head run (entries equal to k[0]), the tail run (equal to k[3]) and up to two
        // entries in between, whose columns no other lane holds
        const bool eq01 = k[0] == k[1], eq12 = k[1] == k[2], eq23 = k[2] == k[3];
        const bool uniform = eq01 && eq12 && eq23;
        double H = p[0];
        if (eq01) H = __dadd_rn(H, p[1]);
        if (eq01 && eq12) H = __dadd_rn(H, p[2]);
        if (uniform) H = __dadd_rn(H, p[3]);
        double T = (eq23 && eq12) ? p[1] : 0.0;
        if (eq23) T = __dadd_rn(T, p[2]);
        T = __dadd_rn(T, p[3]);
        const bool mid1 = !eq01 && !(eq12 && eq23), mid2 = !(eq01 && eq12) && !eq23;
        const unsigned kin = k[0], kout = k[3];
        // S: sum of the run that LEAVES this lane to the right (so far); a uniform lane passes an incoming run through
        double S = uniform ? H : T;
        const unsigned kl = __shfl_up_sync(0xffffffffu, kout, 1);  // key leaving the left neighbour
        bool cont = uniform && lane > 0 && kl == kin;               // this lane extends the neighbour's run
        // segmented inclusive scan over the lanes (Hillis-Steele; `cont` = "the run reaches further back than the lanes
        // folded in so far"); runs are short, so most steps are done after two or three rounds
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          if (!__any_sync(0xffffffffu, cont)) break;
          const double up = __shfl_up_sync(0xffffffffu, S, d);
          const bool upc = __shfl_up_sync(0xffffffffu, (int)cont, d) != 0;
          if (cont && lane >= d) {
            S = __dadd_rn(up, S);
            cont = upc;
          }
        }
        const double Sl = __shfl_up_sync(0xffffffffu, S, 1);
        const unsigned kr = __shfl_down_sync(0xffffffffu, kin, 1);  // key entering the right neighbour
        // (a) a non-uniform lane ends, with its head run, whatever run came in from the left; its in-between entries
        //     are complete runs of their own
        if (!uniform) {
          if (kin != SENT) {
            const double Hf = (lane > 0 && kl == kin) ? __dadd_rn(Sl, H) : H;
            c_s[kin] = __dadd_rn(c_s[kin], Hf);
          }
          if (mid1 && mid2 && eq12) {
            c_s[k[1]] = __dadd_rn(c_s[k[1]], __dadd_rn(p[1], p[2]));
          } else {
            if (mid1) c_s[k[1]] = __dadd_rn(c_s[k[1]], p[1]);
            if (mid2) c_s[k[2]] = __dadd_rn(c_s[k[2]], p[2]);
          }
        }
        __syncwarp();
        // (b) the run leaving this lane ends here when the right neighbour does not take it on
        if (kout != SENT && (lane == 31 || kr != kout)) c_s[kout] = __dadd_rn(c_s[kout], S);
        __syncwarp();
      }
    }
  }
  __syncthreads();
  double* prow = a.partial + (size_t)blockIdx.x * a.m;
  for (int j = tid; j < a.m; j += 1024) prow[j] = c_s[j];
}

}  // namespace

extern "C" {

int64_t sysml_b200_pcsc_slice_entries(int64_t rows, int panel_rows) {
  if (rows < 0 || panel_rows < 1) return -1;
  return ((rows + panel_rows - 1) / panel_rows) * NSLICE + 1;
}

int sysml_b200_csr_to_pcsc(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, int64_t rows, int64_t cols,
                           int64_t nnz, int panel_rows, double* p_val, uint16_t* p_lrow, uint16_t* p_col, int32_t* slice_ptr) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 1 || nnz < 0 || nnz > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_DIM, "csr_to_pcsc: bad dimensions");
  if (cols > 65536) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_pcsc: more than 65536 columns (16-bit column indices)");
  if (panel_rows < 1 || panel_rows > 65536) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_pcsc: panel_rows must be in [1, 65536]");
  if (!slice_ptr) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_to_pcsc: slice_ptr is null");
  const i64 npanels = (rows + panel_rows - 1) / panel_rows;
  if (npanels >= (1ll << 26)) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_pcsc: too many panels");
  const i64 ngroups = npanels * NSLICE;
  if (nnz == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(slice_ptr, 0, (size_t)(ngroups + 1) * sizeof(int), ctx->stream));
    return SYSML_OK;
  }
  if (!rowptr || !colind || !val || !p_val || !p_lrow || !p_col) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_to_pcsc: null pointer");
  int pbits = 1;
  while ((1ll << pbits) < npanels) pbits++;
  size_t sort_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const u64*)nullptr, (u64*)nullptr, (const double*)nullptr,
                                                  (double*)nullptr, (int)nnz, 0, 32 + pbits, ctx->stream));
  const size_t kb = (((size_t)nnz * 8) + 255) & ~(size_t)255;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, 2 * kb + sort_bytes + 256, &ws);
  if (rc) return rc;
  u64* keys = (u64*)ws;
  u64* keys2 = (u64*)((char*)ws + kb);
  void* tmp = (char*)ws + 2 * kb;
  i64 blocks = (rows + 255) / 256;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  pcsc_keys_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(rowptr, colind, rows, panel_rows, keys);
  SYSML_LAUNCH_CHECK(ctx, "pcsc_keys_kernel");
  // stable LSD radix sort: entries with equal (panel, column, row) do not exist, so the order is fully determined
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, sort_bytes, keys, keys2, val, p_val, (int)nnz, 0, 32 + pbits, ctx->stream));
  ctx->launches++;
  const int slice_cols = (int)((cols + NSLICE - 1) / NSLICE);
  blocks = (nnz + 256) / 256;
  if (blocks > cap) blocks = cap;
  pcsc_unpack_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(keys2, nnz, slice_cols, ngroups, p_lrow, p_col, slice_ptr);
  SYSML_LAUNCH_CHECK(ctx, "pcsc_unpack_kernel");
  return SYSML_OK;
}

int sysml_b200_matmult_pcsc_vm(sysml_b200_ctx* ctx, const double* p_val, const uint16_t* p_lrow, const uint16_t* p_col,
                               const int32_t* slice_ptr, int64_t rows, int64_t cols, int panel_rows, const double* y, const double* w,
                               int w_op, double* out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_pcsc_vm: negative dimension");
  if (cols == 0) return SYSML_OK;
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_pcsc_vm: out is null");
  if (rows == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)cols * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (cols > 65536 || panel_rows < 1 || panel_rows > 65536) return sysml_set_error(ctx, SYSML_E_LIMIT, "matmult_pcsc_vm: cols and panel_rows must be <= 65536");
  if (w_op < 0 || w_op > 2 || (w_op && !w)) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_pcsc_vm: bad weight operation");
  if (!p_val || !p_lrow || !p_col || !slice_ptr || !y) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_pcsc_vm: null pointer");
  const size_t smem = ((size_t)cols + (size_t)panel_rows) * sizeof(double);
  if (smem > (size_t)ctx->max_smem_optin)
    return sysml_set_error(ctx, SYSML_E_LIMIT, "matmult_pcsc_vm: cols + panel_rows = %lld doubles do not fit in shared memory (%d bytes)",
                           (long long)(cols + panel_rows), ctx->max_smem_optin);
  SYSML_CUDA(ctx, cudaFuncSetAttribute(pcsc_vm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  PcscArgs a{};
  a.val = p_val; a.lrow = p_lrow; a.col = p_col; a.slice_ptr = slice_ptr; a.y = y; a.w = w; a.wop = w_op;
  a.n = rows; a.m = (int)cols; a.R = panel_rows;
  a.npanels = (rows + panel_rows - 1) / panel_rows;
  i64 grid = ctx->sm_count;
  if (grid > a.npanels) grid = a.npanels;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, (size_t)grid * cols * sizeof(double), &ws);
  if (rc) return rc;
  a.partial = (double*)ws;
  pcsc_vm_kernel<<<(unsigned)grid, 1024, smem, ctx->stream>>>(a);
  SYSML_LAUNCH_CHECK(ctx, "pcsc_vm_kernel");
  return sysml_reduce_partial_rows(ctx, a.partial, (int)grid, (int)cols, out);
}

int sysml_b200_mmchain_csr_pcsc(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, const double* p_val,
                                const uint16_t* p_lrow, const uint16_t* p_col, const int32_t* slice_ptr, int panel_rows, const double* v,
                                const double* w, double* tmp_n, double* out, int64_t n, int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  if (chain_type < 0 || chain_type > 2) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr_pcsc: unknown chain type %d", chain_type);
  if (chain_type != SYSML_CHAIN_XtXv && n > 0 && !w) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr_pcsc: chain type %d needs w", chain_type);
  if (n > 0 && !tmp_n) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr_pcsc: the n-vector temporary is null");
  // X %*% v is a gather over the row-ordered copy; the weights are applied while the panel of the result is parked
  int rc = sysml_b200_matmult_csr_dense(ctx, rowptr, colind, val, v, tmp_n, n, m, m, 1, 0);
  if (rc) return rc;
  return sysml_b200_matmult_pcsc_vm(ctx, p_val, p_lrow, p_col, slice_ptr, n, m, panel_rows, tmp_n, chain_type ? w : nullptr, chain_type, out);
}

}  // extern "C"
