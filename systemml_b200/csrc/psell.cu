// psell.cu -- "panelled sliced ELL": the column-ordered second copy of a constant CSR matrix, laid out so that
//
//     out = t(X) %*% y                       (matrixMultSparseDense VM after the left-transpose rewrite,
//                                             LibMatrixMult.java:1262-1286; AggBinaryOp.java:689-713)
//     out = t(X) %*% (w * (X %*% v)) ...     (matrixMultChainSparse, LibMatrixMult.java:1655-1684)
//
// are gathers with NO cross-lane work: no atomics, no segmented scan, a handful of instructions per non-zero.
//
// Why a second copy.  Row-ordered CSR makes t(X) %*% y a scatter: c[colind[k]] += y[i] * val[k].  On the GPU that is an
// fp64 atomic per non-zero -- shared-memory fp64 atomics are compare-and-swap loops, L2 reductions are a fraction of the
// HBM rate -- and the order of the additions, hence the low bits of the result, changes from run to run.  X is constant
// across the iterations of L2SVM / GLM / LinearRegCG, so it pays to keep a copy in which the scatter is a gather.  A
// plain CSC copy is not that copy: its gathers go to y, an n-vector (400 MB at C3) that no cache holds; hence panels.
// Why this layout.  The first version of the copy kept a panel's entries sorted by (column, row) and let a warp fold
// equal columns in neighbouring lanes with a segmented scan: ~1.8 warp instructions per non-zero, the issue slots ran
// out before the memory did and it streamed at 0.62 of the copy rate (profiles/r02_ncu_pcsc_vm_summary.txt).  Here a
// lane OWNS a column run; issue slots are 29 % busy and what is left is latency (profiles/r02_ncu_psell_vm_summary.txt),
// which an L2 prefetch two groups ahead takes from 1.42 to 1.20 ms at C3 (0.82 of the copy rate).
//
// Layout.  Rows are cut into panels of R rows, the columns into 32 slices (one per warp).  Inside a (panel, slice) the
// entries of one column are a "piece" (a column with more than LMAX entries in the panel is cut into several);
// pieces are sorted by length, longest first, and taken 32 at a time as a GROUP: lane l of the group holds piece l,
// entry j of the piece at   val[(off + j) * 32 + l], lrow[(off + j) * 32 + l]   -- every j is one coalesced 256-byte
// (values) + 64-byte (16-bit rows within the panel) line.  A group is as long as its first piece; shorter pieces are
// padded with value 0 and row R, and y_s[R] is kept at 0, so a padded slot adds +0.0 and needs no test.  Sorting by
// length keeps the padding at a few percent for Poisson-like column counts and bounded by LMAX for any input.
// Per group: 32 column ids (u16, 0xFFFF = no piece) and a descriptor {off, len, dup}; dup marks a group in which two
// lanes hold pieces of the SAME column (a cut column) -- its 32 final adds are then done one lane after the other.
// sgrp_ptr[p*32 + s] .. [p*32 + s + 1] are the groups of slice s of panel p.
//
// Kernel.  One persistent CTA per SM, 32 warps, the CTA's private m-vector and two panels of y in shared memory.
// While panel p is being multiplied every thread already holds its share of panel p+1's y (and w) in registers, so
// the only thing between two panels is one barrier.  Warp s walks the groups of slice s: per entry a lane does
// LDG value, LDG row, LDS y, DMUL, DADD -- products and sums separately rounded, in row order, like the reference's
// loop -- and at the end of the group one read-modify-write of its column in the private vector.  A column belongs
// to one slice, hence to one warp: no atomic anywhere, the order of every addition is fixed, the result is
// bit-identical from run to run.  The per-CTA vectors are summed by the fixed-order compensated reduction the dense
// kernels use.
//
// Measured and dropped (C3, 50M x 10k, ~10 entries per row; this layout: 1.20 ms with the L2 prefetch, 1.42 without):
//  - sorting the pieces of a whole panel instead of per slice (padding 0.45 % instead of 12 %), groups dealt to the warps
//    in rows of 32, one direction or serpentine: 1.25 / 1.33 ms with the prefetch, 1.55 without -- 10 % fewer bytes and
//    slower; a warp's groups are then 90 KB apart instead of one sequential stream per warp;
//  - 16 warps with 128 registers and 16 entries per lane in flight (one round trip per group of the usual length): 1.32 ms
//    without prefetch, 1.63 with it;  the same with up to four short groups sharing a batch: 1.43 ms.
//
// Bytes per product: 10 per stored slot (8 + 2; slots = nnz + padding) + 72 per group + 8*n (y) [+ 8*n (w)] + 8*m.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <stdlib.h>

#include <vector>

#include "common.cuh"

int sysml_reduce_partial_rows(sysml_b200_ctx* ctx, const double* partial, int P, int m, double* out);
extern "C" int sysml_b200_matmult_csr_dense(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                            const double* B, double* C, int64_t a_rows, int64_t a_cols, int64_t b_rows, int64_t b_cols,
                                            int trans_a);

namespace {

typedef unsigned long long u64;
constexpr int NSLICE = 32;
constexpr int LMAX = 64;  // longest piece; a group is at most this long

__global__ void psell_keys_kernel(const int* __restrict__ rp, const int* __restrict__ ci, i64 rows, int R, u64* __restrict__ keys) {
  // one thread per row (rows are short): key = panel << 32 | column << 16 | row-within-panel
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    const u64 hi = ((u64)(r / R) << 32) | (u64)(r % R);
    for (int k = rp[r]; k < rp[r + 1]; k++) keys[k] = hi | ((u64)(unsigned)ci[k] << 16);
  }
}

// position of the head of the (panel, column) run entry e belongs to, for an inclusive max-scan
struct RunHead {
  const u64* keys;
  __host__ __device__ int operator()(int e) const { return (e == 0 || (keys[e] >> 16) != (keys[e - 1] >> 16)) ? e : 0; }
};
// entry e starts a piece: every LMAX-th entry of its run
struct PieceHead {
  const int* runstart;
  __host__ __device__ bool operator()(int e) const { return ((e - runstart[e]) % LMAX) == 0; }
};

// per piece: sort key (panel, slice, longest first) and the piece count of its (panel, slice)
__global__ void psell_piece_keys_kernel(const u64* __restrict__ keys, const int* __restrict__ piece_start, int npieces, int slice_cols,
                                        u64* __restrict__ skey, int* __restrict__ sidx, int* __restrict__ seg_count) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < npieces; i += stride) {
    const int st = piece_start[i], cnt = piece_start[i + 1] - st;
    const u64 k = keys[st];
    const u64 seg = (k >> 32) * NSLICE + ((unsigned)(k >> 16) & 0xffffu) / (unsigned)slice_cols;
    skey[i] = (seg << 8) | (u64)(255 - cnt);
    sidx[i] = (int)i;
    atomicAdd(seg_count + seg, 1);
  }
}

__global__ void psell_group_count_kernel(const int* __restrict__ seg_count, i64 nseg, int* __restrict__ grp_count) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 s = (i64)blockIdx.x * blockDim.x + threadIdx.x; s < nseg; s += stride) grp_count[s] = (seg_count[s] + 31) / 32;
}

// one thread per (panel, slice): the length of each of its groups (= the length of the group's first, longest piece)
__global__ void psell_group_len_kernel(const int* __restrict__ seg_ptr, const int* __restrict__ sgrp_ptr, i64 nseg, const int* __restrict__ sidx,
                                       const int* __restrict__ piece_start, int* __restrict__ glen, int* __restrict__ gseg) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 s = (i64)blockIdx.x * blockDim.x + threadIdx.x; s < nseg; s += stride) {
    const int g0 = sgrp_ptr[s], g1 = sgrp_ptr[s + 1];
    for (int g = g0; g < g1; g++) {
      const int pi = sidx[seg_ptr[s] + 32 * (g - g0)];
      glen[g] = piece_start[pi + 1] - piece_start[pi];
      gseg[g] = (int)s;
    }
  }
}

// one warp per group: lay the 32 pieces out side by side
__global__ void __launch_bounds__(256) psell_fill_kernel(const u64* __restrict__ keys, const double* __restrict__ vals, const int* __restrict__ piece_start,
                                                         const int* __restrict__ sidx, const int* __restrict__ seg_ptr, const int* __restrict__ sgrp_ptr,
                                                         const int* __restrict__ gseg, const int* __restrict__ glen, const i64* __restrict__ goff,
                                                         i64 ngroups, int R, double* __restrict__ s_val, unsigned short* __restrict__ s_lrow,
                                                         unsigned short* __restrict__ s_gcol, uint2* __restrict__ s_gdesc) {
  const int lane = threadIdx.x & 31;
  const i64 nw = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 g = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5; g < ngroups; g += nw) {
    const int s = gseg[g];
    const int q = seg_ptr[s] + 32 * (int)(g - sgrp_ptr[s]) + lane;
    const bool valid = q < seg_ptr[s + 1];
    int st = 0, cnt = 0;
    unsigned col = 0xffffu;
    if (valid) {
      const int pi = sidx[q];
      st = piece_start[pi];
      cnt = piece_start[pi + 1] - st;
      col = (unsigned)(keys[st] >> 16) & 0xffffu;
    }
    const int len = glen[g];
    const i64 off = goff[g];
    for (int j = 0; j < len; j++) {
      const bool in = j < cnt;
      const i64 e = (off + j) * 32 + lane;
      s_val[e] = in ? vals[st + j] : 0.0;
      s_lrow[e] = in ? (unsigned short)(keys[st + j] & 0xffffu) : (unsigned short)R;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? col : 0x10000u + lane);
    const bool dup = __any_sync(0xffffffffu, __popc(peers) > 1);
    s_gcol[g * 32 + lane] = (unsigned short)col;
    if (lane == 0) s_gdesc[g] = make_uint2((unsigned)off, (unsigned)len | (dup ? 0x10000u : 0u));
  }
}

struct PsellArgs {
  const double* val;
  const unsigned short* lrow;
  const unsigned short* gcol;
  const uint2* gdesc;
  const int* sgrp_ptr;
  const double* y;
  const double* w;  // nullable
  int wop;          // 0: y ; 1: y * w ; 2: y - w
  double* partial;  // [gridDim.x][m]
  i64 n;
  int m, R;
  i64 npanels;
  int pf;   // L2 prefetch distance in groups (0: none)
  int pfc;  // groups per prefetch request
};

constexpr int YPT = 8;  // y values per thread per panel held in registers (R <= 1024 * YPT)

__device__ __forceinline__ void psell_fetch_y(const PsellArgs& a, i64 p, int tid, double (&yr)[YPT]) {
  const i64 r0 = p * (i64)a.R;
#pragma unroll
  for (int q = 0; q < YPT; q++) {
    const int i = tid + q * 1024;
    double yv = 0.0;
    if (p < a.npanels && i < a.R && r0 + i < a.n) {
      yv = ldg_stream(a.y + r0 + i);
      if (a.wop == 1) yv = __dmul_rn(yv, ldg_stream(a.w + r0 + i));
      else if (a.wop == 2) yv = __dsub_rn(yv, ldg_stream(a.w + r0 + i));
    }
    yr[q] = yv;
  }
}

__global__ void __launch_bounds__(1024, 1) psell_vm_kernel(const PsellArgs a) {
  extern __shared__ __align__(16) double smem[];
  double* c_s = smem;  // m
  double* y_buf[2] = {smem + a.m, smem + a.m + (a.R + 1)};  // R + 1 each: slot R stays 0 (the row of padded slots)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < a.m; j += 1024) c_s[j] = 0.0;
  if (tid == 0) y_buf[0][a.R] = y_buf[1][a.R] = 0.0;
  double yr[YPT];
  psell_fetch_y(a, blockIdx.x, tid, yr);
  // L2 prefetch cursor: this warp's groups, over all the panels of the CTA, are one stream; the cursor runs a.pf groups
  // ahead of the group being multiplied and asks L2 for each group's two slot ranges with one bulk prefetch each, so
  // the loads below meet L2 latency, not DRAM latency (the kernel is latency-bound: 64 registers per thread leave room for
  // 8 entries per lane in flight)
  i64 pf_p = blockIdx.x;
  int pf_g = 0, pf_end = 0;
  bool pf_live = a.pf > 0 && pf_p < a.npanels;
  if (pf_live) {
    const int* q = a.sgrp_ptr + pf_p * NSLICE + warp;
    pf_g = __ldg(q);
    pf_end = __ldg(q + 1);
  }
  auto pf_step = [&]() {
    while (pf_live && pf_g >= pf_end) {
      pf_p += gridDim.x;
      if (pf_p >= a.npanels) {
        pf_live = false;
        break;
      }
      const int* q = a.sgrp_ptr + pf_p * NSLICE + warp;
      pf_g = __ldg(q);
      pf_end = __ldg(q + 1);
    }
    if (!pf_live) return;
    // a.pfc consecutive groups (contiguous slots) per request: fewer, larger prefetches
    const int gl = (pf_g + a.pfc <= pf_end) ? pf_g + a.pfc - 1 : pf_end - 1;
    if (lane == 0) {
      const uint2 d0 = __ldg(a.gdesc + pf_g);
      const uint2 d1 = __ldg(a.gdesc + gl);
      const unsigned ln = d1.x + (d1.y & 0xffffu) - d0.x;
      if (ln) {
        bulk_prefetch_l2(a.val + (i64)d0.x * 32, ln * 256u);
        bulk_prefetch_l2(a.lrow + (i64)d0.x * 32, ln * 64u);
      }
    }
    pf_g = gl + 1;
  };
  for (int q = 0; q < a.pf; q += a.pfc) pf_step();
  int pf_tick = 0;
  int buf = 0;
  for (i64 p = blockIdx.x; p < a.npanels; p += gridDim.x, buf ^= 1) {
    double* y_s = y_buf[buf];
#pragma unroll
    for (int q = 0; q < YPT; q++) {
      const int i = tid + q * 1024;
      if (i < a.R) y_s[i] = yr[q];
    }
    __syncthreads();  // panel p's y is complete; everybody is done with panel p - 1 (the other buffer)
    psell_fetch_y(a, p + gridDim.x, tid, yr);  // in flight while this panel is multiplied
    const int* sp = a.sgrp_ptr + p * NSLICE + warp;
    const int g0 = __ldg(sp), g1 = __ldg(sp + 1);
    uint2 d = (g0 < g1) ? __ldg(a.gdesc + g0) : make_uint2(0u, 0u);
    for (int g = g0; g < g1; g++) {
      if (++pf_tick == a.pfc) {
        pf_tick = 0;
        pf_step();
      }
      const uint2 dn = (g + 1 < g1) ? __ldg(a.gdesc + g + 1) : make_uint2(0u, 0u);  // the next group's descriptor, ahead of need
      const unsigned col = __ldg(a.gcol + (i64)g * 32 + lane);
      const int len = (int)(d.y & 0xffffu);
      const double* vp = a.val + (i64)d.x * 32 + lane;
      const unsigned short* rp = a.lrow + (i64)d.x * 32 + lane;
      double acc = 0.0;
      int j = 0;
      for (; j + 8 <= len; j += 8) {
        double x[8];
        unsigned r[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          x[u] = ldg_stream(vp + (j + u) * 32);
          r[u] = __ldg(rp + (j + u) * 32);
        }
#pragma unroll
        for (int u = 0; u < 8; u++) acc = __dadd_rn(acc, __dmul_rn(x[u], y_s[r[u]]));
      }
      if (j < len) {
        double x[8];
        unsigned r[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          const bool in = j + u < len;
          x[u] = in ? ldg_stream(vp + (j + u) * 32) : 0.0;
          r[u] = in ? (unsigned)__ldg(rp + (j + u) * 32) : (unsigned)a.R;
        }
#pragma unroll
        for (int u = 0; u < 8; u++) acc = __dadd_rn(acc, __dmul_rn(x[u], y_s[r[u]]));
      }
      if (!(d.y & 0x10000u)) {
        if (col != 0xffffu) c_s[col] = __dadd_rn(c_s[col], acc);
      } else {  // pieces of one column in several lanes: one lane after the other, in lane order
        for (int l = 0; l < 32; l++) {
          if (lane == l && col != 0xffffu) c_s[col] = __dadd_rn(c_s[col], acc);
          __syncwarp();
        }
      }
      d = dn;
    }
  }
  __syncthreads();
  double* prow = a.partial + (size_t)blockIdx.x * a.m;
  for (int j = tid; j < a.m; j += 1024) prow[j] = c_s[j];
}

// frees its device blocks when the build leaves, whichever way
struct Temps {
  std::vector<void*> ptrs;
  cudaError_t err = cudaSuccess;
  template <class T>
  T* get(size_t count) {
    void* p = nullptr;
    if (err == cudaSuccess) err = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
    if (err != cudaSuccess) return nullptr;
    ptrs.push_back(p);
    return (T*)p;
  }
  ~Temps() {
    for (void* p : ptrs) cudaFree(p);
  }
};

inline unsigned blocks_for(sysml_b200_ctx* ctx, i64 items, int threads) {
  i64 b = (items + threads - 1) / threads;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

}  // namespace

extern "C" {

int64_t sysml_b200_psell_segments(int64_t rows, int panel_rows) {
  if (rows < 0 || panel_rows < 1) return -1;
  return ((rows + panel_rows - 1) / panel_rows) * NSLICE + 1;
}

int sysml_b200_csr_to_psell(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, int64_t rows, int64_t cols,
                            int64_t nnz, int panel_rows, int32_t* sgrp_ptr, double** s_val, uint16_t** s_lrow, uint16_t** s_gcol,
                            uint64_t** s_gdesc, int64_t* n_slots, int64_t* n_groups) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 1 || nnz < 0 || nnz > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_DIM, "csr_to_psell: bad dimensions");
  if (cols > 65535) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_psell: more than 65535 columns (16-bit column ids)");
  if (panel_rows < 1 || panel_rows > 1024 * YPT) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_psell: panel_rows must be in [1, %d]", 1024 * YPT);
  if (!sgrp_ptr || !s_val || !s_lrow || !s_gcol || !s_gdesc || !n_slots || !n_groups) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_to_psell: null output");
  *s_val = nullptr;
  *s_lrow = nullptr;
  *s_gcol = nullptr;
  *s_gdesc = nullptr;
  *n_slots = 0;
  *n_groups = 0;
  const i64 npanels = (rows + panel_rows - 1) / panel_rows;
  if (npanels >= (1ll << 24)) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_psell: too many panels");
  const i64 nseg = npanels * NSLICE;
  SYSML_CUDA(ctx, cudaMemsetAsync(sgrp_ptr, 0, (size_t)(nseg + 1) * sizeof(int), ctx->stream));
  if (nnz == 0) return SYSML_OK;
  if (!rowptr || !colind || !val) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_to_psell: null input");
  cudaStream_t st = ctx->stream;
  const int n = (int)nnz;
  int pbits = 1;
  while ((1ll << pbits) < npanels) pbits++;
  int sbits = 1;
  while ((1ll << sbits) < nseg) sbits++;
  Temps T;
  u64* keys = T.get<u64>(nnz);
  u64* keys2 = T.get<u64>(nnz);
  double* vals = T.get<double>(nnz);
  int* runstart = T.get<int>(nnz);
  int* piece_start = T.get<int>(nnz + 1);
  int* d_np = T.get<int>(1);
  int* seg_count = T.get<int>(nseg + 1);
  int* seg_ptr = T.get<int>(nseg + 1);
  int* grp_count = T.get<int>(nseg + 1);
  if (T.err != cudaSuccess) return sysml_set_error(ctx, SYSML_E_OOM, "csr_to_psell: temporaries for %lld entries: %s", (long long)nnz, cudaGetErrorString(T.err));
  // 1. sort the entries by (panel, column, row)
  psell_keys_kernel<<<blocks_for(ctx, rows, 256), 256, 0, st>>>(rowptr, colind, rows, panel_rows, keys);
  SYSML_LAUNCH_CHECK(ctx, "psell_keys_kernel");
  size_t b1 = 0, b2 = 0, b3 = 0, b4 = 0, b5 = 0;
  cub::CountingInputIterator<int> iota(0);
  cub::TransformInputIterator<int, RunHead, cub::CountingInputIterator<int>> heads(iota, RunHead{keys2});
  cub::TransformInputIterator<bool, PieceHead, cub::CountingInputIterator<int>> pflags(iota, PieceHead{runstart});
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, b1, keys, keys2, val, vals, n, 0, 32 + pbits, st));
  SYSML_CUDA(ctx, cub::DeviceScan::InclusiveScan(nullptr, b2, heads, runstart, cub::Max(), n, st));
  SYSML_CUDA(ctx, cub::DeviceSelect::Flagged(nullptr, b3, iota, pflags, piece_start, d_np, n, st));
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, b4, seg_count, seg_ptr, (int)(nseg + 1), st));
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, b5, (const u64*)nullptr, (u64*)nullptr, (const int*)nullptr, (int*)nullptr, n, 0, 8 + sbits, st));
  size_t tb = b1;
  for (size_t b : {b2, b3, b4, b5}) tb = b > tb ? b : tb;
  void* tmp = T.get<char>(tb + 256);
  if (T.err != cudaSuccess) return sysml_set_error(ctx, SYSML_E_OOM, "csr_to_psell: sort scratch: %s", cudaGetErrorString(T.err));
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, b1, keys, keys2, val, vals, n, 0, 32 + pbits, st));
  // 2. runs of equal (panel, column), cut into pieces of at most LMAX entries
  SYSML_CUDA(ctx, cub::DeviceScan::InclusiveScan(tmp, b2, heads, runstart, cub::Max(), n, st));
  SYSML_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, b3, iota, pflags, piece_start, d_np, n, st));
  int np = 0;
  SYSML_CUDA(ctx, cudaMemcpyAsync(&np, d_np, sizeof(int), cudaMemcpyDeviceToHost, st));
  SYSML_CUDA(ctx, cudaStreamSynchronize(st));
  SYSML_CUDA(ctx, cudaMemcpyAsync(piece_start + np, &n, sizeof(int), cudaMemcpyHostToDevice, st));
  // 3. pieces sorted by (panel, slice, longest first); `keys` is free again and holds the sort keys
  u64* skey = keys;
  u64* skey2 = keys + ((size_t)np + 1);
  int* sidx = T.get<int>(np);
  int* sidx2 = T.get<int>(np);
  if (T.err != cudaSuccess) return sysml_set_error(ctx, SYSML_E_OOM, "csr_to_psell: piece index: %s", cudaGetErrorString(T.err));
  if (2 * ((size_t)np + 1) > (size_t)nnz) {  // mostly short runs: the two key arrays do not fit in the old one
    skey = T.get<u64>(np);
    skey2 = T.get<u64>(np);
    if (T.err != cudaSuccess) return sysml_set_error(ctx, SYSML_E_OOM, "csr_to_psell: piece keys: %s", cudaGetErrorString(T.err));
  }
  SYSML_CUDA(ctx, cudaMemsetAsync(seg_count, 0, (size_t)(nseg + 1) * sizeof(int), st));
  const int slice_cols = (int)((cols + NSLICE - 1) / NSLICE);
  psell_piece_keys_kernel<<<blocks_for(ctx, np, 256), 256, 0, st>>>(keys2, piece_start, np, slice_cols, skey, sidx, seg_count);
  SYSML_LAUNCH_CHECK(ctx, "psell_piece_keys_kernel");
  SYSML_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, b5, skey, skey2, sidx, sidx2, np, 0, 8 + sbits, st));
  // 4. groups of 32 pieces per (panel, slice); their lengths and offsets
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, b4, seg_count, seg_ptr, (int)(nseg + 1), st));
  psell_group_count_kernel<<<blocks_for(ctx, nseg, 256), 256, 0, st>>>(seg_count, nseg, grp_count);
  SYSML_LAUNCH_CHECK(ctx, "psell_group_count_kernel");
  SYSML_CUDA(ctx, cudaMemsetAsync(grp_count + nseg, 0, sizeof(int), st));
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, b4, grp_count, sgrp_ptr, (int)(nseg + 1), st));
  int ng = 0;
  SYSML_CUDA(ctx, cudaMemcpyAsync(&ng, sgrp_ptr + nseg, sizeof(int), cudaMemcpyDeviceToHost, st));
  SYSML_CUDA(ctx, cudaStreamSynchronize(st));
  int* glen = T.get<int>((size_t)ng + 1);
  int* gseg = T.get<int>((size_t)ng + 1);
  i64* goff = T.get<i64>((size_t)ng + 1);
  size_t b6 = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, b6, (const int*)nullptr, (i64*)nullptr, ng + 1, st));
  void* tmp6 = b6 <= tb ? tmp : T.get<char>(b6 + 256);
  if (T.err != cudaSuccess) return sysml_set_error(ctx, SYSML_E_OOM, "csr_to_psell: group tables: %s", cudaGetErrorString(T.err));
  psell_group_len_kernel<<<blocks_for(ctx, nseg, 128), 128, 0, st>>>(seg_ptr, sgrp_ptr, nseg, sidx2, piece_start, glen, gseg);
  SYSML_LAUNCH_CHECK(ctx, "psell_group_len_kernel");
  SYSML_CUDA(ctx, cudaMemsetAsync(glen + ng, 0, sizeof(int), st));
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp6, b6, glen, goff, ng + 1, st));
  i64 total = 0;
  SYSML_CUDA(ctx, cudaMemcpyAsync(&total, goff + ng, sizeof(i64), cudaMemcpyDeviceToHost, st));
  SYSML_CUDA(ctx, cudaStreamSynchronize(st));
  if (total >= (1ll << 32)) return sysml_set_error(ctx, SYSML_E_LIMIT, "csr_to_psell: more than 2^37 slots");
  // 5. the result arrays (the caller releases them with sysml_b200_free) and the fill
  const i64 slots = total * 32;
  cudaError_t e = cudaMalloc((void**)s_val, (size_t)(slots ? slots : 1) * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc((void**)s_lrow, (size_t)(slots ? slots : 1) * sizeof(uint16_t));
  if (e == cudaSuccess) e = cudaMalloc((void**)s_gcol, (size_t)(ng ? ng : 1) * 32 * sizeof(uint16_t));
  if (e == cudaSuccess) e = cudaMalloc((void**)s_gdesc, (size_t)(ng ? ng : 1) * sizeof(uint64_t));
  if (e != cudaSuccess) {
    for (void* p : {(void*)*s_val, (void*)*s_lrow, (void*)*s_gcol, (void*)*s_gdesc})
      if (p) cudaFree(p);
    *s_val = nullptr;
    *s_lrow = nullptr;
    *s_gcol = nullptr;
    *s_gdesc = nullptr;
    return sysml_set_error(ctx, SYSML_E_OOM, "csr_to_psell: %lld slots: %s", (long long)slots, cudaGetErrorString(e));
  }
  psell_fill_kernel<<<blocks_for(ctx, (i64)ng * 32, 256), 256, 0, st>>>(keys2, vals, piece_start, sidx2, seg_ptr, sgrp_ptr, gseg, glen, goff, ng, panel_rows,
                                                                       *s_val, *s_lrow, *s_gcol, (uint2*)*s_gdesc);
  SYSML_LAUNCH_CHECK(ctx, "psell_fill_kernel");
  SYSML_CUDA(ctx, cudaStreamSynchronize(st));  // the temporaries are freed on return
  *n_slots = slots;
  *n_groups = ng;
  return SYSML_OK;
}

int sysml_b200_matmult_psell_vm(sysml_b200_ctx* ctx, const double* s_val, const uint16_t* s_lrow, const uint16_t* s_gcol, const uint64_t* s_gdesc,
                                const int32_t* sgrp_ptr, int64_t rows, int64_t cols, int panel_rows, const double* y, const double* w, int w_op,
                                double* out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_psell_vm: negative dimension");
  if (cols == 0) return SYSML_OK;
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_psell_vm: out is null");
  if (rows == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)cols * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (cols > 65535 || panel_rows < 1 || panel_rows > 1024 * YPT) return sysml_set_error(ctx, SYSML_E_LIMIT, "matmult_psell_vm: cols <= 65535 and panel_rows <= %d", 1024 * YPT);
  if (w_op < 0 || w_op > 2 || (w_op && !w)) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_psell_vm: bad weight operation");
  if (!sgrp_ptr || !y) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_psell_vm: null pointer");
  const size_t smem = ((size_t)cols + 2 * ((size_t)panel_rows + 1)) * sizeof(double);
  if (smem > (size_t)ctx->max_smem_optin)
    return sysml_set_error(ctx, SYSML_E_LIMIT, "matmult_psell_vm: cols + 2 * (panel_rows + 1) = %lld doubles do not fit in shared memory (%d bytes)",
                           (long long)(cols + 2 * (panel_rows + 1)), ctx->max_smem_optin);
  SYSML_CUDA(ctx, cudaFuncSetAttribute(psell_vm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  PsellArgs a{};
  a.val = s_val; a.lrow = s_lrow; a.gcol = s_gcol; a.gdesc = (const uint2*)s_gdesc; a.sgrp_ptr = sgrp_ptr; a.y = y; a.w = w; a.wop = w_op;
  a.n = rows; a.m = (int)cols; a.R = panel_rows;
  a.npanels = (rows + panel_rows - 1) / panel_rows;
  a.pf = 2;  // measured at C3: 0 -> 1.42 ms, 1..2 -> 1.20, 3 -> 1.31, 4 -> 1.37, 8 and more -> 1.8 (the prefetched lines push each other out of L2)
  if (const char* e = getenv("SYSML_B200_PSELL_PREFETCH")) a.pf = atoi(e);  // A/B switch for measurements
  if (a.pf < 0 || a.pf > 64) a.pf = 0;
  a.pfc = 1;  // groups per prefetch request (A/B switch: distance 2 with 2 groups per request measured the same 1.20 ms; 4 per request 1.75)
  if (const char* e = getenv("SYSML_B200_PSELL_PREFETCH_CHUNK")) a.pfc = atoi(e);
  if (a.pfc < 1 || a.pfc > 16) a.pfc = 1;
  i64 grid = ctx->sm_count;
  if (grid > a.npanels) grid = a.npanels;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, (size_t)grid * cols * sizeof(double), &ws);
  if (rc) return rc;
  a.partial = (double*)ws;
  psell_vm_kernel<<<(unsigned)grid, 1024, smem, ctx->stream>>>(a);
  SYSML_LAUNCH_CHECK(ctx, "psell_vm_kernel");
  return sysml_reduce_partial_rows(ctx, a.partial, (int)grid, (int)cols, out);
}

int sysml_b200_mmchain_csr_psell(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, const double* s_val,
                                 const uint16_t* s_lrow, const uint16_t* s_gcol, const uint64_t* s_gdesc, const int32_t* sgrp_ptr, int panel_rows,
                                 const double* v, const double* w, double* tmp_n, double* out, int64_t n, int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  if (chain_type < 0 || chain_type > 2) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr_psell: unknown chain type %d", chain_type);
  if (chain_type != SYSML_CHAIN_XtXv && n > 0 && !w) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr_psell: chain type %d needs w", chain_type);
  if (n > 0 && !tmp_n) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr_psell: the n-vector temporary is null");
  // X %*% v is a gather over the row-ordered copy; the weights are applied while the panel of the result is parked
  int rc = sysml_b200_matmult_csr_dense(ctx, rowptr, colind, val, v, tmp_n, n, m, m, 1, 0);
  if (rc) return rc;
  return sysml_b200_matmult_psell_vm(ctx, s_val, s_lrow, s_gcol, s_gdesc, sgrp_ptr, n, m, panel_rows, tmp_n, chain_type ? w : nullptr, chain_type, out);
}

}  // extern "C"
