// sparse_cellwise.cu -- sparse-safe cellwise operations whose operand AND result stay CSR.
//
// Reference (CPU twin; the reference's GPU twin densifies every sparse operand of a cellwise operator,
// LibMatrixCUDA.java:1360-1383): LibMatrixBincell.safeBinaryScalar's SPARSE <- SPARSE branch
// (LibMatrixBincell.java:938-995), safeBinaryMVSparse (:338-440), safeBinaryMMSparseDenseSkip (:760-790),
// safeBinaryMMSparseSparse with mergeForSparseBinary / appendLeft / appendRight (:597-660, 1279-1340), and the
// sparse-safe unary path of MatrixBlock.unaryOperations.  In all of them the value function is evaluated at
// the stored cells only (for two CSR operands: at the union of the stored cells, with inner-join semantics for
// Multiply), and SparseRow.append drops results that are zero -- so the result holds no explicit zeros and
// its nnz equals MatrixBlock.recomputeNonZeros of the dense result.  Which (operator, constant) pairs may take
// this path is the caller's decision (BinaryOperator.sparseSafe, BinaryOperator.java:60-66; ScalarOperator /
// RightScalarOperator / LeftScalarOperator constructors; LibMatrixBincell.isSparseSafeDivide :185-189); the
// scalar and unary entry points additionally refuse an operator whose image of zero is not zero.
//
// Protocol (CSRPointer.allocateForDgeam's, CSRPointer.java:227-299): every entry point is called twice with the
// same operands -- first with c_colind == c_val == NULL: the symbolic pass writes c_rowptr and *c_nnz; the
// caller allocates the two arrays (through its memory manager) and calls again: the numeric pass fills them.
// Nothing is allocated by the library.
//
// Kernels: one warp owns 32 consecutive rows and walks their stored cells as ONE contiguous run, 32 cells per
// step (coalesced 4- and 8-byte loads); a lane finds the row of its cell by a 5-step search over the 32 row
// starts held one per lane (shuffles), results that survive are ranked inside their row from one ballot of the
// run starts (rows ascend with the cell index), and the per-row write cursors live in shared memory.  Two CSR operands: one thread per row merges
// the two sorted runs (rows are ~10 cells at the C3 shape).
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "cellwise_rt.cuh"

namespace {

enum { M_SCALAR_R = 0, M_SCALAR_L = 1, M_UNARY = 2, M_DENSE = 3, M_COL = 4, M_ROW = 5 };

using sysml_rt::binop_rt;
using sysml_rt::unop_rt;

struct CellArgs {
  const int* rp;
  const int* ci;
  const double* va;
  i64 rows, cols;
  int op;
  int left;  // the other operand is the LEFT argument of the value function
  double scalar;
  const double* B;
};

// OPT: the operator when it is known at compile time (Multiply, by far the most common sparse-safe one: inlined), -1 otherwise
template <int MODE, int OPT>
__device__ __forceinline__ double cell_value(const CellArgs& a, double x, i64 row, int col) {
  if (MODE == M_UNARY) return unop_rt(a.op, x);
  double y;
  if (MODE == M_SCALAR_R || MODE == M_SCALAR_L) y = a.scalar;
  else if (MODE == M_DENSE) y = a.B[row * a.cols + col];
  else if (MODE == M_COL) y = a.B[row];
  else y = a.B[col];
  if (OPT == SYSML_OP_MULTIPLY) return __dmul_rn(x, y);
  const bool left = (MODE == M_SCALAR_L) || (MODE >= M_DENSE && a.left);
  return left ? binop_rt(a.op, y, x) : binop_rt(a.op, x, y);
}

constexpr int CW_WARPS = 8;

// COUNT: cnt[r] = number of non-zero results in row r.   FILL: results written at crp[r] + rank inside the row.
template <int MODE, bool FILL, int OPT>
__global__ void __launch_bounds__(CW_WARPS * 32) csr_cell_kernel(CellArgs a, int* __restrict__ cnt, const int* __restrict__ crp,
                                                                  int* __restrict__ cci, double* __restrict__ cva) {
  __shared__ int cursor[CW_WARPS][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const unsigned full = 0xffffffffu, lt = (1u << lane) - 1u;
  const i64 groups = (a.rows + 31) / 32;
  const i64 nwarps = (i64)gridDim.x * CW_WARPS;
  for (i64 g = (i64)blockIdx.x * CW_WARPS + w; g < groups; g += nwarps) {
    const i64 r = g * 32 + lane;
    const int s = a.rp[r < a.rows ? r : a.rows];  // start of this lane's row (rows past the end start at nnz)
    const i64 last = g * 32 + 32;
    const int end = a.rp[last < a.rows ? last : a.rows];
    const int begin = __shfl_sync(full, s, 0);
    cursor[w][lane] = FILL ? ((r < a.rows) ? crp[r] : 0) : 0;
    __syncwarp();
    // the warp's next group of rows is asked into L2 now (one bulk prefetch per array): with ~10 cells per row a lane has
    // only two loads in flight, and this turns their DRAM latency into L2 latency
    if (lane == 0 && g + nwarps < groups) {
      const i64 rn = (g + nwarps) * 32, rl = (rn + 32 < a.rows) ? rn + 32 : a.rows;
      const i64 kb = a.rp[rn], ke = a.rp[rl];
      if (ke > kb) {
        // whole 16-byte pieces inside the arrays: the start rounds down, the end rounds down too (it is only a hint)
        const i64 c0 = kb & ~(i64)3, c1 = ke & ~(i64)3;
        const i64 v0 = kb & ~(i64)1, v1 = ke & ~(i64)1;
        if (c1 > c0) bulk_prefetch_l2(a.ci + c0, (uint32_t)((c1 - c0) * 4));
        if (v1 > v0) bulk_prefetch_l2(a.va + v0, (uint32_t)((v1 - v0) * 8));
      }
    }
    // the next step's cell is loaded before this step's is processed: two loads per lane in flight
    int c_next = 0;
    double x_next = 0.0;
    if (begin + lane < end) {
      c_next = a.ci[begin + lane];
      x_next = a.va[begin + lane];
    }
    for (int k0 = begin; k0 < end; k0 += 32) {
      const int k = k0 + lane;
      const bool active = k < end;
      const int c = c_next;
      const double x = x_next;
      if (k + 32 < end) {
        c_next = a.ci[k + 32];
        x_next = a.va[k + 32];
      }
      // the row of cell k: the largest j with start[j] <= k (equal starts are empty rows; the last one holds the cell)
      int j = 0;
#pragma unroll
      for (int step = 16; step >= 1; step >>= 1) {
        const int t = __shfl_sync(full, s, j + step);
        if (t <= k) j += step;
      }
      double v = 0.0;
      if (active) v = cell_value<MODE, OPT>(a, x, g * 32 + j, c);
      const bool keep = active && (v != 0.0);  // SparseRow.append drops zeros; NaN is kept
      // the lanes of one row are neighbours (rows ascend with k): a row's run starts where j changes
      const int jp = __shfl_up_sync(full, j, 1);
      const unsigned starts = __ballot_sync(full, lane == 0 || jp != j || !active) | 1u;
      const int first = 31 - __clz(starts & (lt | (1u << lane)));          // first lane of this lane's run
      const unsigned after = starts & ~(lt | (1u << lane));                // run starts to the right
      const unsigned run = (after ? ((1u << (__ffs(after) - 1)) - 1u) : full) & ~((1u << first) - 1u);
      const unsigned mine = run & __ballot_sync(full, keep);
      const int base = active ? cursor[w][j] : 0;
      __syncwarp();
      if (active && lane == first) cursor[w][j] = base + __popc(mine);  // one leader per row
      if (FILL && keep) {
        const int pos = base + __popc(mine & lt);
        cci[pos] = c;
        cva[pos] = v;
      }
      __syncwarp();
    }
    if (!FILL && r < a.rows) cnt[r] = cursor[w][lane];
    __syncwarp();
  }
}

// Numeric pass when the symbolic pass found that nothing is dropped (the usual case: X * 0.5, abs(X), X * row vector without
// zeros): the result has the operand's structure, so the cells are independent of their rows -- one flat, fully coalesced
// pass that copies the column indices and writes the values.  (Modes that need the row of a cell keep the row walk.)
template <int MODE, int OPT>
__global__ void __launch_bounds__(256) csr_cell_flat_kernel(CellArgs a, i64 nnz, int* __restrict__ cci, double* __restrict__ cva) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
    const int c = a.ci[k];
    cci[k] = c;
    cva[k] = cell_value<MODE, OPT>(a, ldg_stream(a.va + k), 0, c);
  }
}

// two CSR operands: sorted merge per row.  Multiply: inner join (mergeForSparseBinary :1282-1296); every other
// operator: outer join with op(a,0) / op(0,b) at the cells only one side stores (:1298-1321)
template <bool FILL>
__global__ void csr_merge_kernel(const int* __restrict__ arp, const int* __restrict__ aci, const double* __restrict__ ava,
                                 const int* __restrict__ brp, const int* __restrict__ bci, const double* __restrict__ bva, int op, i64 rows,
                                 int* __restrict__ cnt_or_rp, int* __restrict__ cci, double* __restrict__ cva) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    int i = arp[r], j = brp[r];
    const int ie = arp[r + 1], je = brp[r + 1];
    // the inner join is taken only where BOTH rows hold cells; a row that meets an empty one goes through appendLeft /
    // appendRight like any other operator (LibMatrixBincell.java:622-635) -- visible only for stored NaN / Inf
    const bool inner = (op == SYSML_OP_MULTIPLY) && i < ie && j < je;
    int n = 0;
    const int base = FILL ? cnt_or_rp[r] : 0;
    while (inner ? (i < ie && j < je) : (i < ie || j < je)) {
      const int ca = (i < ie) ? aci[i] : 0x7fffffff, cb = (j < je) ? bci[j] : 0x7fffffff;
      double v;
      int c;
      bool have = true;
      if (ca == cb) {
        v = binop_rt(op, ava[i], bva[j]);
        c = ca;
        i++;
        j++;
      } else if (ca < cb) {
        have = !inner;
        v = have ? binop_rt(op, ava[i], 0.0) : 0.0;
        c = ca;
        i++;
      } else {
        have = !inner;
        v = have ? binop_rt(op, 0.0, bva[j]) : 0.0;
        c = cb;
        j++;
      }
      if (have && v != 0.0) {
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

__global__ void zero_image_kernel(int mode, int op, double scalar, double* out) {
  if (mode == M_UNARY) *out = unop_rt(op, 0.0);
  else *out = (mode == M_SCALAR_L) ? binop_rt(op, scalar, 0.0) : binop_rt(op, 0.0, scalar);
}

// A vector operand is broadcast over cells the CSR operand does not store, and the reference evaluates the operator there
// too (op(0, v2) for every gap of a row, LibMatrixBincell.java:383-398 for a column vector, :411-431 for a row vector) --
// except under Multiply, which skips the gaps of a row vector altogether and the rows that are empty or meet v2 == 0.
// The stored-cells result is the reference's result iff all those gap values are zero: flag = 1 otherwise.
__global__ void vector_gap_kernel(CellArgs a, int col_vector, int* flag) {
  const i64 len = col_vector ? a.rows : a.cols;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  int bad = 0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) {
    const double y = a.B[i];
    if (a.op == SYSML_OP_MULTIPLY && col_vector && (a.rp[i + 1] == a.rp[i] || y == 0.0)) continue;
    const double z = a.left ? binop_rt(a.op, y, 0.0) : binop_rt(a.op, 0.0, y);
    if (!(z == 0.0)) bad = 1;
  }
  if (bad) atomicOr(flag, 1);
}

inline unsigned grid_for(sysml_b200_ctx* ctx, i64 items, int per_block) {
  i64 b = (items + per_block - 1) / per_block;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}
inline size_t padded(size_t bytes) { return ((bytes + 255) / 256) * 256; }

int fetch(sysml_b200_ctx* ctx, void* h, const void* d, size_t bytes) {
  SYSML_CUDA(ctx, cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SYSML_OK;
}

template <int MODE>
int launch_cell(sysml_b200_ctx* ctx, const CellArgs& a, bool fill, int* cnt, const int* crp, int* cci, double* cva) {
  const unsigned g = grid_for(ctx, (a.rows + 31) / 32, CW_WARPS);
  const bool mult = MODE != M_UNARY && a.op == SYSML_OP_MULTIPLY;
  if (fill) {
    if (mult) csr_cell_kernel<MODE, true, SYSML_OP_MULTIPLY><<<g, CW_WARPS * 32, 0, ctx->stream>>>(a, nullptr, crp, cci, cva);
    else csr_cell_kernel<MODE, true, -1><<<g, CW_WARPS * 32, 0, ctx->stream>>>(a, nullptr, crp, cci, cva);
  } else {
    if (mult) csr_cell_kernel<MODE, false, SYSML_OP_MULTIPLY><<<g, CW_WARPS * 32, 0, ctx->stream>>>(a, cnt, nullptr, nullptr, nullptr);
    else csr_cell_kernel<MODE, false, -1><<<g, CW_WARPS * 32, 0, ctx->stream>>>(a, cnt, nullptr, nullptr, nullptr);
  }
  SYSML_LAUNCH_CHECK(ctx, "csr_cell_kernel");
  return SYSML_OK;
}

int launch_cell_mode(sysml_b200_ctx* ctx, int mode, const CellArgs& a, bool fill, int* cnt, const int* crp, int* cci, double* cva) {
  switch (mode) {
    case M_SCALAR_R: return launch_cell<M_SCALAR_R>(ctx, a, fill, cnt, crp, cci, cva);
    case M_SCALAR_L: return launch_cell<M_SCALAR_L>(ctx, a, fill, cnt, crp, cci, cva);
    case M_UNARY: return launch_cell<M_UNARY>(ctx, a, fill, cnt, crp, cci, cva);
    case M_DENSE: return launch_cell<M_DENSE>(ctx, a, fill, cnt, crp, cci, cva);
    case M_COL: return launch_cell<M_COL>(ctx, a, fill, cnt, crp, cci, cva);
    case M_ROW: return launch_cell<M_ROW>(ctx, a, fill, cnt, crp, cci, cva);
  }
  return sysml_set_error(ctx, SYSML_E_INVALID, "csr cellwise: bad mode %d", mode);
}

// counts (rows + 1 ints in the scratch block, last one zero) -> c_rowptr by an exclusive scan; *c_nnz = c_rowptr[rows]
int scan_counts(sysml_b200_ctx* ctx, const int* cnt, void* tmp, size_t tmp_bytes, i64 rows, int32_t* c_rowptr, int64_t* c_nnz) {
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, cnt, c_rowptr, (int)(rows + 1), ctx->stream));
  int nnz = 0;
  int rc = fetch(ctx, &nnz, c_rowptr + rows, sizeof(int));
  if (rc) return rc;
  *c_nnz = nnz;
  return SYSML_OK;
}

struct Scratch {
  int* cnt;
  void* tmp;
  size_t tmp_bytes;
  double* zero;
};
int get_scratch(sysml_b200_ctx* ctx, i64 rows, Scratch* s) {
  size_t scan_bytes = 0;
  SYSML_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const int*)nullptr, (int*)nullptr, (int)(rows + 1), ctx->stream));
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, padded((size_t)(rows + 1) * 4) + padded(scan_bytes) + 768, &ws);
  if (rc) return rc;
  char* p = (char*)ws;
  s->zero = (double*)p;
  p += 256;
  s->cnt = (int*)p;
  p += padded((size_t)(rows + 1) * 4);
  s->tmp = p;
  s->tmp_bytes = scan_bytes;
  return SYSML_OK;
}

int check_dims(sysml_b200_ctx* ctx, const char* who, int64_t rows, int64_t cols, const void* rowptr, int32_t* c_rowptr, int64_t* c_nnz,
               int32_t* c_colind, double* c_val) {
  if (rows < 0 || cols < 0 || rows >= 0x7fffffffLL || cols > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_DIM, "%s: bad dimensions", who);
  if (!c_rowptr || !c_nnz) return sysml_set_error(ctx, SYSML_E_INVALID, "%s: null c_rowptr / c_nnz", who);
  if ((c_colind == nullptr) != (c_val == nullptr)) return sysml_set_error(ctx, SYSML_E_INVALID, "%s: c_colind and c_val must both be NULL (symbolic pass) or both be set (numeric pass)", who);
  if (rows > 0 && !rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "%s: null input", who);
  return SYSML_OK;
}

// both passes of the one-CSR-operand operations
int run_cell(sysml_b200_ctx* ctx, const char* who, int mode, const CellArgs& a, int32_t* c_rowptr, int32_t* c_colind, double* c_val,
             int64_t* c_nnz) {
  const bool fill = c_colind != nullptr;
  if (a.rows == 0) {
    if (!fill) {
      SYSML_CUDA(ctx, cudaMemsetAsync(c_rowptr, 0, sizeof(int), ctx->stream));
      *c_nnz = 0;
    }
    return SYSML_OK;
  }
  if (fill) {
    if (mode == M_SCALAR_R || mode == M_SCALAR_L || mode == M_UNARY || mode == M_ROW) {
      // same number of cells in the result as in the operand <=> every row kept all of its cells
      int na = 0, nc = 0;
      int rc = fetch(ctx, &na, a.rp + a.rows, sizeof(int));
      if (rc) return rc;
      rc = fetch(ctx, &nc, c_rowptr + a.rows, sizeof(int));
      if (rc) return rc;
      if (na == nc) {
        if (na == 0) return SYSML_OK;
        const unsigned g = grid_for(ctx, na, 256 * 4);
        const bool mult = mode != M_UNARY && a.op == SYSML_OP_MULTIPLY;
#define FLAT(M)                                                                                                              \
  case M:                                                                                                                    \
    if (mult) csr_cell_flat_kernel<M, SYSML_OP_MULTIPLY><<<g, 256, 0, ctx->stream>>>(a, (i64)na, c_colind, c_val);          \
    else csr_cell_flat_kernel<M, -1><<<g, 256, 0, ctx->stream>>>(a, (i64)na, c_colind, c_val);                              \
    break;
        switch (mode) {
          FLAT(M_SCALAR_R)
          FLAT(M_SCALAR_L)
          FLAT(M_UNARY)
          FLAT(M_ROW)
        }
#undef FLAT
        SYSML_LAUNCH_CHECK(ctx, "csr_cell_flat_kernel");
        return SYSML_OK;
      }
    }
    return launch_cell_mode(ctx, mode, a, true, nullptr, c_rowptr, c_colind, c_val);
  }
  Scratch s;
  int rc = get_scratch(ctx, a.rows, &s);
  if (rc) return rc;
  if (mode <= M_UNARY) {  // the operator must map zero to zero, else the cells that are not stored would change too
    zero_image_kernel<<<1, 1, 0, ctx->stream>>>(mode, a.op, a.scalar, s.zero);
    SYSML_LAUNCH_CHECK(ctx, "zero_image_kernel");
    double z = 0.0;
    rc = fetch(ctx, &z, s.zero, sizeof(double));
    if (rc) return rc;
    if (!(z == 0.0))
      return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "%s: operator %d is not sparse-safe here (its image of 0 is %g): use the dense form", who, a.op, z);
  }
  if ((mode == M_COL || mode == M_ROW) && !(mode == M_ROW && a.op == SYSML_OP_MULTIPLY)) {
    int* flag = (int*)s.zero;
    SYSML_CUDA(ctx, cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream));
    const i64 len = mode == M_COL ? a.rows : a.cols;
    vector_gap_kernel<<<grid_for(ctx, len, 256), 256, 0, ctx->stream>>>(a, mode == M_COL ? 1 : 0, flag);
    SYSML_LAUNCH_CHECK(ctx, "vector_gap_kernel");
    int bad = 0;
    rc = fetch(ctx, &bad, flag, sizeof(int));
    if (rc) return rc;
    if (bad)
      return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "%s: operator %d is not sparse-safe for this vector (it maps a cell that is not stored to a non-zero): use the dense form", who, a.op);
  }
  SYSML_CUDA(ctx, cudaMemsetAsync(s.cnt + a.rows, 0, sizeof(int), ctx->stream));
  rc = launch_cell_mode(ctx, mode, a, false, s.cnt, nullptr, nullptr, nullptr);
  if (rc) return rc;
  return scan_counts(ctx, s.cnt, s.tmp, s.tmp_bytes, a.rows, c_rowptr, c_nnz);
}

}  // namespace

extern "C" {

int sysml_b200_csr_scalar_op(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, int64_t rows,
                             int64_t cols, double scalar, int op, int scalar_left, int32_t* c_rowptr, int32_t* c_colind, double* c_val,
                             int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  int rc = check_dims(ctx, "csr_scalar_op", rows, cols, rowptr, c_rowptr, c_nnz, c_colind, c_val);
  if (rc) return rc;
  if (op < 0 || op > SYSML_OP_INTDIV) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_scalar_op: unknown opcode %d", op);
  CellArgs a{rowptr, colind, val, rows, cols, op, 0, scalar, nullptr};
  return run_cell(ctx, "csr_scalar_op", scalar_left ? M_SCALAR_L : M_SCALAR_R, a, c_rowptr, c_colind, c_val, c_nnz);
}

int sysml_b200_csr_unary_op(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, int64_t rows,
                            int64_t cols, int un_op, int32_t* c_rowptr, int32_t* c_colind, double* c_val, int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  int rc = check_dims(ctx, "csr_unary_op", rows, cols, rowptr, c_rowptr, c_nnz, c_colind, c_val);
  if (rc) return rc;
  if (un_op < 0 || un_op > SYSML_UN_TANH) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_unary_op: unknown opcode %d", un_op);
  CellArgs a{rowptr, colind, val, rows, cols, un_op, 0, 0.0, nullptr};
  return run_cell(ctx, "csr_unary_op", M_UNARY, a, c_rowptr, c_colind, c_val, c_nnz);
}

int sysml_b200_csr_dense_op(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val, int64_t rows,
                            int64_t cols, const double* B, int bcast, int op, int dense_left, int32_t* c_rowptr, int32_t* c_colind,
                            double* c_val, int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  int rc = check_dims(ctx, "csr_dense_op", rows, cols, rowptr, c_rowptr, c_nnz, c_colind, c_val);
  if (rc) return rc;
  if (op < 0 || op > SYSML_OP_INTDIV) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_dense_op: unknown opcode %d", op);
  if (bcast != SYSML_BCAST_NONE && bcast != SYSML_BCAST_COL && bcast != SYSML_BCAST_ROW)
    return sysml_set_error(ctx, SYSML_E_INVALID, "csr_dense_op: bad broadcast code %d", bcast);
  if (rows > 0 && cols > 0 && !B) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_dense_op: null dense operand");
  // a full dense operand: the reference's sparse result comes from safeBinaryMMSparseDenseSkip (stored cells only), which it
  // takes for Multiply and for Divide by an operand without zeros; And gives the same cells under safeBinaryMMGeneric
  if (bcast == SYSML_BCAST_NONE && op != SYSML_OP_MULTIPLY && op != SYSML_OP_AND && !(op == SYSML_OP_DIVIDE && !dense_left))
    return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "csr_dense_op: operator %d has no sparse form against a dense matrix%s: use the dense form", op,
                           dense_left ? " on its left" : "");
  CellArgs a{rowptr, colind, val, rows, cols, op, dense_left ? 1 : 0, 0.0, B};
  return run_cell(ctx, "csr_dense_op", bcast == SYSML_BCAST_NONE ? M_DENSE : (bcast == SYSML_BCAST_COL ? M_COL : M_ROW), a, c_rowptr, c_colind,
                  c_val, c_nnz);
}

int sysml_b200_csr_csr_op(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val,
                          const int32_t* b_rowptr, const int32_t* b_colind, const double* b_val, int64_t rows, int64_t cols, int op,
                          int32_t* c_rowptr, int32_t* c_colind, double* c_val, int64_t* c_nnz) {
  SYSML_ENTER(ctx);
  int rc = check_dims(ctx, "csr_csr_op", rows, cols, a_rowptr, c_rowptr, c_nnz, c_colind, c_val);
  if (rc) return rc;
  if (rows > 0 && !b_rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_csr_op: null input");
  if (op < 0 || op > SYSML_OP_INTDIV) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_csr_op: unknown opcode %d", op);
  const bool fill = c_colind != nullptr;
  if (rows == 0) {
    if (!fill) {
      SYSML_CUDA(ctx, cudaMemsetAsync(c_rowptr, 0, sizeof(int), ctx->stream));
      *c_nnz = 0;
    }
    return SYSML_OK;
  }
  const unsigned g = grid_for(ctx, rows, 128);
  if (fill) {
    csr_merge_kernel<true><<<g, 128, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, b_rowptr, b_colind, b_val, op, rows, c_rowptr, c_colind, c_val);
    SYSML_LAUNCH_CHECK(ctx, "csr_merge_kernel<fill>");
    return SYSML_OK;
  }
  Scratch s;
  rc = get_scratch(ctx, rows, &s);
  if (rc) return rc;
  zero_image_kernel<<<1, 1, 0, ctx->stream>>>(M_SCALAR_R, op, 0.0, s.zero);  // op(0, 0)
  SYSML_LAUNCH_CHECK(ctx, "zero_image_kernel");
  double z = 0.0;
  rc = fetch(ctx, &z, s.zero, sizeof(double));
  if (rc) return rc;
  if (!(z == 0.0)) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "csr_csr_op: operator %d is not sparse-safe (0 op 0 = %g): use the dense form", op, z);
  SYSML_CUDA(ctx, cudaMemsetAsync(s.cnt + rows, 0, sizeof(int), ctx->stream));
  csr_merge_kernel<false><<<g, 128, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, b_rowptr, b_colind, b_val, op, rows, s.cnt, nullptr, nullptr);
  SYSML_LAUNCH_CHECK(ctx, "csr_merge_kernel<count>");
  return scan_counts(ctx, s.cnt, s.tmp, s.tmp_bytes, rows, c_rowptr, c_nnz);
}

}  // extern "C"
