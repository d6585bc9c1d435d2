// cellwise.cu -- cellwise binary (matrix-matrix with row/column-vector broadcast, matrix-scalar),
// axpy, unary, fill and transpose.  All HBM-bound: one pass, 128-bit accesses where alignment allows,
// grid sized as a multiple of the SM count.
//
// Reference semantics: LibMatrixBincell.bincellOp (LibMatrixBincell.java:87-339) with the value
// functions in runtime/functionobjects; GPU twins matrix_matrix_cellwise_op / matrix_scalar_op /
// daxpy_matrix_vector / fill (src/main/cpp/kernels/SystemML.cu:698,750,627,783) and the unary kernels
// (:1541-1960).  Opcode numbering = LibMatrixCUDA.getBinaryOp (LibMatrixCUDA.java:1551-1575).
//
// Algorithmic bytes: 8*(size(A) + size(B) + rows*cols) for binary, 16*len for scalar/unary.
#include "common.cuh"
#include "cellwise_ops.cuh"
#include "cg_tail.cuh"

namespace {

using namespace sysml_ops;

// ---------------- flat kernels (no broadcast) ----------------
template <int OP>
__global__ void binary_flat_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ C, i64 len, int vec2) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  const i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec2) {
    const i64 n2 = len >> 1;
    const double2* A2 = reinterpret_cast<const double2*>(A);
    const double2* B2 = reinterpret_cast<const double2*>(B);
    double2* C2 = reinterpret_cast<double2*>(C);
    for (i64 i = tid; i < n2; i += stride) {
      const double2 a = ldg_stream2(A2 + i), b = ldg_stream2(B2 + i);
      C2[i] = make_double2(binop<OP>(a.x, b.x), binop<OP>(a.y, b.y));
    }
    if (tid == 0 && (len & 1)) C[len - 1] = binop<OP>(A[len - 1], B[len - 1]);
  } else {
    for (i64 i = tid; i < len; i += stride) C[i] = binop<OP>(ldg_stream(A + i), ldg_stream(B + i));
  }
}

// broadcast kernels: one division per thread at start, then incremental (row, col) stepping
template <int OP, int BC>
__global__ void binary_bcast_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ C, i64 rows, i64 cols) {
  const i64 len = rows * cols;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  i64 r = i / cols, c = i - r * cols;
  const i64 dr = stride / cols, dc = stride - dr * cols;
  for (; i < len; i += stride) {
    const double b = (BC == SYSML_BCAST_COL) ? __ldg(B + r) : __ldg(B + c);
    const double a = (BC == SYSML_BCAST_OUTER) ? __ldg(A + r) : ldg_stream(A + i);  // outer: A is the column vector
    C[i] = binop<OP>(a, b);
    r += dr;
    c += dc;
    if (c >= cols) {
      c -= cols;
      r++;
    }
  }
}

template <int OP, int LEFT>
__global__ void scalar_kernel(const double* __restrict__ A, double s, double* __restrict__ C, i64 len, int vec2) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  const i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec2) {
    const i64 n2 = len >> 1;
    const double2* A2 = reinterpret_cast<const double2*>(A);
    double2* C2 = reinterpret_cast<double2*>(C);
    for (i64 i = tid; i < n2; i += stride) {
      const double2 a = ldg_stream2(A2 + i);
      C2[i] = LEFT ? make_double2(binop<OP>(s, a.x), binop<OP>(s, a.y)) : make_double2(binop<OP>(a.x, s), binop<OP>(a.y, s));
    }
    if (tid == 0 && (len & 1)) C[len - 1] = LEFT ? binop<OP>(s, A[len - 1]) : binop<OP>(A[len - 1], s);
  } else {
    for (i64 i = tid; i < len; i += stride) {
      const double a = ldg_stream(A + i);
      C[i] = LEFT ? binop<OP>(s, a) : binop<OP>(a, s);
    }
  }
}

template <int OP>
__global__ void unary_kernel(const double* __restrict__ A, double* __restrict__ C, i64 len, int vec2) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  const i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec2) {
    const i64 n2 = len >> 1;
    const double2* A2 = reinterpret_cast<const double2*>(A);
    double2* C2 = reinterpret_cast<double2*>(C);
    for (i64 i = tid; i < n2; i += stride) {
      const double2 a = ldg_stream2(A2 + i);
      C2[i] = make_double2(unop<OP>(a.x), unop<OP>(a.y));
    }
    if (tid == 0 && (len & 1)) C[len - 1] = unop<OP>(A[len - 1]);
  } else {
    for (i64 i = tid; i < len; i += stride) C[i] = unop<OP>(ldg_stream(A + i));
  }
}

// C = A + alpha*B (two roundings, like Java's in1 + cnt*in2)
template <int BC>
__global__ void axpy_kernel(const double* __restrict__ A, const double* __restrict__ B, double alpha, double* __restrict__ C, i64 rows, i64 cols) {
  const i64 len = rows * cols;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  i64 r = 0, c = 0, dr = 0, dc = 0;
  if (BC != SYSML_BCAST_NONE) {
    r = i / cols;
    c = i - r * cols;
    dr = stride / cols;
    dc = stride - dr * cols;
  }
  for (; i < len; i += stride) {
    const double b = (BC == SYSML_BCAST_NONE) ? ldg_stream(B + i) : ((BC == SYSML_BCAST_COL) ? __ldg(B + r) : __ldg(B + c));
    C[i] = __dadd_rn(ldg_stream(A + i), __dmul_rn(alpha, b));
    if (BC != SYSML_BCAST_NONE) {
      r += dr;
      c += dc;
      if (c >= cols) {
        c -= cols;
        r++;
      }
    }
  }
}

__global__ void fill_kernel(double* __restrict__ A, i64 len, double v) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) A[i] = v;
}

// 32x32 tiled transpose through padded shared memory
__global__ void transpose_kernel(const double* __restrict__ A, double* __restrict__ C, i64 rows, i64 cols) {
  __shared__ double tile[32][33];
  const i64 tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  for (i64 t = blockIdx.x; t < tiles_c * tiles_r; t += gridDim.x) {
    const i64 tr = t / tiles_c, tc = t - tr * tiles_c;
    const i64 r0 = tr * 32, c0 = tc * 32;
    for (int y = threadIdx.y; y < 32; y += blockDim.y) {
      const i64 r = r0 + y, c = c0 + threadIdx.x;
      if (r < rows && c < cols) tile[y][threadIdx.x] = A[r * cols + c];
    }
    __syncthreads();
    for (int y = threadIdx.y; y < 32; y += blockDim.y) {
      const i64 c = c0 + y, r = r0 + threadIdx.x;  // output row = c, output col = r
      if (r < rows && c < cols) C[c * rows + r] = tile[threadIdx.x][y];
    }
    __syncthreads();
  }
}


// ------------------------------------------------------------------------------------------------
// Fused m-vector tail of one conjugate-gradient iteration (scripts/algorithms/LinearRegCG.dml:183-195,
// the same update GLM's inner CG uses, GLM.dml:1011-1030):
//     q  += lambda * p                       (regularisation, LinearRegCG.dml:183)
//     a   = nr2 / sum(p * q)
//     beta += a * p ;  r += a * q
//     nr2' = sum(r ^ 2) ;  p = -r + (nr2' / nr2) * p ;  nr2 = nr2'
// The reference compiles this into ~10 separate m-sized instructions (dot products as ba+*, +* axpys,
// scalar ops on the host).  On a B200 each of them is launch latency, nothing else: one CTA does them
// all, with compensated block reductions, and leaves nr2 on the device so no host round trip is needed.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) cg_vector_update_kernel(double* __restrict__ p, double* __restrict__ q, double* __restrict__ r,
                                                                double* __restrict__ beta, double* __restrict__ nr2, i64 m, double lambda) {
  __shared__ double sh[64];
  cg_vector_update_body(p, q, r, beta, nr2, m, lambda, sh);
}

inline unsigned grid_for(sysml_b200_ctx* ctx, i64 work_items, int threads, int per_sm) {
  i64 blocks = (work_items + threads - 1) / threads;
  const i64 cap = (i64)ctx->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int OP>
int launch_binary(sysml_b200_ctx* ctx, const double* A, const double* B, double* C, i64 rows, i64 cols, int bcast) {
  const i64 len = rows * cols;
  if (bcast == SYSML_BCAST_NONE) {
    const int vec2 = aligned16(A) && aligned16(B) && aligned16(C);
    binary_flat_kernel<OP><<<grid_for(ctx, vec2 ? len / 2 + 1 : len, 256, 16), 256, 0, ctx->stream>>>(A, B, C, len, vec2);
  } else if (bcast == SYSML_BCAST_COL) {
    binary_bcast_kernel<OP, SYSML_BCAST_COL><<<grid_for(ctx, len, 256, 16), 256, 0, ctx->stream>>>(A, B, C, rows, cols);
  } else if (bcast == SYSML_BCAST_OUTER) {
    binary_bcast_kernel<OP, SYSML_BCAST_OUTER><<<grid_for(ctx, len, 256, 16), 256, 0, ctx->stream>>>(A, B, C, rows, cols);
  } else {
    binary_bcast_kernel<OP, SYSML_BCAST_ROW><<<grid_for(ctx, len, 256, 16), 256, 0, ctx->stream>>>(A, B, C, rows, cols);
  }
  SYSML_LAUNCH_CHECK(ctx, "binary cellwise kernel");
  return SYSML_OK;
}

template <int OP>
int launch_scalar(sysml_b200_ctx* ctx, const double* A, double s, double* C, i64 len, int left) {
  const int vec2 = aligned16(A) && aligned16(C);
  const unsigned g = grid_for(ctx, vec2 ? len / 2 + 1 : len, 256, 16);
  if (left)
    scalar_kernel<OP, 1><<<g, 256, 0, ctx->stream>>>(A, s, C, len, vec2);
  else
    scalar_kernel<OP, 0><<<g, 256, 0, ctx->stream>>>(A, s, C, len, vec2);
  SYSML_LAUNCH_CHECK(ctx, "scalar cellwise kernel");
  return SYSML_OK;
}

template <int OP>
int launch_unary(sysml_b200_ctx* ctx, const double* A, double* C, i64 len) {
  const int vec2 = aligned16(A) && aligned16(C);
  unary_kernel<OP><<<grid_for(ctx, vec2 ? len / 2 + 1 : len, 256, 16), 256, 0, ctx->stream>>>(A, C, len, vec2);
  SYSML_LAUNCH_CHECK(ctx, "unary cellwise kernel");
  return SYSML_OK;
}

}  // namespace

#define SYSML_FOR_ALL_BINOPS(X) \
  X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16) X(17) X(18)
#define SYSML_FOR_ALL_UNOPS(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12)

extern "C" int sysml_b200_binary_mm(sysml_b200_ctx* ctx, const double* A, const double* B, double* C, int64_t rows,
                                    int64_t cols, int bcast, int op) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "binary_mm: negative dimension");
  if (bcast < 0 || bcast > SYSML_BCAST_OUTER) return sysml_set_error(ctx, SYSML_E_INVALID, "binary_mm: bad broadcast code %d", bcast);
  if (rows * cols == 0) return SYSML_OK;
  if (!A || !B || !C) return sysml_set_error(ctx, SYSML_E_INVALID, "binary_mm: null pointer");
  switch (op) {
#define X(o) \
  case o: return launch_binary<o>(ctx, A, B, C, rows, cols, bcast);
    SYSML_FOR_ALL_BINOPS(X)
#undef X
  }
  return sysml_set_error(ctx, SYSML_E_INVALID, "binary_mm: unknown opcode %d", op);
}

extern "C" int sysml_b200_binary_ms(sysml_b200_ctx* ctx, const double* A, double scalar, double* C, int64_t len, int op,
                                    int scalar_left) {
  SYSML_ENTER(ctx);
  if (len < 0) return sysml_set_error(ctx, SYSML_E_DIM, "binary_ms: negative length");
  if (len == 0) return SYSML_OK;
  if (!A || !C) return sysml_set_error(ctx, SYSML_E_INVALID, "binary_ms: null pointer");
  switch (op) {
#define X(o) \
  case o: return launch_scalar<o>(ctx, A, scalar, C, len, scalar_left);
    SYSML_FOR_ALL_BINOPS(X)
#undef X
  }
  return sysml_set_error(ctx, SYSML_E_INVALID, "binary_ms: unknown opcode %d", op);
}

extern "C" int sysml_b200_unary(sysml_b200_ctx* ctx, const double* A, double* C, int64_t len, int op) {
  SYSML_ENTER(ctx);
  if (len < 0) return sysml_set_error(ctx, SYSML_E_DIM, "unary: negative length");
  if (len == 0) return SYSML_OK;
  if (!A || !C) return sysml_set_error(ctx, SYSML_E_INVALID, "unary: null pointer");
  switch (op) {
#define X(o) \
  case o: return launch_unary<o>(ctx, A, C, len);
    SYSML_FOR_ALL_UNOPS(X)
#undef X
  }
  return sysml_set_error(ctx, SYSML_E_INVALID, "unary: unknown opcode %d", op);
}

extern "C" int sysml_b200_axpy(sysml_b200_ctx* ctx, const double* A, const double* B, double alpha, double* C,
                               int64_t rows, int64_t cols, int bcast) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "axpy: negative dimension");
  if (rows * cols == 0) return SYSML_OK;
  if (!A || !B || !C) return sysml_set_error(ctx, SYSML_E_INVALID, "axpy: null pointer");
  const unsigned g = grid_for(ctx, rows * cols, 256, 16);
  if (bcast == SYSML_BCAST_NONE)
    axpy_kernel<SYSML_BCAST_NONE><<<g, 256, 0, ctx->stream>>>(A, B, alpha, C, rows, cols);
  else if (bcast == SYSML_BCAST_COL)
    axpy_kernel<SYSML_BCAST_COL><<<g, 256, 0, ctx->stream>>>(A, B, alpha, C, rows, cols);
  else if (bcast == SYSML_BCAST_ROW)
    axpy_kernel<SYSML_BCAST_ROW><<<g, 256, 0, ctx->stream>>>(A, B, alpha, C, rows, cols);
  else
    return sysml_set_error(ctx, SYSML_E_INVALID, "axpy: bad broadcast code %d", bcast);
  SYSML_LAUNCH_CHECK(ctx, "axpy_kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_fill(sysml_b200_ctx* ctx, double* A, int64_t len, double value) {
  SYSML_ENTER(ctx);
  if (len < 0 || (len > 0 && !A)) return sysml_set_error(ctx, SYSML_E_INVALID, "fill: bad arguments");
  if (len == 0) return SYSML_OK;
  fill_kernel<<<grid_for(ctx, len, 256, 16), 256, 0, ctx->stream>>>(A, len, value);
  SYSML_LAUNCH_CHECK(ctx, "fill_kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_transpose(sysml_b200_ctx* ctx, const double* A, double* C, int64_t rows, int64_t cols) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "transpose: negative dimension");
  if (rows * cols == 0) return SYSML_OK;
  if (!A || !C) return sysml_set_error(ctx, SYSML_E_INVALID, "transpose: null pointer");
  const i64 tiles = ((rows + 31) / 32) * ((cols + 31) / 32);
  i64 blocks = tiles;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  transpose_kernel<<<(unsigned)blocks, dim3(32, 8), 0, ctx->stream>>>(A, C, rows, cols);
  SYSML_LAUNCH_CHECK(ctx, "transpose_kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_cg_vector_update(sysml_b200_ctx* ctx, double* p, double* q, double* r, double* beta, double* nr2,
                                           int64_t m, double lambda) {
  SYSML_ENTER(ctx);
  if (m <= 0 || !p || !q || !r || !beta || !nr2) return sysml_set_error(ctx, SYSML_E_INVALID, "cg_vector_update: bad arguments");
  if (m > (1 << 22)) return sysml_set_error(ctx, SYSML_E_LIMIT, "cg_vector_update: single-CTA kernel, m <= 4M");
  cg_vector_update_kernel<<<1, 1024, 0, ctx->stream>>>(p, q, r, beta, nr2, m, lambda);
  SYSML_LAUNCH_CHECK(ctx, "cg_vector_update_kernel");
  return SYSML_OK;
}
