/*
 * sysml_b200.h -- flat C ABI of libsystemml_b200.so, the sm_100a native library behind SystemML's
 * GPU instruction classes.
 *
 * What this boundary replaces.  In the reference, the GPU instruction classes
 * (the *GPUInstruction.java classes under runtime/instructions/gpu) call static methods of LibMatrixCUDA /
 * LibMatrixCuMatMult, which hand raw device pointers (jcuda.Pointer, owned by GPUObject /
 * GPUMemoryManager) to JCuda: cuBLAS, cuSPARSE and the PTX kernels of src/main/cpp/kernels/SystemML.cu.
 * Every function below takes the place of one such JCuda call site; the citation on each
 * declaration is the reference interface it stands in for.  Paths are relative to
 *   J/  = src/main/java/org/apache/sysml/
 * The JNI stub a maintainer adds on the Java side is shown in INTEGRATION.md; it follows the
 * reference's own native-plugin convention (J/utils/NativeHelper.java:323-327, src/main/cpp/systemml.h).
 *
 * Conventions
 *  - plain pointers and sizes only; all extents are int64_t (the reference GPU path is int32,
 *    LibMatrixCUDA.java:685,1735-1740, and cannot hold 10M x 1000);
 *  - dense matrices are row-major double (DenseBlockDRB layout; GPUObject.java:52-99);
 *  - CSR is 0-based {double val[nnz], int32 rowptr[rows+1], int32 colind[nnz]} exactly as
 *    CSRPointer.java:76-100 holds it (so nnz < 2^31);
 *  - pointers named d_* / without prefix are DEVICE pointers on the context's device; h_* are host;
 *  - every function returns 0 (SYSML_OK) or a negative SYSML_E_* code, never aborts and never throws;
 *    sysml_b200_last_error() returns the message (the Java side turns it into DMLRuntimeException,
 *    as JCudaKernels.checkResult does, JCudaKernels.java:115-119);
 *  - calls are asynchronous on the context's stream unless stated; sysml_b200_sync() is the
 *    "sysml.gpu.sync.postProcess" synchronisation (GPUInstruction.java:202-208);
 *  - a context is NOT thread-safe: one per (host thread, device), like GPUContext
 *    (GPUContext.java:181-184).
 *  - there is no CPU fallback anywhere in this library.
 */
#ifndef SYSML_B200_H
#define SYSML_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define SYSML_OK 0
#define SYSML_E_INVALID (-1)     /* null pointer, bad enum, negative size            */
#define SYSML_E_DIM (-2)         /* dimension mismatch                               */
#define SYSML_E_OOM (-3)         /* cudaMalloc failed (GPUMemoryManager.java:138-148) */
#define SYSML_E_CUDA (-4)        /* any other CUDA runtime error                     */
#define SYSML_E_UNSUPPORTED (-5) /* valid request this build does not implement      */
#define SYSML_E_LIMIT (-6)       /* size beyond a documented fixed limit             */
#define SYSML_E_WORKSPACE (-7)   /* the caller-owned scratch block is too small: see sysml_b200_ctx_set_workspace */

typedef struct sysml_b200_ctx sysml_b200_ctx;

/* ChainType of lops/MapMultChain.java:31-32 / MMChainCPInstruction.java:69-83 */
#define SYSML_CHAIN_XtXv 0  /* t(X) %*% (X %*% v)        */
#define SYSML_CHAIN_XtwXv 1 /* t(X) %*% (w * (X %*% v))  */
#define SYSML_CHAIN_XtXvy 2 /* t(X) %*% ((X %*% v) - y)  */

/* aggregate opcodes (AggregateUnaryGPUInstruction opcodes uak+ uark+ uack+ uasqk+ uamax uarimax ...) */
#define SYSML_AGG_SUM 0
#define SYSML_AGG_SUMSQ 1
#define SYSML_AGG_MAX 2
#define SYSML_AGG_MIN 3
#define SYSML_AGG_MAXINDEX 4 /* rowIndexMax: 1-based, LAST maximal column wins (LibMatrixAgg.java:3081-3089) */
#define SYSML_AGG_MININDEX 5
#define SYSML_AGG_MEAN 6
#define SYSML_AGG_PROD 7
#define SYSML_AGG_VAR 8
/* direction.  Beware the reference's inverted naming (LibMatrixCUDA.java:696):
 * ROW = one result per row (rowSums, index fn ReduceCol), COL = one result per column (colSums, ReduceRow) */
#define SYSML_DIR_ALL 0
#define SYSML_DIR_ROW 1
#define SYSML_DIR_COL 2

/* cellwise binary opcodes: numbering of LibMatrixCUDA.getBinaryOp (LibMatrixCUDA.java:1551-1575) and
 * binaryOp in src/main/cpp/kernels/SystemML.cu:452-512 */
#define SYSML_OP_PLUS 0
#define SYSML_OP_MINUS 1
#define SYSML_OP_MULTIPLY 2
#define SYSML_OP_DIVIDE 3
#define SYSML_OP_POWER 4
#define SYSML_OP_LESS 5
#define SYSML_OP_LESSEQUAL 6
#define SYSML_OP_GREATER 7
#define SYSML_OP_GREATEREQUAL 8
#define SYSML_OP_EQUAL 9
#define SYSML_OP_NOTEQUAL 10
#define SYSML_OP_MIN 11
#define SYSML_OP_MAX 12
#define SYSML_OP_AND 13
#define SYSML_OP_OR 14
#define SYSML_OP_MINUS1MULTIPLY 15
#define SYSML_OP_MINUSNZ 16
#define SYSML_OP_MODULUS 17
#define SYSML_OP_INTDIV 18
/* broadcast code of the second operand: matrix_matrix_cellwise_op's vecStatus (SystemML.cu:698-735) */
#define SYSML_BCAST_NONE 0 /* B is rows x cols          */
#define SYSML_BCAST_COL 1  /* B is rows x 1 (col vector) */
#define SYSML_BCAST_ROW 2  /* B is 1 x cols (row vector) */
#define SYSML_BCAST_OUTER 3 /* A is rows x 1, B is 1 x cols, C[i,j] = A[i] op B[j]: BinaryAccessType.OUTER_VECTOR_VECTOR
                              (LibMatrixBincell.java:150-163; safeBinaryVVGeneric :565-589), binary_mm only */

/* unary opcodes (kernels matrix_exp ... matrix_sigmoid, SystemML.cu:1541-1960;
 * LibMatrixCUDA.java:1969-2290) */
#define SYSML_UN_EXP 0
#define SYSML_UN_SQRT 1
#define SYSML_UN_ROUND 2
#define SYSML_UN_ABS 3
#define SYSML_UN_LOG 4
#define SYSML_UN_FLOOR 5
#define SYSML_UN_CEIL 6
#define SYSML_UN_SIGN 7
#define SYSML_UN_SIGMOID 8
#define SYSML_UN_SIN 9
#define SYSML_UN_COS 10
#define SYSML_UN_TAN 11
#define SYSML_UN_TANH 12

/* ------------------------------------------------------------------------------------------------
 * Context and device memory.
 * Replaces: GPUContextPool.initializeGPU / getDeviceCount (GPUContextPool.java:84-160),
 *           GPUContext.<init>/initializeThread/destroy (GPUContext.java:93-109,133-184),
 *           GPUMemoryManager.malloc/free/cudaMemGetInfo (GPUMemoryManager.java:157-322,521-526),
 *           GPUObject.copyFromHostToDevice / copyFromDeviceToHost (GPUObject.java:755-946).
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_device_count(int* count);
const char* sysml_b200_version(void);
/* Creates a context on `device` with its own non-blocking stream. */
int sysml_b200_ctx_create(int device, sysml_b200_ctx** out);
/* Same, but all work is queued on a stream the caller owns (cudaStream_t passed as void*). */
int sysml_b200_ctx_create_on_stream(int device, void* cuda_stream, sysml_b200_ctx** out);
int sysml_b200_ctx_destroy(sysml_b200_ctx* ctx);
/* Scratch.  Kernels that reduce across CTAs (mmchain, tsmm split-K, aggregates, the sorts behind the CSR conversions)
 * take their temporary device memory from ONE block per context.  By default the library owns that block: it grows it
 * with cudaMalloc when a larger request arrives and keeps it until the context is destroyed.  A host runtime that
 * accounts for every byte on the card itself -- GPUMemoryManager does (GPUMemoryManager.java:64-78, 157-262) -- hands
 * the block over instead:
 *   ctx_set_workspace(ctx, d_ptr, bytes): from now on scratch comes from this caller-owned, 256-byte aligned block and
 *       the library neither allocates nor frees scratch (d_ptr == NULL returns to the library-owned default).  The call
 *       synchronises the stream first.  The block may be replaced at any time between operations.
 *   An operation that needs more than the block holds changes nothing and returns SYSML_E_WORKSPACE;
 *   ctx_workspace_needed(ctx) then tells how many bytes that operation asked for: allocate (evicting if need be), call
 *   ctx_set_workspace again and repeat the operation -- every operation overwrites its outputs, so repeating is safe.
 *   ctx_workspace_bytes(ctx) is the size of the current block, whoever owns it.
 * (Reference: the operator-local temporaries that LibMatrixCUDA obtains with gCtx.allocate and returns with
 * gCtx.cudaFreeHelper, LibMatrixCUDA.java:711,732,926,947.) */
int sysml_b200_ctx_set_workspace(sysml_b200_ctx* ctx, void* d_ptr, int64_t bytes);
int64_t sysml_b200_ctx_workspace_needed(sysml_b200_ctx* ctx);
int64_t sysml_b200_ctx_workspace_bytes(sysml_b200_ctx* ctx);
int sysml_b200_sync(sysml_b200_ctx* ctx);
const char* sysml_b200_last_error(sysml_b200_ctx* ctx); /* ctx may be NULL: last error of this thread */
void* sysml_b200_ctx_stream(sysml_b200_ctx* ctx);       /* the cudaStream_t in use */
int sysml_b200_ctx_device(sysml_b200_ctx* ctx);
int sysml_b200_ctx_sm_count(sysml_b200_ctx* ctx);
int sysml_b200_mem_info(sysml_b200_ctx* ctx, int64_t* free_bytes, int64_t* total_bytes);
/* Number of kernels this context has launched since creation (GPUStatistics.cudaKernelCount analogue). */
int64_t sysml_b200_launch_count(sysml_b200_ctx* ctx);

/* Unlike GPUMemoryManager.malloc (java:265) memory is NOT zero-filled; ops write every output cell. */
int sysml_b200_malloc(sysml_b200_ctx* ctx, int64_t bytes, void** d_ptr);
int sysml_b200_free(sysml_b200_ctx* ctx, void* d_ptr);
int sysml_b200_memset0(sysml_b200_ctx* ctx, void* d_ptr, int64_t bytes);
int sysml_b200_h2d(sysml_b200_ctx* ctx, void* d_dst, const void* h_src, int64_t bytes); /* synchronous */
int sysml_b200_d2h(sysml_b200_ctx* ctx, void* h_dst, const void* d_src, int64_t bytes); /* synchronous */
int sysml_b200_d2d(sysml_b200_ctx* ctx, void* d_dst, const void* d_src, int64_t bytes);
/* Pinned host staging (cudaMallocHost / cudaFreeHost). */
int sysml_b200_host_alloc(sysml_b200_ctx* ctx, int64_t bytes, void** h_ptr);
int sysml_b200_host_free(sysml_b200_ctx* ctx, void* h_ptr);

/* CUDA-event timers on the context's stream, so the Java-side fine-grained timers
 * (GPUInstruction.java:54-146) can be fed device times instead of System.nanoTime(). */
int sysml_b200_timer_start(sysml_b200_ctx* ctx);
int sysml_b200_timer_stop_ms(sysml_b200_ctx* ctx, double* ms); /* records, synchronises, returns elapsed */

/* ------------------------------------------------------------------------------------------------
 * mmchain (the hot op of LinearRegCG / GLM).
 * Replaces: LibMatrixMult.matrixMultChain (LibMatrixMult.java:243-329) as reached from
 * MMChainCPInstruction.processInstruction (cp/MMChainCPInstruction.java:69-83).  The reference has
 * NO GPU twin (AggBinaryOp.isGPUEnabled is false for MAPMM_CHAIN, J/hops/AggBinaryOp.java:159-162);
 * this is the entry a new MMChainGPUInstruction binds.
 *   X: n x m row-major, v: m, w: n (NULL for XtXv), out: m.  X is read from HBM exactly once.
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_mmchain_dense(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w,
                             double* out, int64_t n, int64_t m, int chain_type);
/* CSR twin: LibMatrixMult.matrixMultChainSparse (LibMatrixMult.java:1655-1684). */
int sysml_b200_mmchain_csr(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                           const double* val, const double* v, const double* w, double* out, int64_t n,
                           int64_t m, int chain_type);
/* Host-vector form: X resident on the device (as GPUObject caches it), v / w / out are HOST buffers;
 * copies v (and w when h_w != NULL and d_w == NULL) in, runs the kernel, copies out back and
 * synchronises.  This is the call shape of the reference's native plugin entry points, which take
 * Java heap arrays (NativeHelper.java:323-327; systemml.cpp:78-124). */
int sysml_b200_mmchain_dense_hostvec(sysml_b200_ctx* ctx, const double* X, const double* h_v,
                                     const double* h_w, const double* d_w, double* h_out, int64_t n,
                                     int64_t m, int chain_type);

/* Allocation-free form of sysml_b200_mmchain_dense.  It has NO counterpart in the reference; it exists
 * for callers that must not have device memory allocated on their behalf: a host runtime that budgets
 * the whole card itself (SystemML's GPUMemoryManager is one) cannot have the library call cudaMalloc
 * behind its back, and code running where the CUDA runtime may not be asked for memory cannot either.
 * Same arguments, same result and same error codes as sysml_b200_mmchain_dense, with these differences:
 *  - cols must be <= SYSML_B200_NOALLOC_MAX_COLS, otherwise SYSML_E_LIMIT is returned;
 *  - the scratch is a fixed array inside the loaded library (about 2.4 MiB per GPU, held from load to
 *    process exit); no cudaMalloc / cudaFree or other allocating routine is called;
 *  - all calls on one GPU share that array: calls that overlap on the GPU race with each other;
 *  - the first call on each device resolves the array's device address; issue it before starting any
 *    stream capture;
 *  - CUDA places device code on a GPU lazily by default, so that first call is also the one that makes the
 *    array (and the kernels) resident: it needs about 2.4 MiB plus code of free device memory.  Make one
 *    small warm-up call per device right after creating the context, before the card fills up, or run with
 *    CUDA_MODULE_LOADING=EAGER.  From the second call on, free device memory is the same before and after. */
#define SYSML_B200_NOALLOC_MAX_COLS 1024
int sysml_b200_mmchain_dense_noalloc(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w,
                                     double* out, int64_t n, int64_t m, int chain_type);

/* ------------------------------------------------------------------------------------------------
 * tsmm.  Replaces LibMatrixCUDA.matmultTSMM (LibMatrixCUDA.java:477-524: cublasDsyrk + copy_u2l_dense)
 * called by MMTSJGPUInstruction.processInstruction (gpu/MMTSJGPUInstruction.java:78-89); CPU twin
 * LibMatrixMult.matrixMultTransposeSelf (LibMatrixMult.java:331-399).
 *   left != 0: out = t(X) %*% X (cols x cols); left == 0: out = X %*% t(X) (rows x rows).
 *   h_nnz_out (HOST pointer, may be NULL): exact non-zero count of out, as copyUpperToLowerTriangle
 *   returns it (LibMatrixMult.java:3633-3672); when non-NULL the call synchronises before returning.
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_tsmm_dense(sysml_b200_ctx* ctx, const double* X, double* out, int64_t rows, int64_t cols,
                          int left, int64_t* h_nnz_out);
int sysml_b200_tsmm_csr_left(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                             const double* val, double* out, int64_t rows, int64_t cols,
                             int64_t* h_nnz_out);

/* ------------------------------------------------------------------------------------------------
 * General matrix multiply C = op(A) %*% op(B).
 * Replaces LibMatrixCuMatMult.matmult (LibMatrixCuMatMult.java:139-229) called by
 * AggregateBinaryGPUInstruction.processInstruction (gpu/AggregateBinaryGPUInstruction.java:73-89):
 *   dense x dense : cublasDdot / cublasDgemv / cublasDgemm (LibMatrixCuMatMult.java:362-408)
 *   CSR   x dense : cusparseDcsrmv / cusparseDcsrmm2 + dgeam (java:266-341)
 * a_rows x a_cols and b_rows x b_cols are the STORED shapes; trans flags as in the instruction
 * string "GPU.ba+*.A.B.out.leftT.rightT".  C is row-major (m x n) and fully written.
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_matmult_dense(sysml_b200_ctx* ctx, const double* A, const double* B, double* C,
                             int64_t a_rows, int64_t a_cols, int64_t b_rows, int64_t b_cols, int trans_a,
                             int trans_b);
int sysml_b200_matmult_csr_dense(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                                 const double* val, const double* B, double* C, int64_t a_rows,
                                 int64_t a_cols, int64_t b_rows, int64_t b_cols, int trans_a);

/* Remaining sparsity combinations of ba+* (LibMatrixCuMatMult.matmult :150-195: cusparseDcsrgemm for
 * sparse x sparse, the dense x sparse branch; CPU LibMatrixMult.matrixMultDenseSparse :1105-1168 and
 * matrixMultSparseSparse :1397-1500).  C is always written dense and row-major. */
int sysml_b200_matmult_dense_csr(sysml_b200_ctx* ctx, const double* A, const int32_t* rowptr,
                                 const int32_t* colind, const double* val, double* C, int64_t a_rows,
                                 int64_t a_cols, int64_t b_rows, int64_t b_cols, int trans_a);
int sysml_b200_matmult_csr_csr(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind,
                               const double* a_val, const int32_t* b_rowptr, const int32_t* b_colind,
                               const double* b_val, double* C, int64_t a_rows, int64_t a_cols,
                               int64_t b_rows, int64_t b_cols);
/* Operations whose RESULT is CSR.  No explicit zeros are stored (as the CPU's sparse rows: SparseRow.append skips
 * them), so *c_nnz equals MatrixBlock.recomputeNonZeros of the dense result.
 *   csr_transpose           : CSRPointer.transpose (csr2csc, CSRPointer.java:463-481); nnz is unchanged, the caller
 *                             allocates t_rowptr[cols+1], t_colind[nnz], t_val[nnz].
 *   csr_add                 : C = alpha*A + beta*B, the sparse branch of LibMatrixCUDA.dgeam (cusparseDcsrgeam,
 *                             LibMatrixCUDA.java:1593-1701; CSRPointer.allocateForDgeam :227-299).
 *   matmult_csr_csr_sparse  : LibMatrixCuMatMult.sparseSparseMatMult (cusparseDcsrgemm, LibMatrixCuMatMult.java:150-195;
 *                             CSRPointer.allocateForMatrixMultiply :301-399); per output cell the products are added
 *                             in ascending k like LibMatrixMult.matrixMultSparseSparse (:1397-1500): bit-identical.
 * For the last two the caller passes c_rowptr[rows+1]; the library sizes the result in a symbolic pass, allocates
 * *c_colind / *c_val itself (release them with sysml_b200_free) and returns the count in *c_nnz.
 * matmult_csr_csr_sparse refuses more than 2^28 intermediate products with SYSML_E_LIMIT. */
int sysml_b200_csr_transpose(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                             int64_t rows, int64_t cols, int32_t* t_rowptr, int32_t* t_colind, double* t_val);
int sysml_b200_csr_add(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val,
                       double alpha, const int32_t* b_rowptr, const int32_t* b_colind, const double* b_val,
                       double beta, int64_t rows, int64_t cols, int32_t* c_rowptr, int32_t** c_colind,
                       double** c_val, int64_t* c_nnz);
int sysml_b200_matmult_csr_csr_sparse(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind,
                                      const double* a_val, const int32_t* b_rowptr, const int32_t* b_colind,
                                      const double* b_val, int64_t a_rows, int64_t a_cols, int64_t b_rows,
                                      int64_t b_cols, int32_t* c_rowptr, int32_t** c_colind, double** c_val,
                                      int64_t* c_nnz);
/* Ultra-sparse left operand times a dense matrix with a CSR result: LibMatrixMult.matrixMultUltraSparse, left branch
 * (LibMatrixMult.java:1510-1566; chosen by isUltraSparseMatrixMult :3761-3777, e.g. a selection / permutation matrix times
 * X): only the rows of A that hold cells produce anything -- one cell selects and scales a row of B, more cells add their
 * products in ascending k -- and zero results are not stored.  Two-call protocol as for the sparse cellwise operators
 * below (c_colind == c_val == NULL: symbolic pass). */
int sysml_b200_matmult_ultra_sparse_left(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind,
                                         const double* a_val, const double* B, int64_t a_rows, int64_t a_cols,
                                         int64_t b_cols, int32_t* c_rowptr, int32_t* c_colind, double* c_val,
                                         int64_t* c_nnz);

/* Sparse-safe cellwise operations whose operand AND result stay CSR: the value function is evaluated at the stored
 * cells only and results that are zero are dropped -- LibMatrixBincell.safeBinaryScalar's SPARSE <- SPARSE branch
 * (LibMatrixBincell.java:938-995), safeBinaryMVSparse (:338-440), safeBinaryMMSparseDenseSkip (:760-790),
 * safeBinaryMMSparseSparse + mergeForSparseBinary (:597-660, 1279-1340; inner join for Multiply, outer join with
 * op(a,0) / op(0,b) otherwise) and the sparse-safe unary path of MatrixBlock.unaryOperations.  (The reference's GPU
 * twin has no such path: LibMatrixCUDA converts every sparse operand of a cellwise operator to dense,
 * LibMatrixCUDA.java:1360-1383 -- at the C3 shape that block would be 4 TB.)
 * Whether an (operator, operand) pair is sparse-safe is the caller's decision, by the reference's rules
 * (BinaryOperator.java:60-66, RightScalarOperator.java:42-49, LeftScalarOperator.java:41-46,
 * LibMatrixBincell.isSparseSafeDivide :185-189); csr_scalar_op / csr_unary_op / csr_csr_op also check that the
 * operator maps zero to zero and return SYSML_E_UNSUPPORTED otherwise.
 * Two-call protocol (CSRPointer.allocateForDgeam's, CSRPointer.java:227-299), nothing is allocated by the library:
 *   1. c_colind == c_val == NULL: symbolic pass, writes c_rowptr[rows+1] and *c_nnz;
 *   2. the caller allocates c_colind[*c_nnz], c_val[*c_nnz] and calls again with the same operands and the c_rowptr
 *      of step 1: numeric pass.
 *   csr_scalar_op : C = A op scalar (scalar_left: scalar op A), op = SYSML_OP_*
 *   csr_unary_op  : C = f(A), f = SYSML_UN_*
 *   csr_dense_op  : C = A op B (dense_left: B op A) at A's stored cells; B dense rows x cols, rows x 1 or 1 x cols
 *                   (bcast = SYSML_BCAST_NONE / _COL / _ROW)
 *   csr_csr_op    : C = A op B, both CSR rows x cols */
int sysml_b200_csr_scalar_op(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                             int64_t rows, int64_t cols, double scalar, int op, int scalar_left, int32_t* c_rowptr,
                             int32_t* c_colind, double* c_val, int64_t* c_nnz);
int sysml_b200_csr_unary_op(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                            int64_t rows, int64_t cols, int un_op, int32_t* c_rowptr, int32_t* c_colind,
                            double* c_val, int64_t* c_nnz);
int sysml_b200_csr_dense_op(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                            int64_t rows, int64_t cols, const double* B, int bcast, int op, int dense_left,
                            int32_t* c_rowptr, int32_t* c_colind, double* c_val, int64_t* c_nnz);
int sysml_b200_csr_csr_op(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val,
                          const int32_t* b_rowptr, const int32_t* b_colind, const double* b_val, int64_t rows,
                          int64_t cols, int op, int32_t* c_rowptr, int32_t* c_colind, double* c_val, int64_t* c_nnz);
/* X %*% t(X) for CSR X (LibMatrixMult.prepMatrixMultTransposeSelfInput :3679-3685 + the RIGHT kernel). */
int sysml_b200_tsmm_csr_right(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                              const double* val, double* out, int64_t rows, int64_t cols,
                              int64_t* h_nnz_out);

/* ------------------------------------------------------------------------------------------------
 * Aggregate-unary.  Replaces LibMatrixCUDA.unaryAggregate (LibMatrixCUDA.java:573-894) called by
 * AggregateUnaryGPUInstruction.processInstruction (gpu/AggregateUnaryGPUInstruction.java:67-101);
 * CPU twin LibMatrixAgg.aggregateUnaryMatrix (LibMatrixAgg.java:194-282).
 *   out: 1 (ALL), rows (ROW) or cols (COL) doubles on the device.  MAXINDEX/MININDEX are ROW only and
 *   write the 1-based index as a double (the n x 2 "index,value" block of d_uarimxx collapses to its
 *   first column, which is what the CP instruction keeps: MatrixBlock.java:4080-4125).
 *   Sums are compensated (two-sum) so they track the reference's Kahan result, not the plain tree sum
 *   of the old GPU path (LibMatrixCUDA.java:584-585).
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_agg_unary_dense(sysml_b200_ctx* ctx, const double* X, int64_t rows, int64_t cols, int op,
                               int dir, double* out);
int sysml_b200_agg_unary_csr(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                             const double* val, int64_t rows, int64_t cols, int op, int dir, double* out);

/* ------------------------------------------------------------------------------------------------
 * Cellwise.  Replace LibMatrixCUDA.matrixMatrixArithmetic / matrixMatrixRelational (java:1245-1270,
 * 1209-1230; kernel matrix_matrix_cellwise_op SystemML.cu:698), matrixScalarArithmetic (java:1126-1195;
 * kernel matrix_scalar_op SystemML.cu:750), axpy (java:2295-2350; cublasDaxpy / daxpy_matrix_vector
 * SystemML.cu:627), the unary kernels (SystemML.cu:1541-1960) and fill (SystemML.cu:783).
 * CPU twin: LibMatrixBincell.bincellOp (LibMatrixBincell.java:87-339) + the value functions in runtime/functionobjects.
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_binary_mm(sysml_b200_ctx* ctx, const double* A, const double* B, double* C, int64_t rows,
                         int64_t cols, int bcast, int op);
int sysml_b200_binary_ms(sysml_b200_ctx* ctx, const double* A, double scalar, double* C, int64_t len,
                         int op, int scalar_left);
int sysml_b200_axpy(sysml_b200_ctx* ctx, const double* A, const double* B, double alpha, double* C,
                    int64_t rows, int64_t cols, int bcast);
int sysml_b200_unary(sysml_b200_ctx* ctx, const double* A, double* C, int64_t len, int op);
int sysml_b200_fill(sysml_b200_ctx* ctx, double* A, int64_t len, double value);
/* Fused m-vector tail of one CG iteration of LinearRegCG / GLM (scripts/algorithms/LinearRegCG.dml:183-195):
 *   q += lambda*p; a = nr2 / sum(p*q); beta += a*p; r += a*q; nr2' = sum(r^2); p = -r + (nr2'/nr2)*p; nr2 = nr2'.
 * Stands in for the ~10 m-sized instructions (ba+* dot products, +* axpys, host scalars) the reference
 * compiles this block into; all operands are device pointers (nr2: one double), nothing returns to the host. */
int sysml_b200_cg_vector_update(sysml_b200_ctx* ctx, double* p, double* q, double* r, double* beta, double* nr2,
                                int64_t m, double lambda);
/* Fused cellwise chain: out = f(inputs, consts) cell by cell in ONE pass over HBM, optionally reduced by
 * sum / rowSums / colSums in the same pass.  Stands in for a run of matrixMatrixArithmetic / matrixScalarArithmetic
 * / unaryOp launches and the unaryAggregate after them (LibMatrixCUDA.java:1126-1270, 1969-2290, 573-894), each
 * of which round-trips a temporary through HBM; the reference's own fused form is the code generator's cellwise
 * template (runtime/codegen/SpoofCellwise.java: NO_AGG / FULL_AGG / ROW_AGG / COL_AGG).
 *   inputs[i]      device pointers; input_kinds[i] = SYSML_BCAST_NONE (rows x cols), _COL (rows x 1), _ROW (1 x cols)
 *   program        postfix tokens SYSML_FUSED_TOKEN(kind, arg); binary/unary args are SYSML_OP_* / SYSML_UN_*
 *   agg            SYSML_FUSED_AGG_* (none, or sum / min / max over all cells, per row, per column); out holds
 *                  rows*cols, 1, rows or cols doubles accordingly
 * The chain is compiled to a kernel on first use (NVRTC, sm_100a) and cached by program; every operator has the
 * arithmetic of its unfused kernel, so without an aggregate the result is bit-identical to the unfused sequence.
 * Fails with SYSML_E_UNSUPPORTED when NVRTC or the driver library cannot be loaded (no CPU path). */
#define SYSML_FUSED_IN 0     /* push inputs[arg] at the current cell */
#define SYSML_FUSED_CONST 1  /* push consts[arg] */
#define SYSML_FUSED_BINARY 2 /* pop y, pop x, push x <SYSML_OP_arg> y */
#define SYSML_FUSED_UNARY 3  /* pop x, push SYSML_UN_arg(x) */
#define SYSML_FUSED_PICK 4   /* push a copy of the value arg places below the top (0 = duplicate the top): shared subexpressions */
#define SYSML_FUSED_NIP 5    /* drop the value just below the top (ends the scope of a shared subexpression) */
#define SYSML_FUSED_TOKEN(kind, arg) (((kind) << 24) | (arg))
#define SYSML_FUSED_AGG_NONE 0
#define SYSML_FUSED_AGG_SUM 1
#define SYSML_FUSED_AGG_ROWSUMS 2
#define SYSML_FUSED_AGG_COLSUMS 3
#define SYSML_FUSED_AGG_MIN 4      /* min / max endings (SpoofCellwise.AggOp MIN, MAX; sum of squares is x*x in the chain) */
#define SYSML_FUSED_AGG_ROWMINS 5
#define SYSML_FUSED_AGG_COLMINS 6
#define SYSML_FUSED_AGG_MAX 7
#define SYSML_FUSED_AGG_ROWMAXS 8
#define SYSML_FUSED_AGG_COLMAXS 9
#define SYSML_FUSED_MAX_INPUTS 8
#define SYSML_FUSED_MAX_CONSTS 8
#define SYSML_FUSED_MAX_TOKENS 64
int sysml_b200_fused_cellwise(sysml_b200_ctx* ctx, const double* const* inputs, const int* input_kinds, int n_inputs,
                              const double* consts, int n_consts, const int32_t* program, int n_tokens, int64_t rows,
                              int64_t cols, int agg, double* out);
/* Several full sums over the same operands in ONE pass: the code generator's multi-aggregate template
 * (runtime/codegen/SpoofMultiAggregate.java, hops/codegen/template/TemplateMultiAgg.java), which is what the compiler emits
 * for e.g. l2-svm.dml's line search, sum(out*sv*Y*Xd) and sum(Xd*sv*Xd) over the same three n-vectors.
 * The postfix program leaves n_out values on the stack (result 0 deepest); out[j] = sum over all cells of value j, with
 * the compensated accumulation and the fixed merge order of fused_cellwise's SUM ending (bit-identical to n_out separate
 * calls). 1 <= n_out <= SYSML_FUSED_MAX_RESULTS. */
#define SYSML_FUSED_MAX_RESULTS 4
int sysml_b200_fused_multi_agg(sysml_b200_ctx* ctx, const double* const* inputs, const int* input_kinds, int n_inputs,
                               const double* consts, int n_consts, const int32_t* program, int n_tokens, int64_t rows,
                               int64_t cols, int n_out, double* out);
/* C = t(A); replaces LibMatrixCUDA.transpose (java:1723-1729, cublasDgeam). */
int sysml_b200_transpose(sysml_b200_ctx* ctx, const double* A, double* C, int64_t rows, int64_t cols);

/* ------------------------------------------------------------------------------------------------
 * Sparsity bookkeeping.  Replace cusparseDnnz / cusparseDdense2csr / cusparseDcsr2dense
 * (GPUObject.java:199,223,508; CSRPointer.java:463-481).  Counts use (x != 0): -0.0 is zero, NaN is not.
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_count_nnz(sysml_b200_ctx* ctx, const double* X, int64_t len, int64_t* h_nnz);
/* two-phase like the reference: first call with colind == NULL fills rowptr and *h_nnz. */
int sysml_b200_dense_to_csr(sysml_b200_ctx* ctx, const double* X, int64_t rows, int64_t cols,
                            int32_t* rowptr, int32_t* colind, double* val, int64_t* h_nnz);
int sysml_b200_csr_to_dense(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                            const double* val, int64_t rows, int64_t cols, double* X);

/* ------------------------------------------------------------------------------------------------
 * One Lloyd step of Kmeans (scripts/algorithms/Kmeans.dml:146-166) in ONE pass over X:
 *   D = -2*(X %*% t(C)) + t(c2);  minD = rowMins(D);  P = (D <= minD) / rowSums(D <= minD)
 *   sums[k x m] = t(P) %*% X;  counts[k] = colSums(P);  *sum_min_dist = sum(minD)
 * Stands in for the run ba+* (X %*% t(C)), *, +, uarmin, uak+, <=, uark+, /, uack+, ba+* (t(P) %*% X) that the reference
 * executes as separate instructions (AggregateBinaryGPUInstruction.java:73-89 twice, AggregateUnaryGPUInstruction.java:67-101,
 * MatrixMatrixArithmeticGPUInstruction.java:38-71; CPU: LibMatrixMult.java:880-1103, LibMatrixAgg.java:1624-1927), each
 * of which round-trips an n x k matrix through memory.  c2 = rowSums(C^2) (k doubles) is computed by the caller with
 * agg_unary_dense so that its Kahan arithmetic is the reference's; the division by the counts stays with the caller too.
 * Ties share a record equally, as in the script.  Records containing NaN are left out of the sums (the script would
 * propagate the NaN into every centroid).  Limits: k <= 64, an even number of columns, and C, the accumulators and two
 * 64-row tiles must fit in shared memory (SYSML_E_UNSUPPORTED otherwise: run the step operator by operator). */
int sysml_b200_kmeans_step(sysml_b200_ctx* ctx, const double* X, const double* C, const double* c2, int64_t n,
                           int64_t m, int64_t k, double* sums, double* counts, double* sum_min_dist);

/* ------------------------------------------------------------------------------------------------
 * Column-ordered second copy of a constant CSR matrix for t(X) %*% y and the second half of sparse mmchain
 * ("panelled sliced ELL", csrc/psell.cu).  Replaces the scatter of LibMatrixMult.matrixMultChainSparse
 * (LibMatrixMult.java:1655-1684) and of the VM form of matrixMultSparseDense (:1262-1286) -- and, on the reference GPU
 * path, cusparseDcsrmv with the transpose flag (LibMatrixCuMatMult.java:313-341) -- by a gather without atomics whose
 * result is bit-identical run to run.  Worth building when X is used more than once, as the iterative algorithms do; the
 * CSR entry points above remain the path for one-shot operands.
 * Layout: rows in panels of panel_rows, columns in 32 slices.  Inside a (panel, slice) the entries of a column form a
 * piece (cut at 64 entries), pieces are sorted by length and laid out 32 side by side as a group -- a lane owns a column
 * run, entry j of the 32 pieces is one coalesced line -- so t(X) %*% y costs a load, a shared-memory gather, a multiply
 * and an add per non-zero and nothing else.
 * Storage: 10 bytes per slot (f64 value + u16 row inside the panel; slots = nnz +
 * padding, a few percent for Poisson-like column counts) + 72 bytes per group.
 *   csr_to_psell       : builds the copy.  sgrp_ptr[sysml_b200_psell_segments(rows, panel_rows)] i32 is caller-allocated;
 *                        the slot and group arrays are sized by the build and allocated by the library (release each with
 *                        sysml_b200_free): *s_val f64[*n_slots], *s_lrow u16[*n_slots], *s_gcol u16[32 * *n_groups],
 *                        *s_gdesc u64[*n_groups].  Temporaries (about 30 bytes per non-zero) are taken from and returned
 *                        to the CUDA allocator inside the call.  cols <= 65535, panel_rows <= 8192.
 *   matmult_psell_vm   : out[cols] = t(X) %*% (y [* w | - w]);  w_op 0 none, 1 multiply, 2 subtract.
 *                        cols + 2 * (panel_rows + 1) doubles must fit in shared memory (SYSML_E_LIMIT).
 *   mmchain_csr_psell  : out = t(X) %*% (w * (X %*% v)) etc. (X %*% v on the CSR copy into tmp_n, rows doubles).
 * No atomics: results are bit-identical from run to run.
 * ---------------------------------------------------------------------------------------------- */
int64_t sysml_b200_psell_segments(int64_t rows, int panel_rows);
int sysml_b200_csr_to_psell(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                            int64_t rows, int64_t cols, int64_t nnz, int panel_rows, int32_t* sgrp_ptr, double** s_val,
                            uint16_t** s_lrow, uint16_t** s_gcol, uint64_t** s_gdesc, int64_t* n_slots,
                            int64_t* n_groups);
int sysml_b200_matmult_psell_vm(sysml_b200_ctx* ctx, const double* s_val, const uint16_t* s_lrow, const uint16_t* s_gcol,
                                const uint64_t* s_gdesc, const int32_t* sgrp_ptr, int64_t rows, int64_t cols,
                                int panel_rows, const double* y, const double* w, int w_op, double* out);
int sysml_b200_mmchain_csr_psell(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                 const double* s_val, const uint16_t* s_lrow, const uint16_t* s_gcol,
                                 const uint64_t* s_gdesc, const int32_t* sgrp_ptr, int panel_rows, const double* v,
                                 const double* w, double* tmp_n, double* out, int64_t n, int64_t m, int chain_type);

/* ------------------------------------------------------------------------------------------------
 * Multi-GPU sum exchange over NVLink peer memory (one process per GPU on one box, world <= 8).
 * Replaces the aggregation step of SystemML's distributed tsmm / mmchain plan, RDDAggregateUtils.sumStable
 * (J/runtime/instructions/spark/utils/RDDAggregateUtils.java:58-70, used by MapmmChainSPInstruction and
 * TsmmSPInstruction): X is row-sharded, every rank computes on its shard, the m-sized partials are added.
 * The reference has no GPU collective at all.  The sum is done INSIDE the kernel that finishes the local
 * result (P2P stores into every peer's mailbox + per-CTA flags), not by a separate NCCL call.
 *   comm_create : allocates this rank's mailbox (capacity max_values doubles per message) and returns its
 *                 64-byte cudaIpcMemHandle in handle_out64; exchange the handles of all ranks by any means
 *                 (the bench uses torch.distributed.all_gather) and pass them, in rank order, to comm_connect.
 *                 handle_out64 may be NULL when the ranks will be connected with comm_connect_local.
 *   comm_connect_local : ranks that are contexts of ONE process -- the way the reference itself drives several
 *                 GPUs: every parfor worker thread of one JVM owns a GPUContext and binds its device with
 *                 cudaSetDevice (GPUContext.initializeThread, GPUContext.java:181-184; ParForProgramBlock.java:833).
 *                 No IPC: peer access is enabled between the devices and every rank gets the peers' mailbox
 *                 pointers.  comms[r] must be rank r of n.  Several ranks may share one device (tests).  Each rank
 *                 still issues its own calls on its own context (from its own thread, or one thread issuing the
 *                 asynchronous calls of all ranks in turn before synchronising any of them).
 *   Every rank must issue the same sequence of collective calls with the same sizes.  This is checked: a flag
 *   carries (count, combiner, kind) of the call next to its epoch, and every wait has a time budget (60 s unless
 *   comm_set_timeout_ms changes it).  When a peer issues a different call, or does not arrive in time (it died,
 *   skipped a call, failed validation), the kernel ends without hanging, and the NEXT call on the communicator
 *   -- or comm_check, which first synchronises the context's stream -- returns SYSML_E_CUDA naming the peer
 *   and the epoch.  A communicator that failed stays failed (epochs no longer agree): destroy it.
 * ---------------------------------------------------------------------------------------------- */
typedef struct sysml_b200_comm sysml_b200_comm;
int sysml_b200_comm_create(sysml_b200_ctx* ctx, int rank, int world, int64_t max_values,
                           sysml_b200_comm** out, void* handle_out64);
int sysml_b200_comm_connect(sysml_b200_comm* comm, const void* all_handles);
int sysml_b200_comm_connect_local(sysml_b200_comm** comms, int n);
int sysml_b200_comm_set_timeout_ms(sysml_b200_comm* comm, int64_t ms);
int sysml_b200_comm_check(sysml_b200_comm* comm);
int sysml_b200_comm_destroy(sysml_b200_comm* comm);
/* buf[0..len) <- sum over ranks, in place, identical bits on every rank (buffers longer than max_values go
 * through the mailbox in pieces). */
int sysml_b200_allreduce_sum(sysml_b200_ctx* ctx, sysml_b200_comm* comm, double* buf, int64_t len);
/* Same exchange with sum, max or min as the combiner (colMaxs / colMins / max / min of row-sharded
 * operands; Math.max/min semantics, so the result is exact and identical on every rank). */
#define SYSML_REDUCE_SUM 0
#define SYSML_REDUCE_MAX 1
#define SYSML_REDUCE_MIN 2
int sysml_b200_allreduce(sysml_b200_ctx* ctx, sysml_b200_comm* comm, double* buf, int64_t len, int op);
/* mmchain on the local row shard fused with the sum over ranks (m <= 8192). */
int sysml_b200_mmchain_dense_allreduce(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* X,
                                       const double* v, const double* w, double* out, int64_t n,
                                       int64_t m, int chain_type);
/* mmchain with a row function: out[m] = t(X) %*% f(X %*% v, w, y), X read once.  f is a postfix program of
 * SYSML_FUSED_TOKENs (as for sysml_b200_fused_cellwise) over three inputs per row -- 0: (X %*% v)[i], 1: w[i], 2: y[i]
 * (w / y may be NULL when the program does not read them) -- and up to 4 constants; it leaves one value, or two, in which
 * case the second is written to side_out[i].  Stands in for  ba+* (X %*% v), the cellwise instructions of f, r' and ba+*
 * (t(X) %*% .)  where the reference's MapMultChain has only the three fixed forms XtXv / XtwXv / XtXvy
 * (MapMultChain.java:31-32; LibMatrixMult.matrixMultChain, LibMatrixMult.java:243-268) -- e.g. the outer step of GLM.dml
 * for Poisson / log link: f = y - exp(eta) with exp(eta), the next inner loop's weights, as the second value
 * (scripts/algorithms/GLM.dml:702-706, 1011-1030).  The program is interpreted once per row (a few dozen instructions
 * next to the row's kilobytes).  At most 24 tokens, stack depth 6, cols <= 8192 (SYSML_E_UNSUPPORTED above). */
int sysml_b200_mmchain_rowfn_dense(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w, const double* y,
                                   const int32_t* program, int n_tokens, const double* consts, int n_consts, int n_results,
                                   double* side_out, double* out, int64_t n, int64_t m);
/* One whole conjugate-gradient iteration of LinearRegCG / GLM (scripts/algorithms/LinearRegCG.dml:176-195; inner CG of
 * GLM.dml:1011-1030) on this rank's row shard, in two launches:
 *   q = t(X) %*% (w * (X %*% p))  [chain_type as for mmchain; summed over the ranks of comm unless comm is NULL]
 *   then sysml_b200_cg_vector_update(p, q, r, beta, nr2, m, lambda) -- same arithmetic, same bits --
 *   applied by the last CTA of the finishing kernel instead of a third launch.
 * The reference runs this as MMChainCPInstruction (MMChainCPInstruction.java:69-83) + ~10 m-sized CP instructions.
 * Both kernels are programmatic dependent launches (their set-up overlaps the previous kernel's tail; they touch no
 * global memory before that kernel has completed).  m <= 8192.  Every rank must make the call. */
int sysml_b200_cg_step_dense(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* X, const double* w,
                             double* p, double* q, double* r, double* beta, double* nr2, int64_t n,
                             int64_t m, int chain_type, double lambda);

/* ------------------------------------------------------------------------------------------------
 * SystemML "binaryblock" files (Hadoop SequenceFile, version 6, uncompressed, of MatrixIndexes -> MatrixBlock records;
 * blocks of brlen x bclen, 1000 x 1000 by default) straight into a row-major device matrix.
 * Replaces ReaderBinaryBlock.readBinaryBlockMatrixFromSequenceFile (J/runtime/io/ReaderBinaryBlock.java:124-207) followed
 * by GPUObject.copyFromHostToDevice (GPUObject.java:755-870): no JVM double[] in between (which also caps a dense block
 * array at 2^31 cells), no synchronous whole-matrix copy.  Record layouts: MatrixIndexes.write (MatrixIndexes.java:117-130),
 * MatrixBlock.write / readFields (MatrixBlock.java:1735-2180): dense, sparse, ultra-sparse and empty blocks are all read.
 *   binaryblock_read_dense   : rows [row_begin, row_end) of the matrix -- this rank's shard; row_begin on a block boundary --
 *                              into d_X ((row_end-row_begin) x cols doubles).  `path` is one SequenceFile or a directory of
 *                              part files.  Raw block bytes go pinned slot -> device, a kernel swaps the byte order and
 *                              places the block; file reads overlap the copies.  Blocks absent from the file are zeros.
 *   binaryblock_read_mmchain : the same, and out = t(X) %*% (w * (X %*% v)) [chain_type as for mmchain] of the shard is
 *                              computed WHILE the matrix arrives: as soon as a group of block rows is on the device its
 *                              partial result is added (in row order, so the result is reproducible).  The first pass of
 *                              LinearRegCG / GLM over a matrix that is not resident yet costs the upload, not upload + pass.
 *   binaryblock_write        : host-side writer of the same format (tests, data preparation); dense_only = 0 picks each
 *                              block's format as MatrixBlock.write does (evalSparseFormatOnDisk, MatrixBlock.java:1017-1031,
 *                              isUltraSparseSerialize :897-899).  first_block_row (0-based) is added to the block-row index
 *                              of every record, so that h_X can be one part (a run of whole block rows) of a larger matrix
 *                              written as a directory of part files.  Takes no context.
 * Compressed SequenceFiles are refused (SYSML_E_UNSUPPORTED).
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_binaryblock_write(const char* path, const double* h_X, int64_t rows, int64_t cols, int brlen, int bclen,
                                 int dense_only, int64_t first_block_row);
int sysml_b200_binaryblock_read_dense(sysml_b200_ctx* ctx, const char* path, int64_t rows, int64_t cols, int brlen,
                                      int bclen, int64_t row_begin, int64_t row_end, double* d_X);
int sysml_b200_binaryblock_read_mmchain(sysml_b200_ctx* ctx, const char* path, int64_t rows, int64_t cols, int brlen,
                                        int bclen, int64_t row_begin, int64_t row_end, double* d_X, const double* v,
                                        const double* w, double* out, int chain_type);

/* ------------------------------------------------------------------------------------------------
 * Synthetic data (scripts/datagen/genRandData4LinearRegression.dml uses Rand(min,max,pdf=uniform,seed)).
 * Counter-based: cell i of the GLOBAL array gets hash(seed, offset + i), so any row shard equals the
 * same rows of the full matrix.  No reference GPU counterpart (RandGPU does not exist); the CPU twin is
 * LibMatrixDatagen.  Used by bench.py and the tests.
 * ---------------------------------------------------------------------------------------------- */
int sysml_b200_rand_uniform(sysml_b200_ctx* ctx, double* out, int64_t len, uint64_t seed, int64_t offset,
                            double lo, double hi);
/* Synthetic CSR in the shape of BASELINE config C3 (every cell non-zero with a small probability, i.e. ~Poisson(mean)
 * entries per row; the CPU analogue is Rand(..., sparsity=p), LibMatrixDatagen.java).  Rows [row_offset, row_offset+rows)
 * of the global matrix: a row depends on (seed, global row index) only, so shards compose.  Columns are sorted and
 * distinct within a row, values are U(-1,1) and never zero.  Two calls, like dense_to_csr: with colind == NULL the
 * row pointers (rows+1 entries) and *h_nnz are produced; the second call fills colind / val of that size.
 * No reference GPU counterpart; used by bench.py and the tests (bit-identical CPU twin in oracle/). */
int sysml_b200_rand_csr(sysml_b200_ctx* ctx, int64_t rows, int64_t cols, double mean_per_row, uint64_t seed,
                        int64_t row_offset, int32_t* rowptr, int32_t* colind, double* val, int64_t* h_nnz);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* SYSML_B200_H */
