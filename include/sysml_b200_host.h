/*
 * sysml_b200_host.h -- C ABI of the host-side runtime that mirrors SystemML's GPU execution layer:
 * instruction strings in, matrices out.  It is the C++ counterpart (the reference's toolchain, a JDK,
 * is absent from the build image) of
 *   - GPUInstructionParser.parseSingleInstruction   (J/runtime/instructions/GPUInstructionParser.java:147-205)
 *   - the GPU instruction classes' processInstruction (J/runtime/instructions/gpu/ *GPUInstruction.java)
 *   - ExecutionContext.*ForGPUInstruction            (J/runtime/controlprogram/context/ExecutionContext.java:304-452)
 *   - GPUObject (locks, dirty bit, lazy H2D/D2H)     (J/runtime/instructions/gpu/context/GPUObject.java:528-946)
 *   - GPUMemoryManager (free list by size, eviction) (J/runtime/instructions/gpu/context/GPUMemoryManager.java:157-322)
 * with J/ = src/main/java/org/apache/sysml/.  All numeric work is delegated to the kernels behind
 * include/sysml_b200.h; this layer only does what the Java classes do: parse, look up variables, move
 * data lazily, set output metadata, call the library, release.
 *
 * Instruction strings are exactly what SystemML's compiler emits (lops/Lop.java:82-86): fields separated
 * by OPERAND_DELIMITOR (U+00B0), operands written name<U+00B7>DATATYPE<U+00B7>VALUETYPE[<U+00B7>literal], e.g.
 *   GPU°ba+*°A·MATRIX·DOUBLE°B·MATRIX·DOUBLE°C·MATRIX·DOUBLE°false°false
 *   GPU°tsmm°X·MATRIX·DOUBLE°C·MATRIX·DOUBLE°LEFT
 *   GPU°mmchain°X·MATRIX·DOUBLE°v·MATRIX·DOUBLE°w·MATRIX·DOUBLE°q·MATRIX·DOUBLE°XtwXv      (new on GPU)
 *   GPU°uark+°X·MATRIX·DOUBLE°r·MATRIX·DOUBLE      GPU°uak+°X·MATRIX·DOUBLE°s·SCALAR·DOUBLE
 *   GPU°*°A·MATRIX·DOUBLE°2.5·SCALAR·DOUBLE·true°C·MATRIX·DOUBLE
 *   GPU°+*°A·MATRIX·DOUBLE°0.5·SCALAR·DOUBLE·true°B·MATRIX·DOUBLE°C·MATRIX·DOUBLE
 * ASCII stand-ins are accepted too: ';' for the operand delimiter when no U+00B0 occurs in the string and
 * ':' for U+00B7 likewise.
 *
 * Every function returns 0 or a negative SYSML_E_* code (sysml_b200.h); sysml_b200_ec_last_error() has
 * the message -- the text a DMLRuntimeException would carry.
 */
#ifndef SYSML_B200_HOST_H
#define SYSML_B200_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

typedef struct sysml_b200_ec sysml_b200_ec;

/* mem_util_factor: sysml.gpu.memory.util.factor (DMLConfig.java:138; 0.9 by default; <= 0 selects 0.9).
 * eager_cuda_free : sysml.gpu.eager.cudaFree (false by default: freed pointers go to a size-keyed free list).
 * memory_limit    : > 0 caps the device memory this context may hold (bytes) -- lets tests exercise the
 *                   eviction path on a 180 GB card; 0 = use cudaMemGetInfo * factor. */
int sysml_b200_ec_create(int device, double mem_util_factor, int eager_cuda_free, int64_t memory_limit,
                         sysml_b200_ec** out);
int sysml_b200_ec_destroy(sysml_b200_ec* ec);
const char* sysml_b200_ec_last_error(sysml_b200_ec* ec);

/* Variables (the symbol table of ExecutionContext).  Host data are copied; nothing goes to the device
 * until a GPU instruction acquires the variable (GPUObject.acquireDeviceRead, GPUObject.java:528-545). */
int sysml_b200_ec_set_matrix_dense(sysml_b200_ec* ec, const char* name, const double* h_data, int64_t rows,
                                   int64_t cols);
int sysml_b200_ec_set_matrix_csr(sysml_b200_ec* ec, const char* name, const int32_t* h_rowptr,
                                 const int32_t* h_colind, const double* h_val, int64_t rows, int64_t cols);
/* Binds a dense matrix that already lives on the device (row-major, on this context's device).  The block
 * stays owned by the caller: it is never freed or evicted here (a GPUObject whose pointer came from outside). */
int sysml_b200_ec_set_matrix_device(sysml_b200_ec* ec, const char* name, double* d_data, int64_t rows,
                                    int64_t cols);
int sysml_b200_ec_set_scalar(sysml_b200_ec* ec, const char* name, double value);
int sysml_b200_ec_remove(sysml_b200_ec* ec, const char* name); /* rmvar */

/* Parses and runs ONE instruction (ProgramBlock.executeSingleInstruction, ProgramBlock.java:241-255). */
int sysml_b200_ec_execute(sysml_b200_ec* ec, const char* instruction);
/* Parse only (no device needed): writes "TYPE opcode in1 in2 ... out" into buf.  ec may be NULL. */
int sysml_b200_parse_instruction(const char* instruction, char* buf, int buflen);

int sysml_b200_ec_get_matrix_dims(sysml_b200_ec* ec, const char* name, int64_t* rows, int64_t* cols,
                                  int64_t* nnz);
/* CacheableData.acquireRead -> GPUObject.acquireHostRead: copies device -> host only when the device copy
 * is dirty (GPUObject.java:593-607). */
int sysml_b200_ec_get_matrix_dense(sysml_b200_ec* ec, const char* name, double* h_out);
int sysml_b200_ec_get_scalar(sysml_b200_ec* ec, const char* name, double* value);
/* GPUObject.isSparse (GPUObject.java:416-418): *is_sparse = 1 when the variable's device block is CSR, 0 when it is dense,
 * -1 when it has no device block.  Cellwise instructions keep a CSR operand CSR where the CPU back-end's rules say the
 * operator is sparse-safe (LibMatrixBincell.java:116-119, 195-248; MatrixBlock.estimateSparsityOnBinary :2440-2485) and
 * convert it to dense otherwise, as LibMatrixCUDA does for every cellwise operator (LibMatrixCUDA.java:1360-1383). */
int sysml_b200_ec_is_sparse_on_device(sysml_b200_ec* ec, const char* name, int* is_sparse);

/* sysml.gpu.eviction.policy (DMLScript.EvictionPolicy, api/DMLScript.java:110,175,679): "lru" (default), "lfu",
 * "min_evict" or "mru".  Decides which unlocked matrices leave the device first when an allocation does not fit
 * (GPUMemoryManager.java:229-246 with GPUComparator :531-571; timestamps follow GPUObject.updateReleaseLocks,
 * GPUObject.java:650-667). */
int sysml_b200_ec_set_eviction_policy(sysml_b200_ec* ec, const char* policy);
/* MatrixBlock.evalSparseFormatInMemory(rows, cols, nnz) (MatrixBlock.java:1001-1015): 1 when a block with that many
 * non-zeros is kept in sparse format, 0 when dense.  Pure host arithmetic, no context needed. */
int sysml_b200_eval_sparse_format_in_memory(int64_t rows, int64_t cols, int64_t nnz);
/* The rule by which a cellwise instruction keeps a CSR operand CSR (pure host logic, no device needed): kind 0 = matrix op
 * scalar (ScalarOperator.java:53-59, 84-89; RightScalarOperator.java:42-49; LeftScalarOperator.java:41-46 -- except that
 * Power with a negative constant, which the reference lists although 0^c is infinite, takes the dense form here),
 * kind 1 = matrix op matrix (BinaryOperator.java:60-66), kind 2 = unary builtin (UnaryOperator.java:38-47).
 * opcode as in the instruction string ("*", "/", "^2", "max", "sqrt", ...).  Returns 1 / 0, or -1 for a bad kind. */
int sysml_b200_cellwise_sparse_safe(const char* opcode, int kind, double constant, int scalar_left);
/* MatrixBlock.examSparsity on the host copy of a variable (fetched from the device first when that copy is the
 * valid one), as LibMatrixMult applies it to every matmult result (LibMatrixMult.java:113,149,168,225): the block
 * is rewritten as CSR when evalSparseFormatInMemory says so.  *is_sparse receives the resulting format; the exact
 * nnz is what sysml_b200_ec_get_matrix_dims returns afterwards. */
int sysml_b200_ec_exam_sparsity(sysml_b200_ec* ec, const char* name, int* is_sparse);

/* GPUStatistics counters (J/utils/GPUStatistics.java:40-62), in this order:
 * [0] executed GPU instructions, [1] H2D copies, [2] D2H copies, [3] cudaMalloc calls, [4] free-list reuses,
 * [5] cudaFree calls, [6] evictions, [7] bytes currently allocated, [8] bytes in the free list. */
int sysml_b200_ec_stats(sysml_b200_ec* ec, int64_t* out9);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* SYSML_B200_HOST_H */
