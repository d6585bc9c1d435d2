/*
 * LibMatrixB200Ops -- the LibMatrixCUDA-shaped entry points (ExecutionContext, GPUContext, instName, MatrixObjects,
 * outputName, flags) for the operations whose bodies move from JCuda / cuBLAS / cuSPARSE calls to LibMatrixB200 natives.
 * NEW file.  Each method states the reference method it stands in for; argument meaning, locking discipline (inputs are
 * acquired by the instruction, outputs through getDenseMatrixOutputForGPUInstruction) and error behaviour
 * (DMLRuntimeException) are the reference's.
 *
 * Device addresses: GPUObject keeps its blocks as jcuda.Pointer (GPUObject.java:52-99).  B200Address.of(Pointer) is the
 * one place that turns such a pointer into the 64-bit address the natives take (see integration/README.md, "addresses").
 * int64 extents end to end: the toInt() guards of the reference (LibMatrixCUDA.java:685, 1735-1740) do not apply here.
 */
package org.apache.sysml.runtime.matrix.data;

import org.apache.sysml.lops.MapMultChain.ChainType;
import org.apache.sysml.runtime.DMLRuntimeException;
import org.apache.sysml.runtime.controlprogram.caching.MatrixObject;
import org.apache.sysml.runtime.controlprogram.context.ExecutionContext;
import org.apache.sysml.runtime.instructions.gpu.context.CSRPointer;
import org.apache.sysml.runtime.instructions.gpu.context.GPUContext;
import org.apache.sysml.runtime.instructions.gpu.context.GPUObject;

public class LibMatrixB200Ops extends LibMatrixCUDA {

	/** The native context of a GPUContext (created next to cudaSetDevice in GPUContext.initializeThread, GPUContext.java:181-184). */
	private static long ctx(GPUContext gCtx) {
		return B200Address.context(gCtx);
	}

	private static long dense(GPUContext gCtx, MatrixObject mo, String instName) {
		return B200Address.of(getDensePointer(gCtx, mo, instName));
	}

	/**
	 * mmchain on the GPU; no reference GPU method exists (CPU: LibMatrixMult.matrixMultChain, LibMatrixMult.java:243-329).
	 * X may be dense or CSR; v and w are densified as getDensePointer does for every operator without a sparse form.
	 */
	public static void mmchain(ExecutionContext ec, GPUContext gCtx, String instName, MatrixObject X, MatrixObject v, MatrixObject w,
			String outputName, ChainType type) {
		long n = X.getNumRows(), m = X.getNumColumns();
		int ct = (type == ChainType.XtXv) ? LibMatrixB200.CHAIN_XtXv : (type == ChainType.XtwXv) ? LibMatrixB200.CHAIN_XtwXv : LibMatrixB200.CHAIN_XtXvy;
		if( type != ChainType.XtXv && w == null )
			throw new DMLRuntimeException("GPU : mmchain " + type + " needs a weight vector");
		MatrixObject out = getDenseMatrixOutputForGPUInstruction(ec, instName, outputName, m, 1);
		long c = ctx(gCtx);
		long pv = dense(gCtx, v, instName);
		long pw = (w != null) ? dense(gCtx, w, instName) : 0L;
		long po = dense(gCtx, out, instName);
		GPUObject gx = X.getGPUObject(gCtx);
		if( isInSparseFormat(gCtx, X) ) {
			CSRPointer csr = gx.getJcudaSparseMatrixPtr();
			LibMatrixB200.check(c, LibMatrixB200.mmchainCsr(c, B200Address.of(csr.rowPtr), B200Address.of(csr.colInd), B200Address.of(csr.val), pv, pw, po, n, m, ct));
		}
		else {
			LibMatrixB200.check(c, LibMatrixB200.mmchainDense(c, B200Address.of(gx.getJcudaDenseMatrixPtr()), pv, pw, po, n, m, ct));
		}
	}

	/**
	 * Stands in for LibMatrixCUDA.matmultTSMM (LibMatrixCUDA.java:477-524: cublasDsyrk + copy_u2l_dense); also sets the
	 * exact nnz of the result as the CPU path does (copyUpperToLowerTriangle, LibMatrixMult.java:352-353).
	 */
	public static void matmultTSMM(ExecutionContext ec, GPUContext gCtx, String instName, MatrixObject left, String outputName, boolean isLeftTransposed) {
		long rows = left.getNumRows(), cols = left.getNumColumns();
		long dim = isLeftTransposed ? cols : rows;
		MatrixObject output = getDenseMatrixOutputForGPUInstruction(ec, instName, outputName, dim, dim);
		long c = ctx(gCtx);
		long[] nnz = new long[1];
		GPUObject g = left.getGPUObject(gCtx);
		if( isInSparseFormat(gCtx, left) ) {
			CSRPointer csr = g.getJcudaSparseMatrixPtr();
			long rp = B200Address.of(csr.rowPtr), ci = B200Address.of(csr.colInd), va = B200Address.of(csr.val), po = dense(gCtx, output, instName);
			LibMatrixB200.check(c, isLeftTransposed ? LibMatrixB200.tsmmCsrLeft(c, rp, ci, va, po, rows, cols, nnz) : LibMatrixB200.tsmmCsrRight(c, rp, ci, va, po, rows, cols, nnz));
		}
		else {
			LibMatrixB200.check(c, LibMatrixB200.tsmmDense(c, B200Address.of(g.getJcudaDenseMatrixPtr()), dense(gCtx, output, instName), rows, cols, isLeftTransposed ? 1 : 0, nnz));
		}
		output.getMatrixCharacteristics().setNonZeros(nnz[0]);
	}

	/**
	 * rowIndexMax / rowIndexMin (uarimax / uarimin): new on the GPU (not in GPUInstructionParser.java:111-133); CPU twin
	 * LibMatrixAgg d_uarimxx / d_uarimin (LibMatrixAgg.java:1902-1927): 1-based index of the LAST extremum, NaN never wins.
	 */
	public static void rowIndexExtremum(ExecutionContext ec, GPUContext gCtx, String instName, MatrixObject in, String outputName, boolean max) {
		long rows = in.getNumRows(), cols = in.getNumColumns();
		MatrixObject out = getDenseMatrixOutputForGPUInstruction(ec, instName, outputName, rows, 1);
		long c = ctx(gCtx);
		int op = max ? LibMatrixB200.AGG_MAXINDEX : LibMatrixB200.AGG_MININDEX;
		if( isInSparseFormat(gCtx, in) ) {
			CSRPointer csr = in.getGPUObject(gCtx).getJcudaSparseMatrixPtr();
			LibMatrixB200.check(c, LibMatrixB200.aggUnaryCsr(c, B200Address.of(csr.rowPtr), B200Address.of(csr.colInd), B200Address.of(csr.val), rows, cols, op,
				LibMatrixB200.DIR_ROW, dense(gCtx, out, instName)));
		}
		else {
			LibMatrixB200.check(c, LibMatrixB200.aggUnaryDense(c, dense(gCtx, in, instName), rows, cols, op, LibMatrixB200.DIR_ROW, dense(gCtx, out, instName)));
		}
	}
}
