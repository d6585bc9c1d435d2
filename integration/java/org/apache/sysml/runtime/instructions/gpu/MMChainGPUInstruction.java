/*
 * MMChainGPUInstruction -- t(X) %*% (X %*% v), t(X) %*% (w * (X %*% v)), t(X) %*% ((X %*% v) - w) on the GPU.
 *
 * NEW file (the reference has no GPU twin of MMChainCPInstruction: AggBinaryOp.isGPUEnabled returns false for
 * MAPMM_CHAIN, hops/AggBinaryOp.java:159-162).  Instruction format as the CP one (lops/MapMultChain.java:149-186,
 * runtime/instructions/cp/MMChainCPInstruction.java:45-67):
 *     GPU°mmchain°X°v°out°XtXv[°k]          GPU°mmchain°X°v°w°out°XtwXv|XtXvy[°k]
 * k (the CP thread count) is accepted and ignored.  Registered by patches/0001-GPUInstructionParser-mmchain-uarimax.diff.
 */
package org.apache.sysml.runtime.instructions.gpu;

import org.apache.sysml.lops.MapMultChain.ChainType;
import org.apache.sysml.runtime.DMLRuntimeException;
import org.apache.sysml.runtime.controlprogram.caching.MatrixObject;
import org.apache.sysml.runtime.controlprogram.context.ExecutionContext;
import org.apache.sysml.runtime.instructions.InstructionUtils;
import org.apache.sysml.runtime.instructions.cp.CPOperand;
import org.apache.sysml.runtime.matrix.data.LibMatrixB200Ops;
import org.apache.sysml.runtime.matrix.operators.Operator;
import org.apache.sysml.utils.GPUStatistics;

public class MMChainGPUInstruction extends GPUInstruction {
	private final ChainType _type;
	private final CPOperand _X, _v, _w, _out;

	private MMChainGPUInstruction(CPOperand X, CPOperand v, CPOperand w, CPOperand out, ChainType type, String opcode, String istr) {
		super(new Operator(true), opcode, istr);
		_gputype = GPUINSTRUCTION_TYPE.MMChain;
		_X = X;
		_v = v;
		_w = w;
		_out = out;
		_type = type;
	}

	public static MMChainGPUInstruction parseInstruction(String str) {
		String[] parts = InstructionUtils.getInstructionPartsWithValueType(str);
		String opcode = parts[0];
		if( !opcode.equalsIgnoreCase("mmchain") )
			throw new DMLRuntimeException("Unknown opcode while parsing an MMChainGPUInstruction: " + str);
		// fields after the opcode: X, v, [w,] out, type, [k]
		int n = parts.length - 1;
		boolean hasK = n >= 5 && Character.isDigit(parts[n].charAt(0));
		int nf = hasK ? n - 1 : n;
		if( nf == 4 ) {
			ChainType type = ChainType.valueOf(parts[4]);
			if( type != ChainType.XtXv )
				throw new DMLRuntimeException("ChainType " + type + " needs a weight vector: " + str);
			return new MMChainGPUInstruction(new CPOperand(parts[1]), new CPOperand(parts[2]), null, new CPOperand(parts[3]), type, opcode, str);
		}
		else if( nf == 5 ) {
			ChainType type = ChainType.valueOf(parts[5]);
			if( type == ChainType.XtXv )
				throw new DMLRuntimeException("ChainType XtXv takes no weight vector: " + str);
			return new MMChainGPUInstruction(new CPOperand(parts[1]), new CPOperand(parts[2]), new CPOperand(parts[3]), new CPOperand(parts[4]), type, opcode, str);
		}
		throw new DMLRuntimeException("mmchain expects 4 or 5 fields (+ optional k): " + str);
	}

	@Override
	public void processInstruction(ExecutionContext ec) {
		GPUStatistics.incrementNoOfExecutedGPUInst();
		MatrixObject X = getMatrixInputForGPUInstruction(ec, _X.getName());
		MatrixObject v = getMatrixInputForGPUInstruction(ec, _v.getName());
		MatrixObject w = (_w != null) ? getMatrixInputForGPUInstruction(ec, _w.getName()) : null;
		if( v.getNumRows() * v.getNumColumns() != X.getNumColumns() )
			throw new DMLRuntimeException("mmchain: v must have ncol(X) cells");
		if( w != null && w.getNumRows() * w.getNumColumns() != X.getNumRows() )
			throw new DMLRuntimeException("mmchain: w must have nrow(X) cells");
		ec.setMetaData(_out.getName(), X.getNumColumns(), 1);
		LibMatrixB200Ops.mmchain(ec, ec.getGPUContext(0), getExtendedOpcode(), X, v, w, _out.getName(), _type);
		ec.releaseMatrixInputForGPUInstruction(_X.getName());
		ec.releaseMatrixInputForGPUInstruction(_v.getName());
		if( _w != null )
			ec.releaseMatrixInputForGPUInstruction(_w.getName());
		ec.releaseMatrixOutputForGPUInstruction(_out.getName());
	}

	public ChainType getMMChainType() {
		return _type;
	}
}
