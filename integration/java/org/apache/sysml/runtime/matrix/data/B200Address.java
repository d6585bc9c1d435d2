/*
 * B200Address -- the single adapter between the reference's device-state types and the 64-bit values the natives take.
 * NEW file.
 *
 *  - of(Pointer): the device address behind a jcuda.Pointer.  JCuda keeps it in NativePointerObject.nativePointer, which
 *    has no public accessor; the supported route is a subclass living next to jcuda.Pointer.  When the JCuda dependency is
 *    dropped altogether (the goal of the swap), GPUObject / CSRPointer store plain longs and this method disappears.
 *  - context(GPUContext): the sysml_b200_ctx* of a GPUContext, created lazily on the thread that owns the context
 *    (one context per host thread and device, GPUContext.java:181-184) and destroyed with it (GPUContext.destroy :416-437).
 */
package org.apache.sysml.runtime.matrix.data;

import java.util.concurrent.ConcurrentHashMap;

import org.apache.sysml.runtime.DMLRuntimeException;
import org.apache.sysml.runtime.instructions.gpu.context.GPUContext;

import jcuda.Pointer;

public final class B200Address {
	private static final ConcurrentHashMap<GPUContext, Long> CONTEXTS = new ConcurrentHashMap<>();

	private B200Address() {}

	/** jcuda.Pointer subclass whose only purpose is to read the protected native address of another pointer. */
	private static final class Peek extends Pointer {
		Peek(Pointer other) {
			super(other);
		}
		long address() {
			return getNativePointer() + getByteOffset();
		}
	}

	public static long of(Pointer p) {
		return (p == null) ? 0L : new Peek(p).address();
	}

	public static long context(GPUContext gCtx) {
		return CONTEXTS.computeIfAbsent(gCtx, g -> {
			long[] out = new long[1];
			int rc = LibMatrixB200.ctxCreate(g.getDeviceNum(), out);
			if( rc != LibMatrixB200.OK || out[0] == 0 )
				throw new DMLRuntimeException("GPU : cannot create the native context on device " + g.getDeviceNum() + ": " + LibMatrixB200.lastError(0L));
			return out[0];
		});
	}

	/** Call from GPUContext.destroy(). */
	public static void release(GPUContext gCtx) {
		Long c = CONTEXTS.remove(gCtx);
		if( c != null )
			LibMatrixB200.ctxDestroy(c);
	}
}
