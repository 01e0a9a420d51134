package com.mellanox.hadoop.mapred;

/** Raised by the native side (ThrowNew from csrc/jni/uda_bridge.cc) and by the plugins when neither the
 *  accelerated nor the vanilla shuffle can proceed. The class name is looked up by the native code. */
public class UdaRuntimeException extends RuntimeException {
	private static final long serialVersionUID = 1L;
	public UdaRuntimeException(String msg) { super(msg); }
	public UdaRuntimeException(String msg, Throwable cause) { super(msg, cause); }
}
