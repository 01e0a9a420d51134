package com.mellanox.hadoop.mapred;

import java.io.File;
import java.net.URLDecoder;

import org.apache.commons.logging.Log;

/**
 * Java half of the JNI boundary. The native half is csrc/jni/uda_bridge.cc in libuda.so, which looks
 * these static methods up by name and signature in JNI_OnLoad, so none of the names or signatures below
 * may change:
 *
 *   natives   startNative(Z[Ljava/lang/String;IZ)I, doCommandNative(Ljava/lang/String;)V,
 *             reduceExitMsgNative()V, setLogLevelNative(I)V
 *   up-calls  fetchOverMessage()V, dataFromUda(Ljava/lang/Object;I)V,
 *             getPathUda(Ljava/lang/String;Ljava/lang/String;I)Ljava/lang/Object;,
 *             getConfData(Ljava/lang/String;Ljava/lang/String;)Ljava/lang/String;,
 *             logToJava(Ljava/lang/String;I)V, failureInUda()V
 *
 * (same surface as the reference's plugins/shared/com/mellanox/hadoop/mapred/UdaBridge.java:38-145)
 */
public class UdaBridge {

	/** What the owner of the native side must implement: the targets of the up-calls. */
	public interface UdaCallable {
		void fetchOverMessage();
		void dataFromUda(Object directBuffer, int len) throws Throwable;
		void failureInUda();
	}

	private static UdaCallable owner;
	private static Log log;

	static {
		// libuda.so sits next to the jar that holds this class
		String dir = ".";
		try {
			String jar = UdaBridge.class.getProtectionDomain().getCodeSource().getLocation().getPath();
			dir = new File(URLDecoder.decode(jar, "UTF-8")).getParent();
		} catch (Exception ignored) {
		}
		System.load(dir + File.separator + "libuda.so");
	}

	private static native int startNative(boolean isNetMerger, String[] args, int logLevel, boolean logToUdaFile);
	private static native void doCommandNative(String s);
	private static native void reduceExitMsgNative();
	private static native void setLogLevelNative(int logLevel);

	static void start(boolean isNetMerger, String[] args, Log logger, int logLevel, boolean logToUdaFile, UdaCallable callable) {
		log = logger;
		owner = callable;
		log.info(" +++>>> invoking UdaBridge.startNative: isNetMerger=" + isNetMerger);
		int rc = startNative(isNetMerger, args, logLevel, logToUdaFile);
		log.info(" <<<+++ after UdaBridge.startNative ret=" + rc);
	}

	static void doCommand(String cmd) {
		if (log.isDebugEnabled()) log.debug(" +++>>> invoking UdaBridge.doCommandNative: " + cmd);
		doCommandNative(cmd);
	}

	static void reduceExitMsg() { reduceExitMsgNative(); }

	static void setLogLevel(int level) { setLogLevelNative(level); }

	// ------------------------------------------------------------------ up-calls from C++
	static void fetchOverMessage() throws Throwable {
		owner.fetchOverMessage();
	}

	static void dataFromUda(Object directBuffer, int len) throws Throwable {
		owner.dataFromUda(directBuffer, len);
	}

	/** Provider side: resolve (job, map, reduce) to the MOF path and its index record. */
	static Object getPathUda(String jobId, String mapId, int reduceId) {
		return UdaShuffleProviderPluginShared.getPathIndex(jobId, mapId, reduceId);
	}

	static String getConfData(String name, String defaultValue) {
		return UdaPlugin.getDataFromConf(name, defaultValue);
	}

	/** severity: 1 fatal, 2 error, 3 warn, 4 info, 5 debug, 6 trace (C++ log_severity_t). */
	static void logToJava(String msg, int severity) {
		switch (severity) {
		case 1: log.fatal(msg); break;
		case 2: log.error(msg); break;
		case 3: log.warn(msg); break;
		case 4: log.info(msg); break;
		case 5: log.debug(msg); break;
		case 6: log.trace(msg); break;
		default: log.info(msg); break;
		}
	}

	static void failureInUda() {
		owner.failureInUda();
	}
}
