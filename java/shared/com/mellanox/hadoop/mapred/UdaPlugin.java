package com.mellanox.hadoop.mapred;

import java.io.File;
import java.io.IOException;
import java.nio.ByteBuffer;
import java.util.ArrayList;
import java.util.List;
import java.util.Timer;
import java.util.TimerTask;

import org.apache.commons.logging.Log;
import org.apache.commons.logging.LogFactory;
import org.apache.hadoop.io.DataInputBuffer;
import org.apache.hadoop.io.WritableUtils;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapred.RawKeyValueIterator;
import org.apache.hadoop.mapred.ReduceTask;
import org.apache.hadoop.mapred.Reporter;
import org.apache.hadoop.mapred.TaskAttemptID;
import org.apache.hadoop.mapred.TaskLog;
import org.apache.hadoop.util.DiskChecker;
import org.apache.hadoop.util.Progress;

/**
 * Owner of one native role. Builds the argv handed to startNative, derives the numeric log level from
 * log4j and (provider only) re-pushes it every second, and answers getConfData look-ups from C++.
 * (Surface of the reference's UdaPlugin.java:60-144.)
 */
abstract class UdaPlugin {
	protected final List<String> cmdParams = new ArrayList<String>();
	protected static Log LOG;
	protected static int logLevel = -1;
	protected static JobConf jobConf;

	UdaPlugin(JobConf conf) { jobConf = conf; }

	/** argv for the native side: "-w <wqes> -r <port> [-a <approach>] -m 1 -g <logdir> -s <buf KiB>". */
	protected abstract void buildCmdParams();

	protected static void prepareLog(String loggerName) {
		LOG = LogFactory.getLog(loggerName);
		refreshLogLevel();
	}

	/** 1..6 = fatal..trace, the numbering of the native logger. Returns true when it changed. */
	private static boolean refreshLogLevel() {
		int lvl = (LOG.isFatalEnabled() ? 1 : 0) + (LOG.isErrorEnabled() ? 1 : 0) + (LOG.isWarnEnabled() ? 1 : 0)
				+ (LOG.isInfoEnabled() ? 1 : 0) + (LOG.isDebugEnabled() ? 1 : 0) + (LOG.isTraceEnabled() ? 1 : 0);
		boolean changed = lvl != logLevel;
		logLevel = lvl;
		return changed;
	}

	static String getDataFromConf(String name, String defaultValue) {
		return jobConf.get(name, defaultValue);
	}

	protected void launchCppSide(boolean isNetMerger, UdaBridge.UdaCallable callable) {
		if (!isNetMerger) {
			// the provider lives as long as the TaskTracker / NodeManager: follow log4j level changes
			new Timer(true).schedule(new TimerTask() {
				public void run() {
					if (refreshLogLevel()) UdaBridge.setLogLevel(logLevel);
				}
			}, 0, 1000);
		}
		buildCmdParams();
		LOG.info("going to execute C++ thru JNI with argc=: " + cmdParams.size() + " cmd: " + cmdParams);
		boolean ownLogFile = jobConf.getBoolean("mapred.uda.log.to.unique.file", false);
		try {
			UdaBridge.start(isNetMerger, cmdParams.toArray(new String[0]), LOG, logLevel, ownLogFile, callable);
		} catch (UnsatisfiedLinkError e) {
			LOG.warn("UDA: cannot load libuda.so; java.library.path=" + System.getProperty("java.library.path"));
			throw e;
		}
	}
}

/**
 * Consumer runtime inside a ReduceTask (reference: UdaPlugin.java:146-556). Computes the shuffle memory
 * budget, sends INIT and FETCH commands, receives merged blocks through dataFromUda into two 1 MiB Java
 * buffers and exposes them as a RawKeyValueIterator.
 */
class UdaPluginRT<K, V> extends UdaPlugin implements UdaBridge.UdaCallable {
	static final int KV_BUF_SIZE = 1 << 20;     // one native staging block
	static final int KV_BUF_NUM = 2;
	private final UdaShuffleConsumerPluginShared<K, V> consumer;
	private final ReduceTask reduceTask;
	private final Reporter reporter;
	private final int mapsNeeded;
	private int mapsDone = 0;
	private final KVBuf[] kvBufs = new KVBuf[KV_BUF_NUM];
	private final J2CQueue<K, V> queue = new J2CQueue<K, V>();
	private final List<String> params = new ArrayList<String>();
	private Progress progress = new Progress();

	private static final class KVBuf {
		byte[] bytes = new byte[KV_BUF_SIZE];     // grows when a single record is larger than a staging block
		final DataInputBuffer in = new DataInputBuffer();
		int len = 0;
		boolean full = false;   // true: waiting for the reducer, false: waiting for native data
	}

	UdaPluginRT(UdaShuffleConsumerPluginShared<K, V> consumer, ReduceTask task, JobConf conf, Reporter reporter, int numMaps)
			throws IOException {
		super(conf);
		this.consumer = consumer;
		this.reduceTask = task;
		this.reporter = reporter;
		this.mapsNeeded = numMaps;
		prepareLog("org.apache.hadoop.mapred.ShuffleConsumerPlugin");
		for (int i = 0; i < KV_BUF_NUM; i++) kvBufs[i] = new KVBuf();

		long maxBufKb = conf.getLong("mapred.rdma.buf.size", 1024);
		long minBufKb = conf.getLong("mapred.rdma.buf.size.min", 16);
		long shuffleMem = conf.getLong("mapred.rdma.shuffle.total.size", 0);
		if (shuffleMem <= 0) {
			float pct = conf.getFloat("mapred.job.shuffle.input.buffer.percent", 0.70f);
			if (pct < 0 || pct > 1) throw new IOException("mapred.job.shuffle.input.buffer.percent must be in [0,1]");
			shuffleMem = (long) (Runtime.getRuntime().maxMemory() * pct);
		}
		LOG.info("UDA: shuffle memory = " + shuffleMem + " bytes, buffer " + minBufKb + ".." + maxBufKb + " KiB");

		launchCppSide(true, this);

		TaskAttemptID id = task.getTaskID();
		params.clear();
		params.add(Integer.toString(numMaps));
		params.add(id.getJobID().toString());
		params.add(id.toString());
		params.add(conf.get("mapred.netmerger.hybrid.lpq.size", "0"));
		params.add(Long.toString(maxBufKb * 1024));
		params.add(Long.toString(minBufKb * 1024));
		params.add(conf.getOutputKeyClass().getName());
		String codec = conf.getCompressMapOutput() ? conf.get("mapred.map.output.compression.codec", null) : null;
		params.add(codec);   // literal "null" when map output is not compressed
		String blockSize = Integer.toString(256 * 1024);
		if (codec != null) {
			if (codec.contains("lzo.LzoCodec")) blockSize = conf.get("io.compression.codec.lzo.buffersize", blockSize);
			else if (codec.contains("SnappyCodec")) blockSize = conf.get("io.compression.codec.snappy.buffersize", blockSize);
		}
		params.add(blockSize);
		params.add(Long.toString(shuffleMem));
		List<String> dirs = new ArrayList<String>();
		for (String d : conf.getLocalDirs()) {
			try {
				DiskChecker.checkDir(new File(d.trim()));
				dirs.add(d.trim());
			} catch (DiskChecker.DiskErrorException ignored) {
			}
		}
		params.add(Integer.toString(dirs.size()));
		params.addAll(dirs);
		LOG.info("UDA: sending INIT_COMMAND");
		UdaBridge.doCommand(UdaCmd.formCmd(UdaCmd.INIT_COMMAND, params));
		progress.set(0.5f);
	}

	protected void buildCmdParams() {
		cmdParams.clear();
		cmdParams.add("-w"); cmdParams.add(jobConf.get("mapred.rdma.wqe.per.conn", "256"));
		cmdParams.add("-r"); cmdParams.add(jobConf.get("mapred.rdma.cma.port", "9011"));
		cmdParams.add("-a"); cmdParams.add(jobConf.get("mapred.netmerger.merge.approach", "1"));
		cmdParams.add("-m"); cmdParams.add("1");
		String stdout = TaskLog.getTaskLogFile(reduceTask.getTaskID(), false, TaskLog.LogName.STDOUT).toString();
		cmdParams.add("-g"); cmdParams.add(stdout.substring(0, stdout.lastIndexOf('/') + 1));
		cmdParams.add("-s"); cmdParams.add(jobConf.get("mapred.rdma.buf.size", "1024"));
	}

	/** One completed map output: "host:jobid:mapAttemptId:myPartition". */
	void sendFetchReq(String host, String jobId, String mapAttemptId) {
		params.clear();
		params.add(host);
		params.add(jobId);
		params.add(mapAttemptId);
		params.add(Integer.toString(reduceTask.getPartition()));
		UdaBridge.doCommand(UdaCmd.formCmd(UdaCmd.FETCH_COMMAND, params));
	}

	void close() {
		UdaBridge.reduceExitMsg();
		queue.close();
	}

	RawKeyValueIterator createKVIterator() { return queue; }

	// ---- up-calls -------------------------------------------------------------------------------
	/** Native side reports progress every 20 fetched maps and once at the end. */
	public void fetchOverMessage() {
		mapsDone = Math.min(mapsNeeded, mapsDone + 20);
		reporter.progress();
		if (mapsDone >= mapsNeeded) consumer.notifyFetchCompleted();
	}

	/** A block of whole records arrived in the native staging buffer (a direct ByteBuffer). */
	public void dataFromUda(Object directBuffer, int len) throws Throwable {
		KVBuf buf = kvBufs[queue.producerIndex()];
		synchronized (buf) {
			while (buf.full) buf.wait();
			ByteBuffer src = (ByteBuffer) directBuffer;
			src.position(0);
			if (len > buf.bytes.length) buf.bytes = new byte[len];   // a record larger than 1 MiB arrives as one block
			src.get(buf.bytes, 0, len);
			buf.len = len;
			buf.in.reset(buf.bytes, 0, len);
			buf.full = true;
			buf.notifyAll();
		}
		queue.produced();
	}

	public void failureInUda() {
		consumer.failureInUda(new UdaRuntimeException("Uda Failure in C++"));
	}

	/** Reads "VInt klen, VInt vlen, key, val" records out of the two buffers; negative lengths end the stream. */
	private final class J2CQueue<K2, V2> implements RawKeyValueIterator {
		private final DataInputBuffer key = new DataInputBuffer(), val = new DataInputBuffer();
		private int prod = 0, cons = 0, records = 0;
		private KVBuf cur = null;
		int producerIndex() { return prod; }
		void produced() { prod = (prod + 1) % KV_BUF_NUM; }
		public DataInputBuffer getKey() { return key; }
		public DataInputBuffer getValue() { return val; }
		public Progress getProgress() { return progress; }
		public void close() {}

		public boolean next() throws IOException {
			for (;;) {
				if (cur == null) {
					cur = kvBufs[cons];
					synchronized (cur) {
						try {
							while (!cur.full) cur.wait();
						} catch (InterruptedException e) {
							throw new IOException(e);
						}
					}
				}
				DataInputBuffer in = cur.in;
				if (in.getPosition() < cur.len) {
					int klen = WritableUtils.readVInt(in), vlen = WritableUtils.readVInt(in);
					if (klen < 0 || vlen < 0) return false;   // EOF marker
					int pos = in.getPosition();
					key.reset(cur.bytes, pos, klen);
					val.reset(cur.bytes, pos + klen, vlen);
					in.skip(klen + vlen);
					if (++records % 1000 == 0) reporter.progress();
					return true;
				}
				synchronized (cur) {   // buffer drained: hand it back to the native side
					cur.full = false;
					cur.notifyAll();
				}
				cur = null;
				cons = (cons + 1) % KV_BUF_NUM;
			}
		}
	}
}
