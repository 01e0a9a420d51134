package com.mellanox.hadoop.mapred;

import java.io.IOException;
import java.util.HashSet;
import java.util.Set;

import org.apache.commons.logging.Log;
import org.apache.commons.logging.LogFactory;
import org.apache.hadoop.fs.FileSystem;
import org.apache.hadoop.io.IntWritable;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapred.MapTaskCompletionEventsUpdate;
import org.apache.hadoop.mapred.RawKeyValueIterator;
import org.apache.hadoop.mapred.ReduceTask;
import org.apache.hadoop.mapred.Reporter;
import org.apache.hadoop.mapred.ShuffleConsumerPlugin;
import org.apache.hadoop.mapred.TaskAttemptID;
import org.apache.hadoop.mapred.TaskCompletionEvent;
import org.apache.hadoop.mapred.TaskID;
import org.apache.hadoop.mapred.TaskUmbilicalProtocol;
import org.apache.hadoop.mapred.UdaMapredBridge;
import org.apache.hadoop.util.StringUtils;

/** What a Hadoop-flavour adapter must provide to the version-neutral consumer logic. */
interface UdaConsumerPluginCallable {
	boolean pluginFetchOutputs(ShuffleConsumerPlugin plugin) throws IOException;
	RawKeyValueIterator pluginCreateKVIterator(ShuffleConsumerPlugin plugin, JobConf job, FileSystem fs, Reporter reporter)
			throws IOException, InterruptedException;
	Class getVanillaPluginClass();
	MapTaskCompletionEventsUpdate pluginGetMapCompletionEvents(IntWritable fromEventId, int maxEventsToFetch) throws IOException;
}

/**
 * Hadoop-version-neutral consumer: polls map-completion events, turns them into FETCH commands and, on any
 * failure that happens before merged data reached the reducer, hands the whole shuffle back to the vanilla
 * plugin (reference: UdaShuffleConsumerPluginShared.java:115-603). The two log lines
 * "init - Using UdaShuffleConsumerPlugin" and "====XXX Successfully closed UdaShuffleConsumerPlugin XXX===="
 * are counted by deployment automation and must keep their wording.
 */
class UdaShuffleConsumerPluginShared<K, V> {
	static final Log LOG = LogFactory.getLog(ShuffleConsumerPlugin.class.getCanonicalName());
	private static final int MAX_EVENTS_TO_FETCH = 10000;

	private final UdaConsumerPluginCallable flavour;
	private ReduceTask reduceTask;
	private TaskAttemptID reduceId;
	private TaskUmbilicalProtocol umbilical;
	private JobConf jobConf;
	private Reporter reporter;
	private FileSystem fs;
	private UdaPluginRT<K, V> rdmaChannel;
	private ShuffleConsumerPlugin fallbackPlugin;
	private boolean fallbackFetchDone = false;
	private volatile boolean stopEvents = false;
	private final Object fetchLock = new Object();
	private boolean fetchCompleted = false;

	UdaShuffleConsumerPluginShared(UdaConsumerPluginCallable flavour) { this.flavour = flavour; }

	// ---- ShuffleConsumerPlugin surface ------------------------------------------------------------
	public void init(ReduceTask task, TaskUmbilicalProtocol umbilical, JobConf conf, Reporter reporter, FileSystem fs) throws IOException {
		this.fs = fs;
		try {
			LOG.info("init - Using UdaShuffleConsumerPlugin");
			this.reduceTask = task;
			this.reduceId = task.getTaskID();
			this.umbilical = umbilical;
			this.jobConf = conf;
			this.reporter = reporter;
			this.rdmaChannel = new UdaPluginRT<K, V>(this, task, conf, reporter, task.getNumMaps());
		} catch (Throwable t) {
			doFallbackInit(t);   // unsupported key class, codec, buffers too small, no native library, ...
		}
	}

	public boolean fetchOutputs() throws IOException {
		try {
			if (fallbackPlugin == null) return fetchOutputsInternal();
		} catch (Throwable t) {
			LOG.error("UDA fetch failed, falling back to the vanilla shuffle: " + StringUtils.stringifyException(t));
			doFallbackInit(t);
		}
		return doFallbackFetchOutputs();
	}

	public RawKeyValueIterator createKVIterator(JobConf job, FileSystem fs, Reporter reporter) throws IOException, InterruptedException {
		if (fallbackPlugin != null) return flavour.pluginCreateKVIterator(fallbackPlugin, job, fs, reporter);
		return rdmaChannel.createKVIterator();
	}

	public Throwable getMergeThrowable() { return null; }

	public void close() {
		stopEvents = true;
		if (fallbackPlugin == null) {
			rdmaChannel.close();
			LOG.info("====XXX Successfully closed UdaShuffleConsumerPlugin XXX====");
			return;
		}
		// in fallback the native side may be wedged: give it one second on a side thread, then move on
		Thread closer = new Thread("UdaCloserThread") {
			public void run() { if (rdmaChannel != null) rdmaChannel.close(); }
		};
		closer.setDaemon(true);
		closer.start();
		try { closer.join(1000); } catch (InterruptedException ignored) { }
		LOG.info("====XXX closed UdaShuffleConsumerPlugin after fallback XXX====");
	}

	// ---- called by UdaPluginRT ----------------------------------------------------------------------
	void notifyFetchCompleted() {
		synchronized (fetchLock) { fetchCompleted = true; fetchLock.notifyAll(); }
	}

	void failureInUda(Throwable t) {
		try {
			doFallbackInit(t);
			synchronized (fetchLock) { fetchLock.notifyAll(); }
		} catch (Throwable t2) {
			throw new UdaRuntimeException("Failure in UDA and failure when trying to fallback to vanilla", t2);
		}
	}

	// ---- fallback machinery ---------------------------------------------------------------------------
	private synchronized void doFallbackInit(Throwable t) throws IOException {
		if (fallbackPlugin != null) return;
		stopEvents = true;
		if (jobConf != null && jobConf.getBoolean("mapred.rdma.developer.mode", false)) {
			LOG.fatal("UDA fatal error under mapred.rdma.developer.mode: aborting instead of falling back\n"
					+ (t == null ? "" : StringUtils.stringifyException(t)));
			System.exit(1);
		}
		if (t != null) LOG.error("Critical failure in UdaPlugin - trying vanilla as fallbackPlugin:\n" + StringUtils.stringifyException(t));
		try {
			fallbackPlugin = UdaMapredBridge.getShuffleConsumerPlugin(flavour.getVanillaPluginClass(), reduceTask, umbilical, jobConf, reporter);
			LOG.info("Succesfuly switched to Using fallbackPlugin");
		} catch (ClassNotFoundException e) {
			throw new UdaRuntimeException("Failed to initialize UDA Shuffle and failed to fallback to vanilla Shuffle", e);
		}
	}

	private synchronized boolean doFallbackFetchOutputs() throws IOException {
		if (fallbackFetchDone) return true;
		doFallbackInit(null);
		fallbackFetchDone = flavour.pluginFetchOutputs(fallbackPlugin);
		return fallbackFetchDone;
	}

	private boolean fetchOutputsInternal() throws IOException {
		EventsThread events = new EventsThread();
		events.start();
		LOG.info("fetchOutputs - Using UdaShuffleConsumerPlugin");
		synchronized (fetchLock) {
			while (!fetchCompleted && fallbackPlugin == null && events.failure == null) {
				try { fetchLock.wait(1000); } catch (InterruptedException e) { throw new IOException(e); }
			}
		}
		stopEvents = true;
		if (events.failure != null) throw new IOException("map completion events: " + events.failure, events.failure);
		if (fallbackPlugin != null) throw new UdaRuntimeException("native side failed during fetch");
		return true;
	}

	/** Polls the TaskTracker every second; SUCCEEDED maps become FETCH commands, anything that invalidates an
	 *  already fetched map (FAILED/KILLED/OBSOLETE, JobTracker reset) forces the fallback. */
	private final class EventsThread extends Thread {
		volatile Throwable failure;
		private final IntWritable fromEventId = new IntWritable(0);
		private final Set<TaskID> fetched = new HashSet<TaskID>();
		EventsThread() { setName("Thread for polling Map Completion Events"); setDaemon(true); }

		public void run() {
			try {
				while (!stopEvents) {
					MapTaskCompletionEventsUpdate update = flavour.pluginGetMapCompletionEvents(fromEventId, MAX_EVENTS_TO_FETCH);
					TaskCompletionEvent[] events = update.getMapTaskCompletionEvents();
					if (update.shouldReset()) {
						if (!fetched.isEmpty()) throw new IOException("JobTracker reset after maps were fetched");
						fromEventId.set(0);
					}
					fromEventId.set(fromEventId.get() + events.length);
					for (TaskCompletionEvent e : events) {
						TaskID task = e.getTaskAttemptId().getTaskID();
						switch (e.getTaskStatus()) {
						case SUCCEEDED:
							if (fetched.add(task)) {
								String host = java.net.URI.create(e.getTaskTrackerHttp()).getHost();
								rdmaChannel.sendFetchReq(host, e.getTaskAttemptId().getJobID().toString(), e.getTaskAttemptId().toString());
							}
							break;
						case FAILED: case KILLED: case OBSOLETE:
							if (fetched.contains(task)) throw new IOException("map " + task + " became " + e.getTaskStatus() + " after it was fetched");
							break;
						default:
							break;
						}
					}
					if (!stopEvents) Thread.sleep(1000);
				}
			} catch (InterruptedException ignored) {
			} catch (Throwable t) {
				failure = t;
				synchronized (fetchLock) { fetchLock.notifyAll(); }
			}
		}
	}
}
