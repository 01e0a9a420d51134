package com.mellanox.hadoop.mapred;

import java.io.IOException;

import org.apache.commons.logging.LogFactory;
import org.apache.hadoop.fs.LocalDirAllocator;
import org.apache.hadoop.fs.Path;
import org.apache.hadoop.mapred.IndexCacheBridge;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapred.JobID;
import org.apache.hadoop.mapred.TaskTracker;

/** Hadoop 1 provider: resolves file.out / file.out.index through the TaskTracker's local-dir allocator and
 *  Hadoop's IndexCache (reference: UdaPluginTT1.java:58-143). */
class UdaPluginTT extends UdaShuffleProviderPluginShared {
	private final TaskTracker taskTracker;
	private final LocalDirAllocator dirs = new LocalDirAllocator("mapred.local.dir");
	private final IndexCacheBridge indexCache;
	private final java.util.Map<String, Path> fileCache = java.util.Collections.synchronizedMap(new LruMap<String, Path>(1000));
	private final java.util.Map<String, Path> indexFileCache = java.util.Collections.synchronizedMap(new LruMap<String, Path>(1000));

	private static final class LruMap<A, B> extends java.util.LinkedHashMap<A, B> {
		private final int cap;
		LruMap(int cap) { super(16, 0.75f, true); this.cap = cap; }
		protected boolean removeEldestEntry(java.util.Map.Entry<A, B> e) { return size() > cap; }
	}

	UdaPluginTT(TaskTracker tt, JobConf conf) {
		super(conf);
		this.taskTracker = tt;
		this.indexCache = new IndexCacheBridge(conf);
		LOG = LogFactory.getLog("org.apache.hadoop.mapred.ShuffleProviderPlugin");
		prepareLog("org.apache.hadoop.mapred.ShuffleProviderPlugin");
		launchCppSide(false, null);
	}

	Object pathIndex(String jobId, String mapId, int reduce) {
		try {
			JobConf job = taskTracker.getJobConf(JobID.forName(jobId));
			String user = job.getUser();
			String base = TaskTracker.getIntermediateOutputDir(user, jobId, mapId);
			String key = jobId + "/" + mapId;
			Path index = indexFileCache.get(key);
			if (index == null) { index = dirs.getLocalPathToRead(base + "/file.out.index", jobConf); indexFileCache.put(key, index); }
			Path out = fileCache.get(key);
			if (out == null) { out = dirs.getLocalPathToRead(base + "/file.out", jobConf); fileCache.put(key, out); }
			return indexCache.getIndexInformation(mapId, reduce, index, user, out.toUri().getRawPath());
		} catch (IOException e) {
			LOG.error("getPathIndex failed for " + jobId + " " + mapId + " " + reduce + ": " + e);
			return null;
		}
	}

	void jobOver(String jobId) { /* cached paths age out of the LRU maps */ }
}
