package com.mellanox.hadoop.mapred;

import java.io.IOException;

import org.apache.hadoop.conf.Configuration;
import org.apache.hadoop.fs.LocalDirAllocator;
import org.apache.hadoop.fs.Path;
import org.apache.hadoop.mapred.IndexCacheBridge;
import org.apache.hadoop.mapreduce.JobID;
import org.apache.hadoop.yarn.api.records.ApplicationId;
import org.apache.hadoop.yarn.conf.YarnConfiguration;

/** YARN provider: map outputs live under usercache/<user>/appcache/<appId>/output/<mapId>/
 *  (reference: mlx-2.x/.../UdaPluginSH.java:67-146). */
class UdaPluginSH extends UdaShuffleProviderPluginShared {
	private final UdaShuffleHandler handler;
	private final LocalDirAllocator dirs = new LocalDirAllocator(YarnConfiguration.NM_LOCAL_DIRS);
	private final IndexCacheBridge indexCache;
	private final Configuration conf;

	UdaPluginSH(Configuration conf, UdaShuffleHandler handler) {
		super(UdaShuffleHandler.asJobConf(conf));
		this.conf = conf;
		this.handler = handler;
		this.indexCache = new IndexCacheBridge(jobConf);
		prepareLog("org.apache.hadoop.mapred.ShuffleProviderPlugin");
		launchCppSide(false, null);
	}

	Object pathIndex(String jobId, String mapId, int reduce) {
		try {
			String user = handler.getUser(jobId);
			JobID job = JobID.forName(jobId);
			ApplicationId app = ApplicationId.newInstance(Long.parseLong(job.getJtIdentifier()), job.getId());
			String base = "usercache/" + user + "/appcache/" + app.toString() + "/output/" + mapId;
			Path index = dirs.getLocalPathToRead(base + "/file.out.index", conf);
			Path out = dirs.getLocalPathToRead(base + "/file.out", conf);
			return indexCache.getIndexInformation(mapId, reduce, index, user, out.toUri().getRawPath());
		} catch (IOException e) {
			LOG.error("getPathIndex failed for " + jobId + " " + mapId + " " + reduce + ": " + e);
			return null;
		}
	}
}
