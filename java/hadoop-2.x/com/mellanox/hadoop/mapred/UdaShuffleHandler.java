package com.mellanox.hadoop.mapred;

import java.nio.ByteBuffer;
import java.util.Map;
import java.util.concurrent.ConcurrentHashMap;

import org.apache.hadoop.conf.Configuration;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapreduce.JobID;
import org.apache.hadoop.yarn.api.records.ApplicationId;
import org.apache.hadoop.yarn.server.api.ApplicationInitializationContext;
import org.apache.hadoop.yarn.server.api.ApplicationTerminationContext;
import org.apache.hadoop.yarn.server.api.AuxiliaryService;

/**
 * NodeManager auxiliary service "uda_shuffle":
 *   yarn.nodemanager.aux-services = mapreduce_shuffle,uda_shuffle
 *   yarn.nodemanager.aux-services.uda_shuffle.class = com.mellanox.hadoop.mapred.UdaShuffleHandler
 *   mapreduce.job.shuffle.provider.services = uda_shuffle
 * Keeps the jobId -> user map that the path lookup needs (reference: mlx-2.x/.../UdaShuffleHandler.java:59-151).
 */
public class UdaShuffleHandler extends AuxiliaryService {
	public static final String SERVICE_ID = "uda_shuffle";
	private final Map<String, String> userOfJob = new ConcurrentHashMap<String, String>();
	private UdaPluginSH provider;
	private Configuration config;

	public UdaShuffleHandler() { super("uda_shuffle"); }

	@Override public void initializeApplication(ApplicationInitializationContext ctx) {
		ApplicationId app = ctx.getApplicationId();
		userOfJob.put(new JobID(Long.toString(app.getClusterTimestamp()), app.getId()).toString(), ctx.getUser());
	}

	@Override public void stopApplication(ApplicationTerminationContext ctx) {
		ApplicationId app = ctx.getApplicationId();
		userOfJob.remove(new JobID(Long.toString(app.getClusterTimestamp()), app.getId()).toString());
	}

	@Override protected void serviceInit(Configuration conf) throws Exception {
		this.config = conf;
		super.serviceInit(conf);
	}

	@Override protected void serviceStart() throws Exception {
		provider = new UdaPluginSH(config, this);
		super.serviceStart();
	}

	@Override protected void serviceStop() throws Exception {
		if (provider != null) provider.close();
		super.serviceStop();
	}

	/** The consumer does not need service meta data; an empty buffer avoids YARN-1256. */
	@Override public ByteBuffer getMetaData() { return ByteBuffer.allocate(0); }

	String getUser(String jobId) { return userOfJob.get(jobId); }

	static JobConf asJobConf(Configuration c) { return c instanceof JobConf ? (JobConf) c : new JobConf(c); }
}
