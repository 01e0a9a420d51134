package com.mellanox.hadoop.mapred;

import java.nio.ByteBuffer;
import java.util.Map;
import java.util.concurrent.ConcurrentHashMap;

import org.apache.hadoop.conf.Configuration;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapreduce.JobID;
import org.apache.hadoop.yarn.api.records.ApplicationId;
import org.apache.hadoop.yarn.server.nodemanager.containermanager.AuxServices;
import org.apache.hadoop.yarn.service.AbstractService;

/**
 * The same "uda_shuffle" auxiliary service for the early YARN API (Hadoop 2.0.x alpha), where an auxiliary service is
 * an AbstractService that implements AuxServices.AuxiliaryService with initApp / stopApp instead of extending
 * org.apache.hadoop.yarn.server.api.AuxiliaryService (reference: mlx-2.0.x/.../UdaShuffleHandler.java:58-141).
 * Everything else (UdaPluginSH, the consumer plug-in, UdaMapredBridge) is shared with the hadoop-2.x flavour; the
 * Makefile compiles this file instead of hadoop-2.x's handler.
 */
public class UdaShuffleHandler extends AbstractService implements AuxServices.AuxiliaryService {
	public static final String SERVICE_ID = "uda_shuffle";
	private final Map<String, String> userOfJob = new ConcurrentHashMap<String, String>();
	private UdaPluginSH provider;
	private Configuration config;

	public UdaShuffleHandler() { super("uda_shuffle"); }

	@Override public void initApp(String user, ApplicationId app, ByteBuffer secret) {
		userOfJob.put(new JobID(Long.toString(app.getClusterTimestamp()), app.getId()).toString(), user);
	}

	@Override public void stopApp(ApplicationId app) {
		userOfJob.remove(new JobID(Long.toString(app.getClusterTimestamp()), app.getId()).toString());
	}

	@Override public synchronized void init(Configuration conf) {
		this.config = conf;
		super.init(conf);
	}

	@Override public synchronized void start() {
		provider = new UdaPluginSH(config, this);
		super.start();
	}

	@Override public synchronized void stop() {
		if (provider != null) provider.close();
		super.stop();
	}

	/** The consumer does not need service meta data. */
	@Override public ByteBuffer getMeta() { return ByteBuffer.allocate(0); }

	String getUser(String jobId) { return userOfJob.get(jobId); }

	static JobConf asJobConf(Configuration c) { return c instanceof JobConf ? (JobConf) c : new JobConf(c); }
}
