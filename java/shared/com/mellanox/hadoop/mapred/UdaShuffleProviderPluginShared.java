package com.mellanox.hadoop.mapred;

import java.util.ArrayList;

import org.apache.hadoop.mapred.JobConf;

/** Provider side shared by the TaskTracker (Hadoop 1) and NodeManager (YARN) flavours
 *  (reference: UdaShuffleProviderPluginShared.java:29-52). */
abstract class UdaShuffleProviderPluginShared extends UdaPlugin {
	private static UdaShuffleProviderPluginShared instance;

	UdaShuffleProviderPluginShared(JobConf conf) {
		super(conf);
		instance = this;
	}

	protected void buildCmdParams() {
		cmdParams.clear();
		cmdParams.add("-w"); cmdParams.add(jobConf.get("mapred.rdma.wqe.per.conn", "256"));
		cmdParams.add("-r"); cmdParams.add(jobConf.get("mapred.rdma.cma.port", "9011"));
		cmdParams.add("-m"); cmdParams.add("1");
		cmdParams.add("-g"); cmdParams.add(System.getProperty("hadoop.log.dir"));
		cmdParams.add("-s"); cmdParams.add(jobConf.get("mapred.rdma.buf.size", "1024"));
	}

	/** Resolve a map output to (file.out path, index record); implemented per Hadoop flavour. */
	abstract Object pathIndex(String jobId, String mapId, int reduce);

	static Object getPathIndex(String jobId, String mapId, int reduce) {
		return instance == null ? null : instance.pathIndex(jobId, mapId, reduce);
	}

	void close() {
		UdaBridge.doCommand(UdaCmd.formCmd(UdaCmd.EXIT_COMMAND, new ArrayList<String>()));
	}
}
