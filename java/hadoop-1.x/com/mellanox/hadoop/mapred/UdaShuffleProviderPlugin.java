package com.mellanox.hadoop.mapred;

import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapred.ShuffleProviderPlugin;
import org.apache.hadoop.mapred.TaskTracker;

/** mapreduce.shuffle.provider.plugin.classes=com.mellanox.hadoop.mapred.UdaShuffleProviderPlugin */
public class UdaShuffleProviderPlugin implements ShuffleProviderPlugin {
	private UdaPluginTT provider;

	public void initialize(TaskTracker taskTracker) {
		provider = new UdaPluginTT(taskTracker, taskTracker.getJobConf());
	}

	public void destroy() {
		if (provider != null) provider.close();
	}
}
