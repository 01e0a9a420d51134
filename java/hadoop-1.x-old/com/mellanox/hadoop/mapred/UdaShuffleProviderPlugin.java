package com.mellanox.hadoop.mapred;

import org.apache.hadoop.mapred.ShuffleProviderPlugin;
import org.apache.hadoop.mapred.TaskTracker;

/** Provider plug-in for the first-generation Hadoop 1.x plug-in point (reference: plugins/mlx-1.x-old): the provider
 *  interface is the same in both generations, so this is the hadoop-1.x class built against the older tree. */
public class UdaShuffleProviderPlugin implements ShuffleProviderPlugin {
	private UdaPluginTT provider;

	public void initialize(TaskTracker taskTracker) {
		provider = new UdaPluginTT(taskTracker, taskTracker.getJobConf());
	}

	public void destroy() {
		if (provider != null) provider.close();
	}
}
