package org.apache.hadoop.mapred;

/**
 * Hadoop 1.x plug-in point on the TaskTracker side (reference: plugins/HADOOP-1.x.y-v2.patch:332-383). The TaskTracker
 * instantiates every class listed in {@code mapreduce.shuffle.provider.plugin.classes}, calls {@link #initialize} when it
 * starts serving and {@link #destroy} when it shuts down; a provider serves map outputs to reducers by its own means
 * next to (not instead of) the built-in HTTP servlet.
 */
public interface ShuffleProviderPlugin {
	void initialize(TaskTracker taskTracker);

	void destroy();
}
