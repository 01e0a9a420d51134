package org.apache.hadoop.mapred;

import java.io.IOException;

import org.apache.hadoop.util.ReflectionUtils;

/** Package-private access for the 2.x / 3.x flavour (reference: mlx-2.x/org/apache/hadoop/mapred/UdaMapredBridge.java:30-46). */
public class UdaMapredBridge {
	public static ReduceTask reduceTask(ShuffleConsumerPlugin.Context ctx) { return (ReduceTask) ctx.getReduceTask(); }

	@SuppressWarnings("unchecked")
	public static ShuffleConsumerPlugin getShuffleConsumerPlugin(Class clazz, ReduceTask task, TaskUmbilicalProtocol umbilical,
			JobConf conf, Reporter reporter) throws ClassNotFoundException, IOException {
		ShuffleConsumerPlugin plugin = (ShuffleConsumerPlugin) ReflectionUtils.newInstance(clazz, conf);
		plugin.init(task.shuffleContext);   // the Context the ReduceTask built for the configured plugin
		return plugin;
	}
}
