package com.mellanox.hadoop.mapred;

import java.io.IOException;

import org.apache.hadoop.fs.FileSystem;
import org.apache.hadoop.io.IntWritable;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapred.MapTaskCompletionEventsUpdate;
import org.apache.hadoop.mapred.RawKeyValueIterator;
import org.apache.hadoop.mapred.ReduceTask;
import org.apache.hadoop.mapred.Reporter;
import org.apache.hadoop.mapred.ShuffleConsumerPlugin;
import org.apache.hadoop.mapred.TaskUmbilicalProtocol;
import org.apache.hadoop.mapred.UdaMapredBridge;

/** mapreduce.job.reduce.shuffle.consumer.plugin.class=com.mellanox.hadoop.mapred.UdaShuffleConsumerPlugin
 *  (Hadoop 1.x / CDH3 / HDP1 flavour; signature differences between those are hidden in UdaMapredBridge). */
public class UdaShuffleConsumerPlugin<K, V> implements ShuffleConsumerPlugin, UdaConsumerPluginCallable {
	private final UdaShuffleConsumerPluginShared<K, V> shared = new UdaShuffleConsumerPluginShared<K, V>(this);
	private ReduceTask reduceTask;
	private TaskUmbilicalProtocol umbilical;

	public void init(ShuffleConsumerPlugin.Context context) throws IOException {
		this.reduceTask = context.getReduceTask();
		this.umbilical = context.getUmbilical();
		shared.init(reduceTask, umbilical, UdaMapredBridge.getJobConf(context), context.getReporter(), null);
	}

	public boolean fetchOutputs() throws IOException { return shared.fetchOutputs(); }
	public Throwable getMergeThrowable() { return shared.getMergeThrowable(); }
	public RawKeyValueIterator createKVIterator(JobConf job, FileSystem fs, Reporter reporter) throws IOException {
		try {
			return shared.createKVIterator(job, fs, reporter);
		} catch (InterruptedException e) {
			throw new IOException(e);
		}
	}
	public void close() { shared.close(); }

	// ---- UdaConsumerPluginCallable ---------------------------------------------------------------
	public boolean pluginFetchOutputs(ShuffleConsumerPlugin plugin) throws IOException { return plugin.fetchOutputs(); }
	public RawKeyValueIterator pluginCreateKVIterator(ShuffleConsumerPlugin plugin, JobConf job, FileSystem fs, Reporter reporter)
			throws IOException {
		return UdaMapredBridge.createKVIterator(plugin, job, fs, reporter);
	}
	public Class getVanillaPluginClass() { return UdaMapredBridge.vanillaConsumerClass(); }
	public MapTaskCompletionEventsUpdate pluginGetMapCompletionEvents(IntWritable from, int max) throws IOException {
		return umbilical.getMapCompletionEvents(reduceTask.getJobID(), from.get(), max, reduceTask.getTaskID(), UdaMapredBridge.jvmContext(reduceTask));
	}
}
