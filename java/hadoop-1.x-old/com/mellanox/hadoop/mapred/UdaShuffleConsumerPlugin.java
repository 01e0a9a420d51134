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

/**
 * Consumer plug-in for Hadoop 1.x trees patched with the FIRST generation of the plug-in point
 * (reference: plugins/HADOOP-1.x.y-old.patch, plugins/mlx-1.x-old): there ShuffleConsumerPlugin is an abstract class whose
 * init() takes the four objects directly instead of a Context, there is no getMergeThrowable(), and no vanilla plug-in
 * class to fall back to (getVanillaPluginClass() returns null: a failure fails the task, as in that generation).
 */
public class UdaShuffleConsumerPlugin<K, V> extends ShuffleConsumerPlugin implements UdaConsumerPluginCallable {
	private final UdaShuffleConsumerPluginShared<K, V> shared = new UdaShuffleConsumerPluginShared<K, V>(this);
	private ReduceTask reduceTask;
	private TaskUmbilicalProtocol umbilical;

	@Override
	public void init(ReduceTask reduceTask, TaskUmbilicalProtocol umbilical, JobConf conf, Reporter reporter) throws IOException {
		this.reduceTask = reduceTask;
		this.umbilical = umbilical;
		shared.init(reduceTask, umbilical, conf, reporter, null);
	}

	@Override
	public boolean fetchOutputs() throws IOException { return shared.fetchOutputs(); }

	@Override
	public RawKeyValueIterator createKVIterator(JobConf job, FileSystem fs, Reporter reporter) throws IOException {
		try {
			return shared.createKVIterator(job, fs, reporter);
		} catch (InterruptedException e) {
			throw new IOException(e);
		}
	}

	@Override
	public void close() { shared.close(); }

	// ---- UdaConsumerPluginCallable ---------------------------------------------------------------
	public boolean pluginFetchOutputs(ShuffleConsumerPlugin plugin) throws IOException { return plugin.fetchOutputs(); }
	public RawKeyValueIterator pluginCreateKVIterator(ShuffleConsumerPlugin plugin, JobConf job, FileSystem fs, Reporter reporter)
			throws IOException, InterruptedException {
		return plugin.createKVIterator(job, fs, reporter);
	}
	public Class getVanillaPluginClass() { return null; }
	public MapTaskCompletionEventsUpdate pluginGetMapCompletionEvents(IntWritable from, int max) throws IOException {
		return umbilical.getMapCompletionEvents(reduceTask.getJobID(), from.get(), max, reduceTask.getTaskID(), reduceTask.getJvmContext());
	}
}
