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

/** mapreduce.job.reduce.shuffle.consumer.plugin.class for Hadoop 2.x / 3.x (init(Context), run(), close());
 *  the vanilla fallback is org.apache.hadoop.mapreduce.task.reduce.Shuffle. */
public class UdaShuffleConsumerPlugin<K, V> implements ShuffleConsumerPlugin<K, V>, UdaConsumerPluginCallable {
	private final UdaShuffleConsumerPluginShared<K, V> shared = new UdaShuffleConsumerPluginShared<K, V>(this);
	private ShuffleConsumerPlugin.Context<K, V> context;
	private ReduceTask reduceTask;
	private TaskUmbilicalProtocol umbilical;

	public void init(ShuffleConsumerPlugin.Context<K, V> context) {
		this.context = context;
		this.reduceTask = UdaMapredBridge.reduceTask(context);
		this.umbilical = context.getUmbilical();
		try {
			shared.init(reduceTask, umbilical, context.getJobConf(), context.getReporter(), context.getLocalFS());
		} catch (IOException e) {
			throw new UdaRuntimeException("init failed", e);
		}
	}

	public RawKeyValueIterator run() throws IOException, InterruptedException {
		shared.fetchOutputs();
		return shared.createKVIterator(context.getJobConf(), context.getLocalFS(), context.getReporter());
	}

	public void close() { shared.close(); }

	public boolean pluginFetchOutputs(ShuffleConsumerPlugin plugin) { return true; }   // vanilla Shuffle fetches inside run()
	public RawKeyValueIterator pluginCreateKVIterator(ShuffleConsumerPlugin plugin, JobConf job, FileSystem fs, Reporter reporter)
			throws IOException, InterruptedException {
		return plugin.run();
	}
	public Class getVanillaPluginClass() { return org.apache.hadoop.mapreduce.task.reduce.Shuffle.class; }
	public MapTaskCompletionEventsUpdate pluginGetMapCompletionEvents(IntWritable from, int max) throws IOException {
		return umbilical.getMapCompletionEvents(reduceTask.getJobID(), from.get(), max, reduceTask.getTaskID());
	}
}
