package org.apache.hadoop.mapred;

import java.io.IOException;

import org.apache.hadoop.fs.FileSystem;
import org.apache.hadoop.mapred.Task.TaskReporter;

/**
 * Hadoop 1.x plug-in point on the reduce side (the interface the reference adds to Hadoop with
 * plugins/HADOOP-1.x.y-v2.patch:226-331; upstream since Hadoop 2 as MAPREDUCE-4049). ReduceTask instantiates the class
 * named by {@code mapreduce.job.reduce.shuffle.consumer.plugin.class} (default: the built-in ReduceCopier wrapper),
 * calls {@link #init}, {@link #fetchOutputs}, asks for the merge error if that failed, then {@link #createKVIterator}
 * and finally {@link #close}.
 *
 * Only needed when building against a Hadoop 1.x tree that does not carry the patch; see ../../../../README.md.
 */
public interface ShuffleConsumerPlugin {
	/** Called once, right after construction through the configured class name. */
	void init(Context context) throws ClassNotFoundException, IOException;

	/** Brings every map output of this reducer's partition in. False = failed; the cause is getMergeThrowable(). */
	boolean fetchOutputs() throws IOException;

	/** The error of the merge side, if any. */
	Throwable getMergeThrowable();

	/** The merged, sorted stream of raw key/value pairs the reducer iterates over. */
	RawKeyValueIterator createKVIterator(JobConf job, FileSystem fs, Reporter reporter) throws IOException;

	/** Releases everything the plug-in holds. */
	void close();

	/** What ReduceTask hands to a plug-in. */
	class Context {
		private final ReduceTask reduceTask;
		private final TaskUmbilicalProtocol umbilical;
		private final JobConf conf;
		private final TaskReporter reporter;

		public Context(ReduceTask reduceTask, TaskUmbilicalProtocol umbilical, JobConf conf, TaskReporter reporter) {
			this.reduceTask = reduceTask;
			this.umbilical = umbilical;
			this.conf = conf;
			this.reporter = reporter;
		}

		public ReduceTask getReduceTask() { return reduceTask; }
		public JobConf getConf() { return conf; }
		public TaskUmbilicalProtocol getUmbilical() { return umbilical; }
		public TaskReporter getReporter() { return reporter; }
	}
}
