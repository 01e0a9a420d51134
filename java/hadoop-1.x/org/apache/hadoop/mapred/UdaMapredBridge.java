package org.apache.hadoop.mapred;

import java.io.IOException;
import java.lang.reflect.Constructor;
import java.lang.reflect.Method;

import org.apache.hadoop.fs.FileSystem;
import org.apache.hadoop.util.ReflectionUtils;

/** Reaches into Hadoop's mapred package for the vanilla shuffle (ReduceTask.ReduceCopier) and papers over
 *  the vanilla / CDH / HDP differences by reflection: getConf vs getJobConf on the Context,
 *  createKVIterator arity, Context constructor argument order, ReduceCopier as inner vs static class
 *  (reference: mlx-1.x/org/apache/hadoop/mapred/UdaMapredBridge.java:34-85). */
public class UdaMapredBridge {
	public static Class vanillaConsumerClass() {
		try {
			return Class.forName("org.apache.hadoop.mapred.ReduceTask$ReduceCopier");
		} catch (ClassNotFoundException e) {
			throw new RuntimeException(e);
		}
	}

	public static JobConf getJobConf(ShuffleConsumerPlugin.Context ctx) {
		for (String name : new String[] {"getConf", "getJobConf"}) {
			try {
				Method m = ctx.getClass().getMethod(name);
				return (JobConf) m.invoke(ctx);
			} catch (Exception ignored) {
			}
		}
		throw new RuntimeException("ShuffleConsumerPlugin.Context exposes neither getConf nor getJobConf");
	}

	public static ShuffleConsumerPlugin getShuffleConsumerPlugin(Class clazz, ReduceTask task, TaskUmbilicalProtocol umbilical,
			JobConf conf, Reporter reporter) throws ClassNotFoundException, IOException {
		ShuffleConsumerPlugin plugin;
		try {
			plugin = (ShuffleConsumerPlugin) ReflectionUtils.newInstance(clazz, conf);
		} catch (RuntimeException e) {
			// ReduceCopier as a non-static inner class needs the enclosing ReduceTask
			try {
				Constructor c = clazz.getDeclaredConstructor(ReduceTask.class);
				c.setAccessible(true);
				plugin = (ShuffleConsumerPlugin) c.newInstance(task);
			} catch (Exception e2) {
				throw new ClassNotFoundException("cannot instantiate " + clazz, e2);
			}
		}
		ShuffleConsumerPlugin.Context ctx = null;
		for (Constructor c : ShuffleConsumerPlugin.Context.class.getConstructors()) {
			try {
				Class[] p = c.getParameterTypes();
				Object[] args = new Object[p.length];
				for (int i = 0; i < p.length; i++) {
					if (p[i].isInstance(task)) args[i] = task;
					else if (p[i].isInstance(umbilical)) args[i] = umbilical;
					else if (p[i].isInstance(conf)) args[i] = conf;
					else if (p[i].isInstance(reporter)) args[i] = reporter;
				}
				ctx = (ShuffleConsumerPlugin.Context) c.newInstance(args);
				break;
			} catch (Exception ignored) {
			}
		}
		if (ctx == null) throw new IOException("no usable ShuffleConsumerPlugin.Context constructor");
		plugin.init(ctx);
		return plugin;
	}

	public static RawKeyValueIterator createKVIterator(ShuffleConsumerPlugin plugin, JobConf job, FileSystem fs, Reporter reporter)
			throws IOException {
		try {
			for (Method m : plugin.getClass().getMethods()) {
				if (!m.getName().equals("createKVIterator")) continue;
				if (m.getParameterTypes().length == 3) return (RawKeyValueIterator) m.invoke(plugin, job, fs, reporter);
				if (m.getParameterTypes().length == 0) return (RawKeyValueIterator) m.invoke(plugin);
			}
		} catch (Exception e) {
			throw new IOException(e);
		}
		throw new IOException("vanilla plugin has no createKVIterator");
	}

	public static JvmContext jvmContext(ReduceTask task) { return task.jvmContext; }
}
