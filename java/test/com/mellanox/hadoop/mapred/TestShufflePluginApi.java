package com.mellanox.hadoop.mapred;

import static org.junit.Assert.assertEquals;
import static org.junit.Assert.assertFalse;
import static org.junit.Assert.assertNotNull;
import static org.junit.Assert.assertTrue;

import java.io.IOException;
import java.lang.reflect.Method;
import java.lang.reflect.Modifier;

import org.apache.hadoop.fs.FileSystem;
import org.apache.hadoop.mapred.JobConf;
import org.apache.hadoop.mapred.RawKeyValueIterator;
import org.apache.hadoop.mapred.Reporter;
import org.apache.hadoop.mapred.ShuffleConsumerPlugin;
import org.apache.hadoop.mapred.ShuffleProviderPlugin;
import org.apache.hadoop.mapred.TaskTracker;
import org.apache.hadoop.util.ReflectionUtils;
import org.junit.Test;

/**
 * Keeps the API that third-party shuffle plug-ins compile against (purpose of the reference's TestShufflePlugin,
 * plugins/HADOOP-1.x.y-v2.patch:592-785, and TestShuffleProvider, plugins/MAPREDUCE-5329.patch:64-227): the plug-in
 * points can be implemented from outside org.apache.hadoop.mapred, are loadable by class name through the two
 * configuration keys, and the UDA plug-ins implement them with the JNI names the native bridge looks up.
 */
public class TestShufflePluginApi {
	/** A consumer that only checks it can reach what a real plug-in needs from the Context. */
	public static class ProbeConsumer implements ShuffleConsumerPlugin {
		boolean inited, closed;
		public void init(ShuffleConsumerPlugin.Context context) {
			context.getReduceTask();
			context.getConf();
			context.getUmbilical();
			context.getReporter();
			inited = true;
		}
		public boolean fetchOutputs() throws IOException { return true; }
		public Throwable getMergeThrowable() { return null; }
		public RawKeyValueIterator createKVIterator(JobConf job, FileSystem fs, Reporter reporter) throws IOException { return null; }
		public void close() { closed = true; }
	}

	public static class ProbeProvider implements ShuffleProviderPlugin {
		static int initialized, destroyed;
		public void initialize(TaskTracker tt) { ++initialized; }
		public void destroy() { ++destroyed; }
	}

	@Test
	public void consumerIsLoadableThroughItsConfigurationKey() throws Exception {
		JobConf job = new JobConf();
		job.setClass("mapreduce.job.reduce.shuffle.consumer.plugin.class", ProbeConsumer.class, ShuffleConsumerPlugin.class);
		Class<? extends ShuffleConsumerPlugin> c = job.getClass("mapreduce.job.reduce.shuffle.consumer.plugin.class", null, ShuffleConsumerPlugin.class);
		ShuffleConsumerPlugin p = ReflectionUtils.newInstance(c, job);
		p.init(new ShuffleConsumerPlugin.Context(null, null, job, null));
		assertTrue(p.fetchOutputs());
		p.close();
		assertTrue(((ProbeConsumer) p).inited && ((ProbeConsumer) p).closed);
	}

	@Test
	public void providersAreLoadableAsAList() throws Exception {
		JobConf job = new JobConf();
		job.set("mapreduce.shuffle.provider.plugin.classes", ProbeProvider.class.getName() + "," + ProbeProvider.class.getName());
		Class<?>[] classes = job.getClasses("mapreduce.shuffle.provider.plugin.classes");
		assertEquals(2, classes.length);
		for (Class<?> c : classes) {
			ShuffleProviderPlugin p = (ShuffleProviderPlugin) ReflectionUtils.newInstance(c, job);
			p.initialize(null);
			p.destroy();
		}
		assertEquals(2, ProbeProvider.initialized);
		assertEquals(2, ProbeProvider.destroyed);
	}

	@Test
	public void udaPluginsImplementThePluginPoints() {
		assertTrue(ShuffleConsumerPlugin.class.isAssignableFrom(UdaShuffleConsumerPlugin.class));
		assertTrue(ShuffleProviderPlugin.class.isAssignableFrom(UdaShuffleProviderPlugin.class));
		assertFalse(Modifier.isAbstract(UdaShuffleConsumerPlugin.class.getModifiers()));
	}

	/** The native bridge finds these by name and signature (csrc/jni/uda_bridge.cc:163-167,136). */
	@Test
	public void upCallTargetsHaveTheSignaturesTheNativeSideLooksUp() throws Exception {
		Class<?> b = UdaBridge.class;
		assertStatic(b.getDeclaredMethod("fetchOverMessage"));
		assertStatic(b.getDeclaredMethod("dataFromUda", Object.class, int.class));
		Method m = b.getDeclaredMethod("getPathUda", String.class, String.class, int.class);
		assertStatic(m);
		assertEquals(Object.class, m.getReturnType());
		m = b.getDeclaredMethod("getConfData", String.class, String.class);
		assertStatic(m);
		assertEquals(String.class, m.getReturnType());
		assertStatic(b.getDeclaredMethod("logToJava", String.class, int.class));
		assertStatic(b.getDeclaredMethod("failureInUda"));
		for (String n : new String[] {"startNative", "doCommandNative", "reduceExitMsgNative", "setLogLevelNative"}) {
			Method nat = null;
			for (Method x : b.getDeclaredMethods()) if (x.getName().equals(n)) nat = x;
			assertNotNull(n, nat);
			assertTrue(n, Modifier.isNative(nat.getModifiers()) && Modifier.isStatic(nat.getModifiers()));
		}
	}

	private static void assertStatic(Method m) { assertTrue(m.getName(), Modifier.isStatic(m.getModifiers())); }
}
