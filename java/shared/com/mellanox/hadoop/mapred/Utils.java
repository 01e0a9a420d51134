package com.mellanox.hadoop.mapred;

import java.lang.reflect.Constructor;
import java.lang.reflect.InvocationTargetException;
import java.lang.reflect.Method;

/**
 * Reflection helpers for the places where Hadoop distributions differ in signatures that the plug-ins must call
 * (vanilla vs CDH vs HDP: TaskTracker.getJobConf / getConf, ReduceTask.createKVIterator argument lists, ...).
 * Counterpart of the reference's plugins/shared/.../Utils.java:7-100. Failures surface as UdaRuntimeException so that
 * the caller's fallback logic sees one exception type.
 */
public final class Utils {
	private Utils() {}

	/** Finds a declared or inherited method by name and exact parameter types; null when the class has none. */
	public static Method findMethod(Class<?> owner, String name, Class<?>... params) {
		for (Class<?> c = owner; c != null; c = c.getSuperclass()) {
			try {
				Method m = c.getDeclaredMethod(name, params);
				m.setAccessible(true);
				return m;
			} catch (NoSuchMethodException e) {
				// keep walking up
			}
		}
		return null;
	}

	/** Calls owner.name(args) on target (null target for a static method). */
	public static Object invoke(Class<?> owner, String name, Class<?>[] params, Object target, Object... args) {
		Method m = findMethod(owner, name, params);
		if (m == null) throw new UdaRuntimeException("no method " + owner.getName() + "." + name);
		try {
			return m.invoke(target, args);
		} catch (InvocationTargetException e) {
			throw new UdaRuntimeException("exception in " + owner.getName() + "." + name, e.getCause());
		} catch (IllegalAccessException e) {
			throw new UdaRuntimeException("cannot access " + owner.getName() + "." + name, e);
		}
	}

	/** Calls the first of several alternative method names that exists (e.g. "getJobConf" then "getConf"). */
	public static Object invokeFirst(Class<?> owner, String[] names, Class<?>[] params, Object target, Object... args) {
		for (String n : names)
			if (findMethod(owner, n, params) != null) return invoke(owner, n, params, target, args);
		throw new UdaRuntimeException("none of the expected methods exists in " + owner.getName());
	}

	/** new owner(args) through a possibly non-public constructor. */
	public static <T> T construct(Class<T> owner, Class<?>[] params, Object... args) {
		try {
			Constructor<T> k = owner.getDeclaredConstructor(params);
			k.setAccessible(true);
			return k.newInstance(args);
		} catch (InvocationTargetException e) {
			throw new UdaRuntimeException("exception in the constructor of " + owner.getName(), e.getCause());
		} catch (ReflectiveOperationException e) {
			throw new UdaRuntimeException("cannot construct " + owner.getName(), e);
		}
	}

	/** new owner(arg) with a single argument of the given declared type. */
	public static <T> T constructWithArg(Class<T> owner, Class<?> argType, Object arg) {
		return construct(owner, new Class<?>[] {argType}, arg);
	}
}
