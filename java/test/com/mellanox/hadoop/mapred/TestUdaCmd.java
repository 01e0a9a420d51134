package com.mellanox.hadoop.mapred;

import static org.junit.Assert.assertEquals;

import java.util.ArrayList;
import java.util.Arrays;

import org.junit.Test;

/** The command strings must be exactly what the native parser expects (csrc/runtime/protocol.cc, goldens in
 *  tests/test_runtime_cpu.py::test_command_golden): "count:header:p1:...:pn", count = header + parameters. */
public class TestUdaCmd {
	@Test
	public void exitHasNoParameters() {
		assertEquals("1:0", UdaCmd.formCmd(UdaCmd.EXIT_COMMAND, new ArrayList<String>()));
	}

	@Test
	public void fetchCarriesHostJobMapReduce() {
		assertEquals("5:4:host7:job_1:attempt_m_3:2",
				UdaCmd.formCmd(UdaCmd.FETCH_COMMAND, Arrays.asList("host7", "job_1", "attempt_m_3", "2")));
	}

	@Test
	public void lastParameterMayContainColons() {
		assertEquals("3:7:a:/x:y", UdaCmd.formCmd(UdaCmd.INIT_COMMAND, Arrays.asList("a", "/x:y")));
	}

	@Test
	public void headerValuesAreTheNativeEnum() {
		int[] want = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9};
		int[] got = {UdaCmd.EXIT_COMMAND, UdaCmd.NEW_MAP_COMMAND, UdaCmd.FINAL_MERGE_COMMAND, UdaCmd.RESULT_COMMAND, UdaCmd.FETCH_COMMAND,
				UdaCmd.FETCH_OVER_COMMAND, UdaCmd.JOB_OVER_COMMAND, UdaCmd.INIT_COMMAND, UdaCmd.MORE_COMMAND, UdaCmd.NETLEV_REDUCE_LAUNCHED};
		for (int i = 0; i < want.length; ++i) assertEquals(want[i], got[i]);
	}
}
