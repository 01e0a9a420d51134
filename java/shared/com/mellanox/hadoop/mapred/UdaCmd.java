package com.mellanox.hadoop.mapred;

import java.util.List;

/** Command strings understood by the native parser (csrc/runtime/protocol.cc):
 *  "&lt;n+1&gt;:&lt;cmd&gt;:&lt;p1&gt;:...:&lt;pn&gt;"; the last parameter may itself contain ':'. */
final class UdaCmd {
	static final int EXIT_COMMAND = 0, NEW_MAP_COMMAND = 1, FINAL_MERGE_COMMAND = 2, RESULT_COMMAND = 3, FETCH_COMMAND = 4,
			FETCH_OVER_COMMAND = 5, JOB_OVER_COMMAND = 6, INIT_COMMAND = 7, MORE_COMMAND = 8, NETLEV_REDUCE_LAUNCHED = 9;

	private UdaCmd() {}

	static String formCmd(int cmd, List<String> params) {
		StringBuilder sb = new StringBuilder().append(params.size() + 1).append(':').append(cmd);
		for (String p : params) sb.append(':').append(p);
		return sb.toString();
	}
}
