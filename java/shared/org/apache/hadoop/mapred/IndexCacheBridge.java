package org.apache.hadoop.mapred;

import java.io.IOException;

import org.apache.hadoop.fs.Path;

/** Access to Hadoop's package-private IndexCache / IndexRecord from the com.mellanox package. */
public class IndexCacheBridge {
	private final IndexCache cache;

	public IndexCacheBridge(JobConf conf) { cache = new IndexCache(conf); }

	public IndexRecordBridge getIndexInformation(String mapId, int reduce, Path indexFile, String expectedOwner, String mofPath)
			throws IOException {
		IndexRecord r = cache.getIndexInformation(mapId, reduce, indexFile, expectedOwner);
		return new IndexRecordBridge(r.startOffset, r.rawLength, r.partLength, mofPath);
	}

	public void removeMap(String mapId) { cache.removeMap(mapId); }
}
