package org.apache.hadoop.mapred;

/** Plain holder read field-by-field from C++ (GetFieldID on startOffset/rawLength/partLength: J and
 *  pathMOF: Ljava/lang/String;). Lives in Hadoop's package because IndexRecord is package-private there. */
public class IndexRecordBridge {
	public long startOffset;
	public long rawLength;
	public long partLength;
	public String pathMOF;

	public IndexRecordBridge() {}

	public IndexRecordBridge(long startOffset, long rawLength, long partLength, String pathMOF) {
		this.startOffset = startOffset;
		this.rawLength = rawLength;
		this.partLength = partLength;
		this.pathMOF = pathMOF;
	}
}
