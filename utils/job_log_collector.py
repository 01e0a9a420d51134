#!/usr/bin/env python
"""Collects, for one job directory, the log lines that matter for post-mortems and for the regression analyzer
(counterpart of utils/master/job-log-collector.sh:22-60 + utils/slave/job-log-collector.awk:22-51, which pull
per-job snippets of tracker logs and userlogs from every slave).

Sources: stdout/stderr of the job, the native runtime's own log files (udaNetMerger.log,
hadoop-<user>-udaMOFSupplier-<host>.log) when ``mapred.uda.log.to.unique.file`` was on. Kept lines: version lines,
buffer registration, open/close markers, warnings/errors, fallback, the JSON result."""
from __future__ import annotations

import argparse
import glob
import os
import re
import sys

KEEP = re.compile(r"UDA version is|The version is|After RDMA buffers registration|Using UdaShuffle|Successfully closed|fallback|"
                  r"ERROR|FATAL|WARN|Traceback|Error|MPRESULT|^\{.*\}$|CUDA error|timed? ?out", re.I)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--dir", required=True)
    ap.add_argument("--out", default="-")
    ap.add_argument("--log-dirs", default="/tmp", help="comma separated directories searched for the native log files")
    a = ap.parse_args()
    files = sorted(glob.glob(os.path.join(a.dir, "*.log")))
    for d in a.log_dirs.split(","):
        files += sorted(glob.glob(os.path.join(d, "udaNetMerger.log"))) + sorted(glob.glob(os.path.join(d, "hadoop-*-udaMOFSupplier-*.log")))
    out = []
    for f in files:
        if os.path.abspath(f) == os.path.abspath(a.out):
            continue
        try:
            with open(f, errors="replace") as fh:
                hits = [l.rstrip("\n") for l in fh if KEEP.search(l)]
        except OSError:
            continue
        if hits:
            out.append(f"==== {f} ({len(hits)} lines)")
            out += hits[-200:]
    text = "\n".join(out) + "\n"
    if a.out == "-":
        sys.stdout.write(text)
    else:
        with open(a.out, "w") as fh:
            fh.write(text)
    return 0


if __name__ == "__main__":
    sys.exit(main())
