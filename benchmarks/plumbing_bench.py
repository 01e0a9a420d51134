#!/usr/bin/env python
"""BASELINE.json config 1: protocol plumbing on CPU — 4 mappers x 2 reducers over the TCP loop-back transport,
TeraSort-shaped records, native provider + consumer roles driven through the same C API the JNI bridge uses.

This is also the CPU-merger baseline of BASELINE.md: the consumer's heap merge is the restatement of the reference's
single merge thread per reducer (src/Merger/MergeQueue.h:250-269, src/Merger/StreamRW.cc:151-225), fed 1 MiB chunks
by the provider's chunk pool (src/MOFServer/IndexInfo.cc:237-335); the reference itself cannot be built on this box
(no verbs / rdmacm / libaio / JDK).

Prints one JSON line; the value is merged bytes delivered to the "reducers" per second of wall clock, all reducers
together (they run concurrently, one native merge thread each, like one reduce task per JVM).
"""
from __future__ import annotations

import argparse
import json
import os
import struct
import sys
import tempfile
import time
import zlib

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np


def make_mof(path: str, n: int, reducers: int, rng) -> np.ndarray:
    """Writes file.out with `reducers` sorted partitions of 100-byte records; returns the index [reducers, 3]."""
    keys = rng.integers(0, 256, size=(n, 10), dtype=np.uint8)
    part = (keys[:, 0].astype(np.int64) * reducers) >> 8
    order = np.lexsort(tuple(keys[:, c] for c in range(9, -1, -1)) )      # total order on the 10 key bytes
    order = order[np.argsort(part[order], kind="stable")]
    rec = np.empty((n, 102), dtype=np.uint8)
    rec[:, 0], rec[:, 1] = 10, 90
    rec[:, 2:12] = keys
    rec[:, 12:] = rng.integers(0, 256, size=(n, 90), dtype=np.uint8)
    rec = rec[order]
    counts = np.bincount(part, minlength=reducers)
    index = np.zeros((reducers, 3), dtype=np.int64)
    pos = 0
    with open(path, "wb") as f:
        start = 0
        for r in range(reducers):
            body = rec[start:start + counts[r]].tobytes() + b"\xff\xff"
            seg = body + struct.pack(">I", zlib.crc32(body) & 0xFFFFFFFF)
            f.write(seg)
            index[r] = (pos, len(body), len(seg))
            pos += len(seg)
            start += counts[r]
    return index


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=float, default=1e9, help="total record bytes over all maps")
    ap.add_argument("--maps", type=int, default=4)
    ap.add_argument("--reducers", type=int, default=2)
    ap.add_argument("--approach", type=int, default=1, help="1 online merge, 2 hybrid (LPQ spill files + final merge)")
    ap.add_argument("--buf-kb", type=int, default=1024)
    a = ap.parse_args()
    from uda_b200 import bridge

    n_per_map = int(a.bytes / 100 / a.maps)
    rng = np.random.default_rng(7)
    with tempfile.TemporaryDirectory(prefix="uda_plumb_") as tmp:
        mofs = {}
        for m in range(a.maps):
            p = os.path.join(tmp, f"map_{m}.out")
            mofs[f"attempt_m_{m:06d}_0"] = (p, make_mof(p, n_per_map, a.reducers, rng))

        def get_path(job, mapid, reduce):
            p, idx = mofs[mapid]
            return int(idx[reduce, 0]), int(idx[reduce, 1]), int(idx[reduce, 2]), p

        prov = bridge.NativeRole(False, bridge.build_argv(port=0, buf_kb=a.buf_kb, provider=True), get_path=get_path)
        cons = []
        want = [sum(int(idx[r, 1]) - 2 for _, idx in mofs.values()) // 102 for r in range(a.reducers)]
        t0 = time.perf_counter()
        for r in range(a.reducers):
            c = bridge.NativeRole(True, bridge.build_argv(port=prov.port, approach=a.approach, buf_kb=a.buf_kb))
            init = [str(a.maps), "job_1", f"attempt_r_{r:06d}_0", "0", str(a.buf_kb * 1024), str(16 * 1024), "org.apache.hadoop.io.LongWritable",
                    "null", "0", str(1 << 30), "1", tmp]
            assert c.command(bridge.form_cmd(bridge.INIT, init)) == 0, c.last_error
            for mapid in mofs:
                assert c.command(bridge.form_cmd(bridge.FETCH, ["127.0.0.1", "job_1", mapid, str(r)])) == 0
            cons.append(c)
        ok = True
        for r, c in enumerate(cons):
            while c.stat("records_emitted") < want[r] and not c.failed and time.perf_counter() - t0 < 600:
                time.sleep(0.005)
            ok = ok and not c.failed and c.stat("records_emitted") == want[r]
        phase = [{k: c.stat(k) for k in ("device_merges", "fetch_ms", "device_ms", "fetch_stalls", "fetch_stall_us")} for c in cons]
        dt = time.perf_counter() - t0
        # validation: order inside every reducer's stream + byte count (TeraValidate-style, done after the clock stops)
        out_bytes = 0
        for r, c in enumerate(cons):
            time.sleep(0.05)
            s = np.frombuffer(c.stream(), dtype=np.uint8)
            out_bytes += len(s) - 2
            ok = ok and len(s) == want[r] * 102 + 2 and bytes(s[-2:]) == b"\xff\xff"
            k = s[: want[r] * 102].reshape(-1, 102)[:, 2:12]
            a8 = k[:, :8].copy().view(">u8").ravel()
            t2 = k[:, 8:10].copy().view(">u2").ravel()
            ok = ok and bool(np.all((a8[1:] > a8[:-1]) | ((a8[1:] == a8[:-1]) & (t2[1:] >= t2[:-1]))))
            c.reduce_exit()
            c.close()
        if os.environ.get("UDA_BENCH_LOGS"):
            for c in cons:
                for sev, m in c.logs:
                    if "device merge" in m: print(m, file=sys.stderr)
        prov.command(bridge.form_cmd(bridge.EXIT, []))
        prov.close()
    rec_bytes = sum(want) * 100
    print(json.dumps({"metric": "protocol plumbing on CPU: provider -> TCP loop-back -> heap merge -> dataFromUda blocks (wall clock)",
                      "value": rec_bytes / dt / 1e9, "unit": "GB/s", "seconds": dt, "record_bytes": rec_bytes, "stream_bytes": out_bytes,
                      "ok": bool(ok), "validated": bool(ok),
                      "config": {"maps": a.maps, "reducers": a.reducers, "approach": a.approach, "buf_kb": a.buf_kb, "transport": "tcp loopback",
                                 "merge_threads": a.reducers}, "consumer_stats": phase}))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
