#!/usr/bin/env python
"""The plug-in surface end to end without Hadoop or a GPU: Python plays the Java side of the JNI bridge.

One provider role (MOFSupplier) serves two map outputs from local files; one consumer role (NetMerger) is told about
the job (INIT), asked to fetch both maps (FETCH) and hands back the merged stream in blocks (dataFromUda). This is the
flow of SURVEY §3.1-3.4 through the C API that the JNI bridge uses (include/uda_rt_api.h).

    python examples/fake_hadoop_plumbing.py [--approach 1|2|3]
"""
import argparse
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from uda_b200 import bridge
from uda_b200.config import UdaConfig
from uda_b200.oracle import kway_merge, parse_stream
from uda_b200.utils import ifile


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--approach", type=int, default=1, help="1 online, 2 hybrid, 3 device merge (CPU fallback without a GPU)")
    a = ap.parse_args()
    conf = UdaConfig.from_dict({"mapred.netmerger.merge.approach": a.approach, "mapred.rdma.buf.size": 64,
                                "mapred.mapoutput.key.class": "org.apache.hadoop.io.LongWritable"})
    with tempfile.TemporaryDirectory() as tmp:
        # two map outputs, one partition each: file.out + file.out.index exactly as Hadoop writes them
        runs = [sorted((bytes([i, j]) * 4, b"value-%d-%d" % (m, i)) for i in range(40) for j in range(3)) for m in range(2)]
        mofs = {}
        for m, recs in enumerate(runs):
            data, index = ifile.write_mof([recs])
            path = os.path.join(tmp, f"map_{m}.out")
            open(path, "wb").write(data)
            mofs[f"attempt_m_{m:06d}_0"] = (path, ifile.read_index(index))

        def get_path(job, map_id, reduce):          # Java: UdaPluginSH/TT.getPathIndex -> IndexRecordBridge
            path, idx = mofs[map_id]
            r = idx[reduce]
            return r.start_offset, r.raw_length, r.part_length, path

        # port 0 = let the provider pick a free port (a real deployment uses mapred.rdma.cma.port on both sides)
        provider = bridge.NativeRole(False, conf.replace(cma_port=1).provider_argv(tmp)[:3] + ["0"] + conf.provider_argv(tmp)[4:],
                                     get_path=get_path)
        consumer = bridge.NativeRole(True, conf.replace(cma_port=provider.port).consumer_argv(tmp), conf=conf.to_dict())
        init = conf.init_params(num_maps=2, job_id="job_1", reduce_attempt_id="attempt_r_000000_0", local_dirs=[tmp])
        assert consumer.command(bridge.form_cmd(bridge.INIT, init)) == 0, consumer.last_error
        for map_id in mofs:
            assert consumer.command(bridge.form_cmd(bridge.FETCH, ["127.0.0.1", "job_1", map_id, "0"])) == 0
        want = list(kway_merge(runs))
        deadline = time.time() + 30
        while time.time() < deadline and consumer.stat("records_emitted") < len(want) and not consumer.failed:
            time.sleep(0.01)
        time.sleep(0.05)
        got = parse_stream(consumer.stream())
        print(f"approach {a.approach}: {len(got)} records in {len(consumer.blocks)} block(s), fetchOverMessage x{consumer.fetch_over}, "
              f"device merges: {consumer.stat('device_merges')}")
        for sev, msg in consumer.logs:
            if "version" in msg or "device merge" in msg:
                print("   log:", msg)
        ok = sorted(got) == sorted(want) and [k for k, _ in got] == [k for k, _ in want]
        consumer.reduce_exit()
        consumer.close()
        provider.command(bridge.form_cmd(bridge.EXIT, []))
        provider.close()
    print("OK" if ok else "MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
