"""Typed configuration and the failure -> fallback controller."""
import ctypes as C
import logging

import pytest

from uda_b200.config import UdaConfig
from uda_b200.runtime.failure import FallbackController, FallbackImpossible, ShuffleTimeout, UdaFailure


def test_defaults_match_reference_table():
    c = UdaConfig()
    d = c.to_dict()
    assert d["mapred.rdma.wqe.per.conn"] == "256" and d["mapred.rdma.cma.port"] == "9011"
    assert d["mapred.netmerger.merge.approach"] == "1"
    assert d["mapred.rdma.buf.size"] == "1024" and d["mapred.rdma.buf.size.min"] == "16"
    assert d["mapred.rdma.shuffle.total.size"] == "0" and d["mapred.job.shuffle.input.buffer.percent"] == "0.7"
    assert d["mapred.netmerger.hybrid.lpq.size"] == "0" and d["mapred.rdma.num.parallel.lpqs"] == "0"
    assert d["mapred.rdma.compression.buffer.ratio"] == "0.2" and d["mapred.rdma.mem.use.contig.pages"] == "1"
    assert d["mapred.uda.log.to.unique.file"] == "false" and d["mapred.rdma.developer.mode"] == "false"
    assert d["mapred.uda.provider.blocked.threads.per.disk"] == "1"
    assert c.parallel_lpqs() == 3 and c.lpq_size(100) == 10 and c.lpq_size(1) == 1


def test_from_dict_types_and_unknown_keys():
    c = UdaConfig.from_dict({"mapred.rdma.buf.size": "512", "mapred.rdma.developer.mode": "TRUE", "mapred.local.dir": "/a,/b",
                             "mapred.child.java.opts": "-server -Xmx2g -Dx=y", "some.other.key": "v",
                             "uda.b200.shuffle.subpartitions": 4})
    assert c.buf_size_kb == 512 and c.developer_mode is True and c.local_dirs == ["/a", "/b"] and c.subparts == 4
    assert c.get("some.other.key") == "v" and c.get("missing", "dflt") == "dflt"
    assert c.shuffle_memory_bytes() == int((2 << 30) * 0.7)
    assert c.replace(shuffle_total_size=12345).shuffle_memory_bytes() == 12345
    with pytest.raises(ValueError):
        UdaConfig.from_dict({"mapred.rdma.buf.size": "8"})          # below the 16 KiB minimum
    with pytest.raises(ValueError):
        UdaConfig.from_dict({"mapred.netmerger.merge.approach": "7"})
    assert UdaConfig.from_dict({"mapred.netmerger.merge.approach": "3"}).consumer_argv("/l")[5] == "3"   # device merge


def test_from_hadoop_xml(tmp_path):
    a, b = tmp_path / "mapred-site.xml", tmp_path / "override.xml"
    a.write_text("<configuration><property><name>mapred.rdma.cma.port</name><value>9100</value></property>"
                 "<property><name>mapred.compress.map.output</name><value>true</value></property>"
                 "<property><name>mapred.map.output.compression.codec</name>"
                 "<value>org.apache.hadoop.io.compress.SnappyCodec</value></property></configuration>")
    b.write_text("<configuration><property><name>mapred.rdma.cma.port</name><value> 9200 </value></property></configuration>")
    c = UdaConfig.from_hadoop_xml(str(a), str(b))
    assert c.cma_port == 9200 and c.codec_short == "snappy" and c.codec_block_size() == 256 * 1024


def test_argv_and_init_params_parse_in_the_native_runtime(built):
    """The argv / INIT strings built here are accepted by the C++ parsers the JNI path uses."""
    from uda_b200 import bridge

    L = bridge.lib()
    c = UdaConfig.from_dict({"mapred.rdma.wqe.per.conn": 64, "mapred.rdma.cma.port": 9555, "mapred.netmerger.merge.approach": 2,
                             "mapred.rdma.buf.size": 256, "mapred.mapoutput.key.class": "org.apache.hadoop.io.LongWritable"})
    argv = c.consumer_argv("/tmp/logs", log_threshold=5)
    assert argv == ["-w", "64", "-r", "9555", "-a", "2", "-m", "1", "-g", "/tmp/logs", "-s", "256", "-t", "5"]
    assert c.provider_argv("/var/log")[:4] == ["-w", "64", "-r", "9555"] and "-a" not in c.provider_argv("/var/log")
    init = c.init_params(40, "job_1", "attempt_1_r_0", ["/d0", "/d1"])
    assert init[0] == "40" and init[3] == "0" and init[4] == str(256 * 1024) and init[5] == str(16 * 1024)
    assert init[6].endswith("LongWritable") and init[7] == "null" and init[10:] == ["2", "/d0", "/d1"]
    cmd = bridge.form_cmd(bridge.INIT, init)
    assert cmd.startswith(f"{len(init) + 1}:{bridge.INIT}:40:job_1:")
    out = C.create_string_buffer(4096)
    n = L.uda_rt_parse_cmd(cmd.encode(), out, len(out)) if hasattr(L, "uda_rt_parse_cmd") else None
    if n is not None:
        assert n >= 0 and b"attempt_1_r_0" in out.value


def test_fallback_controller_contract(caplog):
    calls = []
    fc = FallbackController(on_failure=calls.append)
    with caplog.at_level(logging.ERROR, logger="org.apache.hadoop.mapred.ShuffleConsumerPlugin"):
        with fc.guard("shuffle") as active:
            assert active
            raise ShuffleTimeout("flags stuck")
    assert fc.failed and len(calls) == 1 and isinstance(calls[0], UdaFailure)
    assert "will try to fallback to vanilla" in caplog.text            # the line the reference's automation greps
    with fc.guard("merge") as active:                                   # commands after a failure are ignored
        assert active is False
    assert len(calls) == 1
    dev = FallbackController(on_failure=calls.append, developer_mode=True)
    with pytest.raises(ShuffleTimeout):
        with dev.guard("shuffle"):
            raise ShuffleTimeout("x")
    assert len(calls) == 1                                              # developer mode aborts instead of falling back
    late = FallbackController(on_failure=calls.append)
    late.mark_delivered()
    with pytest.raises(FallbackImpossible):
        with late.guard("merge"):
            raise RuntimeError("CUDA error after the first block was delivered")


def test_affinity_helper_is_safe_without_a_gpu():
    from uda_b200.utils.affinity import _parse_cpulist, bind_to_gpu_numa
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11} and _parse_cpulist("") == set()
    import os
    before = os.sched_getaffinity(0)
    info = bind_to_gpu_numa(0)
    assert info["bound"] is False and os.sched_getaffinity(0) == before


def test_log_collector_and_pre_req(tmp_path):
    """utils/job_log_collector.py keeps the lines post-mortems need; scripts/regression/pre_req.py reports the host state."""
    import json, subprocess, sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    (tmp_path / "stdout.log").write_text('noise\nUDA version is b200-1.0.1\n{"metric": "x", "value": 1}\nmore noise\n')
    (tmp_path / "stderr.log").write_text("blah\nUDA has encountered a critical error and will try to fallback to vanilla MSG=x\nTraceback (most recent call last):\n")
    out = tmp_path / "collected.log"
    p = subprocess.run([sys.executable, str(root / "utils" / "job_log_collector.py"), "--dir", str(tmp_path), "--out", str(out), "--log-dirs", str(tmp_path)],
                       capture_output=True, text=True, timeout=60)
    assert p.returncode == 0
    text = out.read_text()
    assert "UDA version is" in text and "fallback" in text and '"metric"' in text and "Traceback" in text and "noise" not in text
    q = subprocess.run([sys.executable, str(root / "scripts" / "regression" / "pre_req.py"), "--json", "--min-free-ram-gb", "0.001", "--min-free-disk-gb", "0.001"],
                       capture_output=True, text=True, timeout=60)
    r = json.loads(q.stdout)
    assert q.returncode == 0 and r["failed"] == [] and r["cpus"] >= 1 and r["dir_writable"] is True


def test_job_kwargs_from_config():
    """Zeros in the uda.b200.* keys resolve to the defaults bench.py measured (docs/TUNING.md)."""
    c = UdaConfig()
    base = {"merge_ctas_per_sm": 0, "arena_slots": 0, "device_cuts": False, "local_copy": "auto"}
    assert c.terasort_job_kwargs(1) == {"subparts": 4, "push_mode": "lsu", "push_ctas": 296, **base}
    assert c.terasort_job_kwargs(2) == {"subparts": 8, "push_mode": "tma", "push_ctas": 32, **base}
    assert c.terasort_job_kwargs(4) == {"subparts": 12, "push_mode": "tma", "push_ctas": 36, **base}
    assert c.terasort_job_kwargs(8) == {"subparts": 12, "push_mode": "tma", "push_ctas": 48, **base}
    c2 = UdaConfig.from_dict({"uda.b200.shuffle.subpartitions": 4, "uda.b200.shuffle.push.ctas": 24, "uda.b200.shuffle.arena.slots": 2,
                              "uda.b200.shuffle.device.cuts": "true", "uda.b200.shuffle.local.copy": "ce"})
    assert c2.terasort_job_kwargs(8) == {"subparts": 4, "push_mode": "tma", "push_ctas": 24, "merge_ctas_per_sm": 0, "arena_slots": 2,
                                         "device_cuts": True, "local_copy": "ce"}
    # one GPU without pipelining: the push kernel moves everything, then the merge runs
    assert UdaConfig.from_dict({"uda.b200.shuffle.local.copy": "push"}).terasort_job_kwargs(1)["subparts"] == 1
    assert UdaConfig.from_dict({"uda.b200.shuffle.subpartitions": 1}).terasort_job_kwargs(1)["push_ctas"] == 0
    import pytest
    with pytest.raises(ValueError):
        c2.terasort_job_kwargs(1)                                   # a receive ring needs the copy-engine push
    with pytest.raises(ValueError):
        UdaConfig.from_dict({"uda.b200.shuffle.subpartitions": 4, "uda.b200.shuffle.arena.slots": 5})
    with pytest.raises(ValueError):
        UdaConfig.from_dict({"uda.b200.shuffle.local.copy": "dma"})
