"""Static cross-checks between the Java plug-in sources and the native side (no JDK on the build boxes: the Java code is never
compiled here, so the names, signatures and constants both sides must agree on are compared as text)."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
JAVA = ROOT / "java"


def _read(p):
    return (ROOT / p).read_text()


def test_command_codes_match_the_native_enum():
    java = _read("java/shared/com/mellanox/hadoop/mapred/UdaCmd.java")
    native = _read("csrc/runtime/protocol.h")
    jvals = dict(re.findall(r"(\w+_COMMAND|NETLEV_REDUCE_LAUNCHED)\s*=\s*(\d+)", java))
    nvals = dict(re.findall(r"(\w+_MSG|RESULT|NETLEV_REDUCE_LAUNCHED)\s*=\s*(\d+)", native))
    pairs = {"EXIT_COMMAND": "EXIT_MSG", "NEW_MAP_COMMAND": "NEW_MAP_MSG", "FINAL_MERGE_COMMAND": "FINAL_MSG", "RESULT_COMMAND": "RESULT",
             "FETCH_COMMAND": "FETCH_MSG", "FETCH_OVER_COMMAND": "FETCH_OVER_MSG", "JOB_OVER_COMMAND": "JOB_OVER_MSG", "INIT_COMMAND": "INIT_MSG",
             "MORE_COMMAND": "MORE_MSG", "NETLEV_REDUCE_LAUNCHED": "NETLEV_REDUCE_LAUNCHED"}
    for j, n in pairs.items():
        assert jvals[j] == nvals[n], (j, n)


_JNI_TYPES = {"V": "void", "I": "int", "Z": "boolean", "Ljava/lang/String;": "String", "Ljava/lang/Object;": "Object", "[Ljava/lang/String;": "String[]"}


def _parse_sig(sig):
    args, ret = re.match(r"\((.*)\)(.*)", sig).groups()
    out = []
    while args:
        m = re.match(r"(\[?L[^;]+;|\[?[VIZJ])", args)
        out.append(_JNI_TYPES[m.group(1)])
        args = args[m.end():]
    return out, _JNI_TYPES[ret]


def test_upcall_signatures_match_what_the_bridge_looks_up():
    bridge = _read("csrc/jni/uda_bridge.cc")
    java = _read("java/shared/com/mellanox/hadoop/mapred/UdaBridge.java")
    lookups = re.findall(r'GetStaticMethodID, g_cls, "(\w+)", "([^"]+)"', bridge)
    assert {n for n, _ in lookups} == {"fetchOverMessage", "dataFromUda", "getPathUda", "getConfData", "logToJava", "failureInUda"}
    for name, sig in lookups:
        args, ret = _parse_sig(sig)
        m = re.search(r"static\s+(\w+(?:\[\])?)\s+" + name + r"\(([^)]*)\)", java)
        assert m, name
        jret = m.group(1)
        jargs = [a.strip().split()[0] for a in m.group(2).split(",") if a.strip()]
        assert jret == ret and jargs == args, (name, sig, jret, jargs)


def test_native_methods_are_exported_with_the_jni_names():
    bridge = _read("csrc/jni/uda_bridge.cc")
    java = _read("java/shared/com/mellanox/hadoop/mapred/UdaBridge.java")
    natives = re.findall(r"private static native \w+ (\w+)\(", java)
    assert set(natives) == {"startNative", "doCommandNative", "reduceExitMsgNative", "setLogLevelNative"}
    for n in natives:
        assert f"Java_com_mellanox_hadoop_mapred_UdaBridge_{n}(" in bridge, n
    assert "package com.mellanox.hadoop.mapred;" in java


def test_block_size_and_buffer_count_agree():
    """Java copies every dataFromUda block into one of two 1 MiB buffers; the native consumer never emits a larger block
    (a single larger record travels alone and the Java buffer grows)."""
    java = _read("java/shared/com/mellanox/hadoop/mapred/UdaPlugin.java")
    assert re.search(r"KV_BUF_SIZE\s*=\s*1\s*<<\s*20", java) and re.search(r"KV_BUF_NUM\s*=\s*2", java)
    assert "buf.bytes = new byte[len]" in java        # grows for a record larger than a block
    consumer = _read("csrc/runtime/consumer.cc") + _read("csrc/runtime/consumer.h")
    assert re.search(r"1\s*<<\s*20|1048576|kStagingBytes|kBlockBytes", consumer)


def test_every_flavour_has_its_plugin_classes_and_the_makefile_knows_it():
    mk = _read("java/Makefile")
    for flavour, need in {"hadoop-1.x": ["UdaShuffleConsumerPlugin.java", "UdaShuffleProviderPlugin.java", "UdaPluginTT.java"],
                          "hadoop-1.x-old": ["UdaShuffleConsumerPlugin.java", "UdaShuffleProviderPlugin.java"],
                          "hadoop-2.0.x": ["UdaShuffleHandler.java"],
                          "hadoop-2.x": ["UdaShuffleHandler.java", "UdaPluginSH.java", "UdaShuffleConsumerPlugin.java"]}.items():
        have = {p.name for p in (JAVA / flavour).rglob("*.java")}
        assert set(need) <= have, (flavour, have)
        assert f"SRC_{flavour}" in mk
    for t in ("TestUdaCmd.java", "TestShufflePluginApi.java"):
        assert (JAVA / "test" / "com" / "mellanox" / "hadoop" / "mapred" / t).exists()
    for f in ("ShuffleConsumerPlugin.java", "ShuffleProviderPlugin.java"):
        assert (JAVA / "plugin-points" / "hadoop-1.x" / "org" / "apache" / "hadoop" / "mapred" / f).exists()


def test_configuration_keys_in_java_are_known_to_the_typed_config():
    from uda_b200.config import UdaConfig

    known = set(UdaConfig().as_dict().keys()) if hasattr(UdaConfig(), "as_dict") else None
    text = "".join(p.read_text() for p in (JAVA / "shared").rglob("*.java"))
    keys = set(re.findall(r'"(mapred\.(?:rdma|netmerger|uda)\.[a-z0-9.]+)"', text))
    assert {"mapred.rdma.wqe.per.conn", "mapred.rdma.cma.port", "mapred.rdma.buf.size"} <= keys
    if known is not None:
        assert keys <= known, keys - known
