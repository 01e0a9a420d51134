"""Merge approach 3 (csrc/runtime/device_merge.cc) on a machine without a GPU, against the host-only mock of the kernel
library (tests/native/mock/mock_kernels.cc, selected with UDA_KERNELS_LIB).

The device-merge back-end is a process-wide singleton that binds its library once, so every scenario runs in a
subprocess: the outer tests (collected by the normal CPU suite) start `pytest` on the inner tests of this file or on
tests/test_device_merge_gpu.py with the environment that selects the mock. The mock is synchronous host code: these tests
check the host logic of the device merge (upload layout, fixed-shape detection, windowed delivery, fall-back rules), not
stream ordering or the kernels (GPU suite).
"""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(Path(__file__).resolve().parent))
INNER = os.environ.get("UDA_MOCK_INNER") == "1"
inner = pytest.mark.skipif(not INNER, reason="runs in a subprocess started by the outer tests of this file")
outer = pytest.mark.skipif(INNER, reason="outer test: not inside its own subprocess")


def _mock_path() -> Path:
    rt = os.environ.get("UDA_B200_RUNTIME_LIB")                      # sanitizer runs: the mock built with the same flags
    if rt and (Path(rt).parent / "libuda_kernels_mock.so").exists():
        return Path(rt).parent / "libuda_kernels_mock.so"
    return ROOT / "build" / "bin" / "libuda_kernels_mock.so"


def _pytest_in_subprocess(args, extra_env=None, timeout=600):
    env = dict(os.environ, UDA_KERNELS_LIB=str(_mock_path()), UDA_B200_RUN_GPU_MARKED="1", UDA_MOCK_INNER="1")
    env.update(extra_env or {})
    p = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-p", "no:cacheprovider", *args], cwd=str(ROOT), env=env,
                       capture_output=True, text=True, timeout=timeout)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    last = [l for l in p.stdout.splitlines() if " passed" in l or " failed" in l or " skipped" in l][-1]
    assert " passed" in last and "failed" not in last, tail
    return last


# ------------------------------------------------------------------------------------------------------ outer tests
@outer
def test_the_gpu_device_merge_tests_pass_against_the_mock(built):
    """Every scenario of tests/test_device_merge_gpu.py (general path, fused fixed-shape path incl. Text keys, Snappy
    segments) through provider -> TCP -> consumer -> device_merge.cc -> mock library."""
    assert _mock_path().exists(), "tools/build.py builds build/bin/libuda_kernels_mock.so"
    last = _pytest_in_subprocess(["tests/test_device_merge_gpu.py"])
    assert "skipped" not in last, last


@outer
def test_mock_scenarios(built):
    last = _pytest_in_subprocess([__file__, "-k", "inner and not no_device"])
    assert "skipped" not in last.split("deselected")[0] or "passed" in last


@outer
def test_without_a_device_the_consumer_merges_on_the_cpu(built):
    _pytest_in_subprocess([__file__, "-k", "inner_no_device"], {"UDA_MOCK_DEVICES": "0"})


# ------------------------------------------------------------------------------------------------------ inner tests
def _job(tmp_path, **kw):
    import test_runtime_cpu as T
    return T, T._run_job(str(tmp_path), **kw)


def _check(T, streams, truth, exact=True):
    from uda_b200.oracle.merge import kway_merge
    for r, s in enumerate(streams):
        want = list(kway_merge(truth[r]))
        got = T.parse_stream(s)
        if exact:
            assert got == want and s == T.merged_stream(want)
        else:
            assert [k for k, _ in got] == [k for k, _ in want] and sorted(got) == sorted(want)


@inner
def test_inner_general_path_and_forced_general_path(tmp_path):
    os.environ.pop("UDA_MOCK_NO_FUSED", None)
    T, (streams, truth, cons, _) = _job(tmp_path / "a", n_maps=6, n_reduces=2, recs=500, approach=3, key_len=24, seed=11)
    _check(T, streams, truth)
    assert all(c.final_stats["device_merges"] == 1 for c in cons)
    fused_before = cons[-1].final_stats["device_fused_merges"]
    # fixed-shape records while the library says it has no fused merge for them: the general path must serve them
    os.environ["UDA_MOCK_NO_FUSED"] = "1"
    try:
        T, (streams, truth, cons, _) = _job(tmp_path / "b", n_maps=4, n_reduces=1, recs=800, approach=3, key_len=10, val_len=30, seed=5)
    finally:
        del os.environ["UDA_MOCK_NO_FUSED"]
    _check(T, streams, truth, exact=False)
    assert cons[0].final_stats["device_merges"] == 1 and cons[0].final_stats["device_fused_merges"] == fused_before
    # and with the fused merge back
    T, (streams, truth, cons, _) = _job(tmp_path / "c", n_maps=4, n_reduces=1, recs=800, approach=3, key_len=10, val_len=30, seed=5)
    _check(T, streams, truth, exact=False)
    assert cons[0].final_stats["device_fused_merges"] == fused_before + 1


@inner
def test_inner_record_larger_than_a_block_and_empty_partitions(tmp_path):
    import test_runtime_cpu as T
    big = bytes(range(256)) * 5000
    parts = [[[(b"a" * 10, b"x"), (b"m" * 10, big), (b"z" * 10, b"y")], []],
             [[(b"b" * 10, b"p" * 100), (b"n" * 10, b"q" * 100)], []],
             [[(b"c" * 10, b""), (b"m" * 10 + b"!", b"after the big one")], []]]
    (tmp_path / "d").mkdir()
    streams, truth, cons, _ = T._run_job(str(tmp_path / "d"), 3, 2, 0, approach=3, prebuilt=T._prebuilt_mofs(str(tmp_path / "d"), parts),
                                         max_block=len(big) + 64)
    _check(T, streams, truth)
    assert streams[1] == b"\xff\xff"                                  # a reducer without any record still gets its EOF marker
    sizes = [len(b) for b in cons[0].blocks]
    assert sum(1 for x in sizes if x > (1 << 20)) == 1


@inner
def test_inner_failures_before_the_first_block_fall_back_to_the_cpu_merge(tmp_path):
    """Allocation failure and too little free memory are reported before anything was delivered: the consumer logs why
    and heap-merges the segments it holds; the stream is the same."""
    import ctypes as C
    mock = C.CDLL(os.environ["UDA_KERNELS_LIB"])
    mock.uda_mock_live_allocations.restype = C.c_longlong
    mock.uda_mock_d2h_count.restype = C.c_longlong
    for knob, value, needle in (("UDA_MOCK_FREE_BYTES", "4096", "MiB free"),
                                ("UDA_MOCK_FAIL_MALLOC_AT", lambda: _allocations_so_far(mock) + 3, "allocation failed"),   # the third allocation of the merge
                                ("UDA_MOCK_FAIL_D2H_AT", lambda: int(mock.uda_mock_d2h_count()) + 2, "illegal memory access")):  # the first window's copy
        os.environ[knob] = str(value() if callable(value) else value)
        try:
            T, (streams, truth, cons, _) = _job(tmp_path / knob, n_maps=5, n_reduces=1, recs=300, approach=3, key_len=24, seed=3)
        finally:
            del os.environ[knob]
        _check(T, streams, truth)
        assert cons[0].final_stats["device_merges"] == 0
        assert any("device merge not possible" in m and needle in m for _, m in cons[0].logs), [m for _, m in cons[0].logs][-5:]
        assert mock.uda_mock_live_allocations() == 0                  # nothing leaks on the failure path


@inner
def test_inner_kernel_errors_before_delivery(tmp_path):
    """A kernel entry point that reports an error: in the general path the whole device merge is abandoned for the CPU merge
    (nothing was delivered yet); a failing fused merge only demotes the partition to the general device path."""
    for fn in ("uda_extract_keys_var", "uda_tiebreak_var", "uda_gather_var"):
        os.environ["UDA_MOCK_FAIL_FN"] = fn
        try:
            T, (streams, truth, cons, _) = _job(tmp_path / fn, n_maps=4, n_reduces=1, recs=300, approach=3, key_len=24, seed=9)
        finally:
            del os.environ["UDA_MOCK_FAIL_FN"]
        _check(T, streams, truth)
        assert cons[0].final_stats["device_merges"] == 0
        assert any("device merge not possible" in m and fn in m for _, m in cons[0].logs), [m for _, m in cons[0].logs][-4:]
    os.environ["UDA_MOCK_FAIL_FN"] = "uda_merge_records_fixed"
    try:
        T, (streams, truth, cons, _) = _job(tmp_path / "fused", n_maps=4, n_reduces=1, recs=500, approach=3, key_len=10, val_len=20, seed=2)
    finally:
        del os.environ["UDA_MOCK_FAIL_FN"]
    _check(T, streams, truth, exact=False)
    assert cons[0].final_stats["device_merges"] == 1
    assert any("fused record merge not used" in m for _, m in cons[0].logs)


def _allocations_so_far(mock):
    import ctypes as C
    mock.uda_mock_allocation_count.restype = C.c_longlong
    return int(mock.uda_mock_allocation_count())


@inner
def test_inner_failure_after_the_first_block_fails_the_task(tmp_path):
    """A copy that fails once blocks have been handed to Java cannot be hidden by a CPU re-merge (the stream cannot be
    restarted): the task fails over to the vanilla shuffle through failureInUda."""
    import ctypes as C
    mock = C.CDLL(os.environ["UDA_KERNELS_LIB"])
    mock.uda_mock_d2h_count.restype = C.c_longlong
    import test_runtime_cpu as T
    # 80 MiB of records: three 32 MiB windows. Copies of this merge: record offsets, window 0, window 1 (started before
    # window 0 is delivered), window 2 (started after window 0 went out): the fourth copy fails with blocks already delivered
    val = b"v" * 4000
    parts = [[[(bytes([65 + m]) + i.to_bytes(9, "big"), val) for i in range(7000)]] for m in range(3)]
    (tmp_path / "e").mkdir()
    os.environ["UDA_MOCK_FAIL_D2H_AT"] = str(int(mock.uda_mock_d2h_count()) + 4)
    try:
        streams, _truth, cons, _ = T._run_job(str(tmp_path / "e"), 3, 1, 0, approach=3, prebuilt=T._prebuilt_mofs(str(tmp_path / "e"), parts),
                                              expect_failure=True)
    finally:
        del os.environ["UDA_MOCK_FAIL_D2H_AT"]
    assert 0 < len(streams[0]) < 3 * 7000 * 4010                      # some blocks were delivered, not all


def _big_parts(n_maps, recs, val, var_len):
    """Sorted map outputs with random 10-byte keys, generated quickly (os.urandom): ~recs * (len(val) + 14) bytes per map."""
    parts = []
    for m in range(n_maps):
        blob = os.urandom(recs * 10)
        keys = sorted(blob[i * 10:(i + 1) * 10] for i in range(recs))
        parts.append([[(k, val[: len(val) - (k[0] % 7 if var_len else 0)]) for k in keys]])
    return parts


@inner
@pytest.mark.parametrize("var_len", [True, False], ids=["general-path", "fused-path"])
def test_inner_partition_beyond_device_memory_is_merged_in_groups(tmp_path, var_len):
    """Free device memory below what the whole partition needs, above what half of it needs: the segments are merged on
    the device in groups into host runs (the hybrid merge's LPQs with the device as the merger) and a heap merges the
    runs; with memory for no group at all the CPU merges everything."""
    import test_runtime_cpu as T
    parts = _big_parts(8, 12_000, b"v" * 100, var_len)               # 8 x 1.4 MB
    total = sum(len(k) + len(v) + 2 for p in parts for k, v in p[0])
    d = tmp_path / "grp"
    d.mkdir()
    prebuilt = T._prebuilt_mofs(str(d), parts)
    # the fused path needs ~2x the data (input + output), the general path ~3x: pick a budget between half and all of it
    need_all = total * (3.6 if var_len else 2.1)
    os.environ["UDA_MOCK_FREE_BYTES"] = str(int(need_all * 0.62 / 0.9))
    try:
        streams, truth, cons, _ = T._run_job(str(d), 8, 1, 0, approach=3, prebuilt=prebuilt)
    finally:
        del os.environ["UDA_MOCK_FREE_BYTES"]
    _check(T, streams, truth, exact=False)
    st = cons[0].final_stats
    assert st["device_groups"] in (2, 4) and st["device_merges"] == st["device_groups"], st
    assert any("merging in groups" in m for _, m in cons[0].logs) and any("final merge of the" in m for _, m in cons[0].logs)
    # memory for nothing: every grouping is refused, the CPU merges
    os.environ["UDA_MOCK_FREE_BYTES"] = "100000"
    try:
        streams, truth, cons, _ = T._run_job(str(d), 8, 1, 0, approach=3, prebuilt=prebuilt)
    finally:
        del os.environ["UDA_MOCK_FREE_BYTES"]
    _check(T, streams, truth, exact=False)
    assert cons[0].final_stats["device_merges"] == 0
    assert any("no grouping fits either" in m and "merging on the CPU" in m for _, m in cons[0].logs)


@inner
def test_inner_no_device(tmp_path):
    T, (streams, truth, cons, _) = _job(tmp_path, n_maps=4, n_reduces=1, recs=200, approach=3, seed=8)
    _check(T, streams, truth)
    assert cons[0].final_stats["device_merges"] == 0
    assert any("device merge not possible" in m and "no CUDA device visible" in m for _, m in cons[0].logs)
