"""CPU tests of the C++ host runtime (libuda.so) through its C API: protocol goldens with the reference's
quirks, queues, comparators, memory planning, Snappy, and the provider<->consumer plumbing path over TCP
loop-back with Python playing Hadoop (BASELINE.json config 1 in miniature)."""
import ctypes as C
import os
import random
import struct
import time

import pytest

from uda_b200 import bridge
from uda_b200.oracle import kway_merge, merged_stream, parse_stream
from uda_b200.utils import ifile, vint


@pytest.fixture(scope="module")
def L(built):
    return bridge.lib()


def _buf(n=4096):
    return C.create_string_buffer(n)


# ------------------------------------------------------------------------------------------ vint
@pytest.mark.parametrize("v", [0, 1, 127, 128, -1, -112, -113, 255, 256, 65535, 65536, -129, 2**31 - 1, -(2**31), 2**63 - 1, -(2**63), 10, 90])
def test_vint_matches_python(L, v):
    out = _buf(16)
    n = L.uda_rt_vint_encode(v, out)
    assert out.raw[:n] == vint.encode_vlong(v) and n == L.uda_rt_vint_size(v)
    got = C.c_longlong()
    assert L.uda_rt_vint_decode(out.raw[:n], n, C.byref(got)) == n and got.value == v
    assert L.uda_rt_vint_decode(out.raw[:n], max(0, n - 1), C.byref(got)) == 0  # truncated


# ------------------------------------------------------------------------------------------ commands
def _parse(L, s):
    out = _buf()
    n = L.uda_rt_parse_cmd(s.encode(), out, len(out))
    if n < 0:
        return None
    parts = out.value.decode().split("\n")
    return int(parts[0]), int(parts[1]), parts[2:]


def test_command_golden(L):
    assert _parse(L, "") == (1, 0, [])                       # empty string == EXIT
    assert _parse(L, "1:0") == (1, 0, [])
    assert _parse(L, "5:4:host1:job_1:attempt_1_m_0:3") == (5, 4, ["host1", "job_1", "attempt_1_m_0", "3"])
    # the last parameter swallows the rest of the string, colons included
    assert _parse(L, "3:7:a:/path/with:colons:inside") == (3, 7, ["a", "/path/with:colons:inside"])
    assert _parse(L, "nocolon") is None
    assert _parse(L, "4:7:onlyone") is None                  # announces 3 params, provides 1
    s = bridge.form_cmd(bridge.FETCH, ["h", "j", "m", "2"])
    assert s == "5:4:h:j:m:2"
    arr = (C.c_char_p * 4)(b"h", b"j", b"m", b"2")
    out = _buf()
    L.uda_rt_form_cmd(4, 4, arr, out, len(out))
    assert out.value.decode() == s


def test_options_golden(L):
    def parse(argv):
        arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
        out = (C.c_longlong * 7)()
        d = _buf(512)
        L.uda_rt_parse_options(len(argv), arr, out, d, len(d))
        return list(out), d.value.decode()

    # Java passes no program name, so getopt skips "-w": its value never takes effect (reference quirk)
    o, d = parse(["-w", "99", "-r", "9012", "-a", "2", "-m", "1", "-g", "/logs/", "-s", "1024"])
    assert o[0] == 256 and o[1] == 9012 and o[2] == 2 and o[3] == 1 and o[4] == 1024 * 1024 and d == "/logs/"
    o, _ = parse(["prog", "-w", "99", "-s", "5"])            # with a program name -w is honoured
    assert o[0] == 99 and o[4] == 4096                       # 5 KiB -> rounded down to 4 KiB
    o, _ = parse(["prog", "-s", "1"])
    assert o[4] == 4096                                      # minimum one 4 KiB block
    o, _ = parse(["prog", "--dataport", "777", "--rdmabufsize=33", "-t", "6", "-b", "x"])
    assert o[1] == 777 and o[4] == 32 * 1024 and o[5] == 6


def test_fetch_request_and_ack_golden(L):
    req = "job_201_0001:attempt_201_0001_m_000003_0:1048576:7:140000:42:1048576:8192:/data/x/file.out:5000:5004"
    out = _buf()
    assert L.uda_rt_fetch_request_roundtrip(req.encode(), out, len(out)) == len(req) and out.value.decode() == req
    assert L.uda_rt_fetch_request_roundtrip(b"a:b:c", out, len(out)) == -1
    long_path = "/p" * 400
    assert L.uda_rt_fetch_request_roundtrip(f"j:m:0:1:2:3:4:5:{long_path}:6:7".encode(), out, len(out)) == -3  # >= 800 bytes
    ack = "5000:5004:4096:8192:/data/x/file.out:"
    assert L.uda_rt_fetch_ack_roundtrip(ack.encode(), out, len(out)) == len(ack) and out.value.decode() == ack
    assert L.uda_rt_fetch_ack_roundtrip(b"1:2:3", out, len(out)) == -1


# ------------------------------------------------------------------------------------------ comparators / queues / planning
def test_comparators(L):
    assert L.uda_rt_key_kind(b"org.apache.hadoop.io.Text") == 1
    assert L.uda_rt_key_kind(b"org.apache.hadoop.io.LongWritable") == 0
    assert L.uda_rt_key_kind(b"org.apache.hadoop.hbase.io.ImmutableBytesWritable") == 2
    assert L.uda_rt_key_kind(b"org.apache.hadoop.io.DoubleWritable") == -1
    t = lambda w: vint.encode_vlong(len(w)) + w
    cmp = lambda k, a, b: L.uda_rt_compare_keys(k, a, len(a), b, len(b))
    assert cmp(1, t(b"ab"), t(b"abc")) < 0 and cmp(1, t(b"b"), t(b"abc")) > 0 and cmp(1, t(b"x" * 200), t(b"x" * 200)) == 0
    assert cmp(0, b"\x00\x01", b"\x00\x02") < 0 and cmp(0, b"\xff", b"\x01") > 0   # unsigned bytes
    assert cmp(2, b"\x00\x00\x00\x09zz", b"\x00\x00\x00\x01zza") < 0


def test_queues_selftest(L):
    assert L.uda_rt_selftest_queues() == 0


def test_thread_wrapper_and_shuffles(L):
    """N4: a throwing native thread raises exactly one failure up-call; the shuffles are permutations and
    deterministic for a seed (src/CommUtils/UdaUtil.cc:39-94; vector_shuffle / list_shuffle_in_vector)."""
    import ctypes as C
    n = 97
    a, b = (C.c_int * n)(), (C.c_int * n)()
    assert L.uda_rt_selftest_threads(C.c_ulonglong(42), n, a) == 0
    assert L.uda_rt_selftest_threads(C.c_ulonglong(42), n, b) == 0
    assert list(a) == list(b) and sorted(a) == list(range(n)) and list(a) != list(range(n))
    assert L.uda_rt_selftest_threads(C.c_ulonglong(43), n, b) == 0
    assert list(a) != list(b)


def test_memory_plan(L):
    out = (C.c_longlong * 4)()
    MiB = 1 << 20
    assert L.uda_rt_plan_memory(10, MiB, 16 * 1024, 1 << 30, 0, 0.2, out) == 0
    assert list(out) == [MiB, 20, MiB, MiB]                  # (num_maps + 10) pairs of full-size buffers
    # shuffle memory too small for 1 MiB buffers: shrink, page aligned
    assert L.uda_rt_plan_memory(100, MiB, 16 * 1024, 100 * MiB, 0, 0.2, out) == 0
    assert out[0] == (100 * MiB // 200) // 4096 * 4096
    assert L.uda_rt_plan_memory(1000, MiB, 64 * 1024, 10 * MiB, 0, 0.2, out) == 1     # "Not enough memory for rdma buffers"
    assert L.uda_rt_plan_memory(1, 1000, 16, 1 << 30, 0, 0.2, out) == 2              # "RDMA Buffer is too small" (< one page)
    assert L.uda_rt_plan_memory(10, MiB, 16 * 1024, 1 << 30, 1, 0.2, out) == 0
    assert out[0] == int(2 * MiB * 0.2) // 4096 * 4096


def test_snappy_builtin(L):
    data = bytes(random.Random(0).getrandbits(8) for _ in range(70000))
    comp = _buf(len(data) + 1024)
    n = L.uda_rt_snappy_store(data, len(data), comp, len(comp))
    out = _buf(len(data))
    assert L.uda_rt_snappy_uncompress(comp.raw[:n], n, out, len(data)) == 0 and out.raw == data
    # hand-made stream with copies: "abcd" + copy(offset 4, len 8) + 2-byte-offset copy
    stream = bytes([16]) + bytes([3 << 2]) + b"abcd" + bytes([((8 - 4) << 2) | 1, 4]) + bytes([(4 - 1) << 2 | 2, 12, 0])
    out = _buf(16)
    assert L.uda_rt_snappy_uncompress(stream, len(stream), out, 16) == 0 and out.raw == b"abcdabcdabcdabcd"
    assert L.uda_rt_snappy_uncompress(stream[:-1], len(stream) - 1, out, 16) == -1


# ------------------------------------------------------------------------------------------ plumbing end to end
def _make_mofs(tmp, n_maps, n_reduces, recs_per_part, seed, key_len=10, compress=None, val_len=None, dirs=None):
    rng = random.Random(seed)
    mofs = {}
    truth = [[] for _ in range(n_reduces)]
    for m in range(n_maps):
        parts = []
        for r in range(n_reduces):
            n = rng.randint(0, recs_per_part)
            recs = sorted(((bytes(rng.getrandbits(8) for _ in range(key_len)), bytes(rng.getrandbits(8) for _ in range(rng.randint(0, 200) if val_len is None else val_len)))
                           for _ in range(n)), key=lambda kv: kv[0])
            parts.append(recs)
            truth[r].append(recs)
        if compress is None:
            data, index = ifile.write_mof(parts)
        else:
            out, idx = bytearray(), []
            for recs in parts:
                plain = ifile.encode_segment(recs, with_crc=False)
                seg = bytearray()
                for off in range(0, len(plain), 3000):      # several codec blocks per segment
                    blk = plain[off:off + 3000]
                    seg += struct.pack(">II", len(blk), len(compress(blk))) + compress(blk)
                seg += b"\0\0\0\0"                           # trailing checksum bytes (ignored by readers)
                idx.append(ifile.IndexRecord(len(out), len(plain), len(seg)))
                out += seg
            data, index = bytes(out), ifile.write_index(idx)
        where = dirs[m % len(dirs)] if dirs else tmp       # map outputs spread over the "disks" of mapred.local.dir
        os.makedirs(where, exist_ok=True)
        path = os.path.join(where, f"map_{m}.out")
        with open(path, "wb") as f:
            f.write(data)
        mofs[f"attempt_m_{m:06d}_0"] = (path, ifile.read_index(index))
    return mofs, truth


def _run_job(tmp, n_maps, n_reduces, recs, approach=1, buf_kb=8, key_class="org.apache.hadoop.io.LongWritable", key_len=10,
             lpq_size=0, compress=None, codec="null", seed=1, val_len=None, provider_conf=None, dirs=None, prebuilt=None,
             max_block=1 << 20, expect_failure=False):
    # (_make_mofs is replaced by callers that need other record shapes: only pass `dirs` when it is used)
    mofs, truth = prebuilt or _make_mofs(tmp, n_maps, n_reduces, recs, seed, key_len, compress, val_len, **({"dirs": dirs} if dirs else {}))

    def get_path(job, mapid, reduce):
        path, idx = mofs[mapid]
        r = idx[reduce]
        return r.start_offset, r.raw_length, r.part_length, path

    prov = bridge.NativeRole(False, bridge.build_argv(port=0, buf_kb=buf_kb, provider=True), get_path=get_path, conf=provider_conf)
    assert not prov.failed, prov.last_error
    port = prov.port
    assert port > 0
    streams, consumers = [], []
    try:
        for r in range(n_reduces):
            cons = bridge.NativeRole(True, bridge.build_argv(port=port, approach=approach, buf_kb=buf_kb))
            consumers.append(cons)
            init = [str(n_maps), "job_1", f"attempt_r_{r:06d}_0", str(lpq_size), str(buf_kb * 1024), str(4096), key_class, codec,
                    str(256 * 1024), str(64 << 20), "1", tmp]
            assert cons.command(bridge.form_cmd(bridge.INIT, init)) == 0, cons.last_error
            for mapid in mofs:
                assert cons.command(bridge.form_cmd(bridge.FETCH, ["127.0.0.1", "job_1", mapid, str(r)])) == 0
        for r, cons in enumerate(consumers):
            deadline = time.time() + 60
            want = sum(len(x) for x in truth[r])
            while time.time() < deadline and not cons.failed:
                s = cons.stream()
                if s.endswith(b"\xff\xff") and cons.stat("records_emitted") == want:
                    break
                time.sleep(0.02)
            if expect_failure:
                assert cons.failed and cons.failures >= 1
                streams.append(cons.stream())
                continue
            assert not cons.failed and cons.failures == 0, cons.last_error
            cons.final_stats = {k: cons.stat(k) for k in ("records_emitted", "lpqs_spilled", "device_merges", "device_fused_merges", "device_groups")}
            streams.append(cons.stream())
            assert all(len(b) <= max_block for b in cons.blocks)
            assert cons.fetch_over >= 1
        return streams, truth, consumers, prov
    finally:
        for c in consumers:
            c.reduce_exit()
            c.close()
        prov.command(bridge.form_cmd(bridge.EXIT, []))
        prov.close()


def test_plumbing_online_4_maps_2_reducers(built, tmp_path):
    streams, truth, consumers, prov = _run_job(str(tmp_path), 4, 2, 400)
    for r, s in enumerate(streams):
        want = list(kway_merge(truth[r]))
        assert parse_stream(s) == want
        assert s == merged_stream(want)


@pytest.mark.parametrize("threads_per_disk", [1, 3])
def test_plumbing_blocked_reader_backend(built, tmp_path, threads_per_disk):
    """The provider's second disk back-end (reference: AsyncReaderManager, src/MOFServer/AsyncReaderManager.cc:19-20 — a
    queue and `mapred.uda.provider.blocked.threads.per.disk` pread threads per local dir); map outputs on two "disks" and
    one outside every configured dir (served by the fallback queue)."""
    d0, d1 = os.path.join(str(tmp_path), "disk0"), os.path.join(str(tmp_path), "disk1")
    conf = {"mapred.uda.provider.reader": "blocked", "mapred.uda.provider.blocked.threads.per.disk": str(threads_per_disk),
            "mapred.local.dir": f"{d0},{d1}"}
    streams, truth, _, prov = _run_job(str(tmp_path), 5, 2, 350, buf_kb=16, seed=12, provider_conf=conf,
                                       dirs=[d0, d1, os.path.join(str(tmp_path), "elsewhere")])
    for r, s in enumerate(streams):
        assert parse_stream(s) == list(kway_merge(truth[r]))
    assert any("reader=blocked" in m for _, m in prov.logs)


@pytest.mark.parametrize("lie", ["sent-larger-than-payload", "partition-shrinks", "negative-length", "chunk-larger-than-buffer"])
def test_consumer_rejects_impossible_fetch_acks(built, lie):
    """The other direction of the hostile-request test: a provider (or anything else on that port) whose acks cannot be
    true must fail the fetch -- and through it the task, which falls back to the vanilla shuffle -- not make the reducer
    parse bytes that never arrived."""
    import socket
    import threading

    srv = socket.socket()
    srv.bind(("127.0.0.1", 0))
    srv.listen(4)
    port = srv.getsockname()[1]
    rec = ifile.encode_record(b"k" * 10, b"v" * 30)

    def serve():
        try:
            conn, _ = srv.accept()
        except OSError:
            return
        with conn:
            try:
                while True:
                    hdr = b""
                    while len(hdr) < 24:
                        c = conn.recv(24 - len(hdr))
                        if not c:
                            return
                        hdr += c
                    _t, _c, _p, tot, _cookie, _pl = struct.unpack("<BBHIQQ", hdr)
                    text = b""
                    while len(text) < tot:
                        text += conn.recv(tot - len(text))
                    cookie = int(text.decode().split(":")[5])
                    payload = rec * 3
                    raw = part = len(payload) + 2
                    sent = len(payload)
                    if lie == "sent-larger-than-payload":
                        sent = 500_000
                    elif lie == "partition-shrinks":
                        raw = part = 10
                    elif lie == "negative-length":
                        raw = -5
                    elif lie == "chunk-larger-than-buffer":
                        payload = rec * 2000                  # 88 KB into the 8 KiB buffers this consumer asked to fill
                        sent, raw, part = len(payload), len(payload) + 2, len(payload) + 2
                    ack = f"{raw}:{part}:{sent}:0:/some/file.out:".encode()
                    conn.sendall(struct.pack("<BBHIQQ", 0x10, 1, 0, len(ack), cookie, len(payload)) + ack + payload)
            except (OSError, ValueError, IndexError):
                return

    t = threading.Thread(target=serve, daemon=True)
    t.start()
    cons = bridge.NativeRole(True, bridge.build_argv(port=port, approach=1, buf_kb=8))
    try:
        init = ["1", "job_1", "attempt_r_000000_0", "0", str(8 * 1024), "4096", "org.apache.hadoop.io.LongWritable", "null", "0", str(1 << 20), "1", "/tmp"]
        assert cons.command(bridge.form_cmd(bridge.INIT, init)) == 0, cons.last_error
        assert cons.command(bridge.form_cmd(bridge.FETCH, ["127.0.0.1", "job_1", "attempt_m_000000_0", "0"])) == 0
        deadline = time.time() + 30
        while time.time() < deadline and not cons.failed:
            time.sleep(0.02)
        assert cons.failed and cons.failures >= 1
        assert cons.stream() == b"" or not cons.stream().endswith(b"\xff\xff")      # no complete stream was fabricated
    finally:
        srv.close()
        cons.reduce_exit()
        cons.close()


def test_log_to_unique_file_writes_the_provider_log(built, tmp_path):
    """mapred.uda.log.to.unique.file = true: the native side logs to <log dir>/hadoop-<user>-udaMOFSupplier-<host>.log
    instead of the host's logger (reference: src/CommUtils/IOUtility.cc log_to_unique_file, build/README 'log files')."""
    logdir = str(tmp_path)
    prov = bridge.NativeRole(False, bridge.build_argv(port=0, buf_kb=64, provider=True, log_dir=logdir), log_to_unique_file=True)
    try:
        assert not prov.failed, prov.last_error
        prov.command(bridge.form_cmd(bridge.EXIT, []))
    finally:
        prov.close()
    files = [f for f in os.listdir(logdir) if "udaMOFSupplier" in f and f.endswith(".log")]
    assert len(files) == 1 and files[0].startswith("hadoop-")
    text = open(os.path.join(logdir, files[0])).read()
    assert "MOFSupplier ready" in text and "INFO" in text and "provider.cc" in text
    assert not any("MOFSupplier ready" in m for _, m in prov.logs)      # nothing went to the host's logger
    # the next role in this process logs through the host again
    prov2 = bridge.NativeRole(False, bridge.build_argv(port=0, buf_kb=64, provider=True, log_dir=logdir))
    try:
        assert any("MOFSupplier ready" in m for _, m in prov2.logs)
        prov2.command(bridge.form_cmd(bridge.EXIT, []))
    finally:
        prov2.close()


def test_provider_refuses_an_unknown_reader_backend(built):
    prov = bridge.NativeRole(False, bridge.build_argv(port=0, buf_kb=64, provider=True), conf={"mapred.uda.provider.reader": "mmap"})
    try:
        assert prov.failed and "unknown reader kind" in (prov.last_error or "") + " ".join(m for _, m in prov.logs)
    finally:
        prov.close()


def _prebuilt_mofs(tmp, per_map_parts):
    """MOFs from explicit record lists: per_map_parts[m][r] = sorted records of map m for reducer r."""
    mofs, truth = {}, [[] for _ in per_map_parts[0]]
    for m, parts in enumerate(per_map_parts):
        data, index = ifile.write_mof(parts)
        path = os.path.join(tmp, f"map_{m}.out")
        with open(path, "wb") as f:
            f.write(data)
        mofs[f"attempt_m_{m:06d}_0"] = (path, ifile.read_index(index))
        for r, recs in enumerate(parts):
            truth[r].append(recs)
    return mofs, truth


@pytest.mark.parametrize("approach", [1, 2])
def test_plumbing_record_larger_than_a_block(built, tmp_path, approach):
    """A record above 1 MiB cannot share a block: it is delivered alone, in a block of its own size, between ordinary
    blocks of whole records (reference: write_kv_to_stream refuses nothing, the Java side grows its buffer)."""
    big = bytes(range(256)) * 5000                                    # 1.28 MB value
    parts = [[[(b"a" * 10, b"x"), (b"m" * 10, big), (b"z" * 10, b"y")]],
             [[(b"b" * 10, b"p" * 100), (b"n" * 10, b"q" * 100)]],
             [[(b"c" * 10, b""), (b"m" * 10 + b"!", b"after the big one")]]]
    streams, truth, consumers, _ = _run_job(str(tmp_path), 3, 1, 0, approach=approach, prebuilt=_prebuilt_mofs(str(tmp_path), parts),
                                            max_block=len(big) + 64)
    assert parse_stream(streams[0]) == list(kway_merge(truth[0]))
    sizes = [len(b) for b in consumers[0].blocks]
    assert max(sizes) > (1 << 20) and sum(1 for x in sizes if x > (1 << 20)) == 1


def test_unsupported_merge_approach_fails_over(built, tmp_path):
    """An approach the native side does not know ends in one failureInUda up-call (-> vanilla shuffle), not in a hang."""
    parts = [[[(b"k" * 10, b"v")]]]
    _run_job(str(tmp_path), 1, 1, 0, approach=7, prebuilt=_prebuilt_mofs(str(tmp_path), parts), expect_failure=True)


def test_truncated_segment_is_a_failure_not_garbage(built, tmp_path):
    """The index promises more bytes than the segment's records fill (EOF marker missing): the consumer must not emit a
    silently shortened stream (reference: BaseSegment::nextKVInternal bounds checks, src/Merger/StreamRW.cc:334-404)."""
    recs = [(bytes([65 + i]) * 10, b"v" * 50) for i in range(20)]
    body = ifile.encode_segment(recs, with_crc=False)[:-2]           # records without the EOF marker
    cut = body[:len(body) - 30]                                      # ... and the last record cut short
    path = os.path.join(str(tmp_path), "map_0.out")
    with open(path, "wb") as f:
        f.write(cut + b"\0\0\0\0")
    mofs = {"attempt_m_000000_0": (path, [ifile.IndexRecord(0, len(cut), len(cut) + 4)])}
    _run_job(str(tmp_path), 1, 1, 0, prebuilt=(mofs, [[recs]]), expect_failure=True)


def test_plumbing_small_buffers_straddling_records(built, tmp_path):
    # 4 KiB chunks with values up to 200 B: many records straddle chunk boundaries
    streams, truth, _, _ = _run_job(str(tmp_path), 7, 1, 900, buf_kb=4, seed=5)
    assert parse_stream(streams[0]) == list(kway_merge(truth[0]))


def test_plumbing_hybrid_spills_lpqs(built, tmp_path):
    streams, truth, consumers, _ = _run_job(str(tmp_path), 9, 1, 300, approach=2, seed=9)
    assert parse_stream(streams[0]) == list(kway_merge(truth[0]))
    assert not [f for f in os.listdir(tmp_path) if ".lpq-" in f]       # spill files are deleted once consumed


def test_plumbing_text_keys(built, tmp_path):
    rng = random.Random(3)
    # Text keys: ordering must ignore the VInt length prefix
    def text(n):
        w = bytes(rng.choice(b"abc") for _ in range(n))
        return vint.encode_vlong(len(w)) + w
    tmp = str(tmp_path)
    runs = [sorted(((text(rng.randint(0, 5)), b"v") for _ in range(50)), key=lambda kv: kv[0][1:]) for _ in range(3)]
    mofs = {}
    for m, recs in enumerate(runs):
        data, index = ifile.write_mof([recs])
        p = os.path.join(tmp, f"t{m}.out")
        open(p, "wb").write(data)
        mofs[f"m{m}"] = (p, ifile.read_index(index))
    prov = bridge.NativeRole(False, bridge.build_argv(port=0, provider=True),
                             get_path=lambda j, m, r: (mofs[m][1][r].start_offset, mofs[m][1][r].raw_length, mofs[m][1][r].part_length, mofs[m][0]))
    cons = bridge.NativeRole(True, bridge.build_argv(port=prov.port))
    try:
        init = ["3", "job_1", "attempt_r_0", "0", str(1 << 20), "4096", "org.apache.hadoop.io.Text", "null", "262144", str(64 << 20), "0"]
        assert cons.command(bridge.form_cmd(bridge.INIT, init)) == 0
        for m in mofs:
            cons.command(bridge.form_cmd(bridge.FETCH, ["127.0.0.1", "job_1", m, "0"]))
        t0 = time.time()
        while not cons.stream().endswith(b"\xff\xff") and time.time() - t0 < 30:
            time.sleep(0.02)
        got = parse_stream(cons.stream())
        assert [k[1:] for k, _ in got] == sorted(k[1:] for run in runs for k, _ in run)
    finally:
        cons.reduce_exit(); cons.close(); prov.close()


def test_plumbing_snappy_compressed_segments(built, tmp_path, L):
    def store(b):
        out = _buf(len(b) + 64)
        n = L.uda_rt_snappy_store(bytes(b), len(b), out, len(out))
        return out.raw[:n]
    streams, truth, _, _ = _run_job(str(tmp_path), 3, 1, 200, compress=store, codec="org.apache.hadoop.io.compress.SnappyCodec", seed=4)
    assert parse_stream(streams[0]) == list(kway_merge(truth[0]))


def test_unsupported_key_class_fails_init_for_fallback(built):
    cons = bridge.NativeRole(True, bridge.build_argv(port=1))
    try:
        init = ["1", "job_1", "attempt_r_0", "0", str(1 << 20), "4096", "org.apache.hadoop.io.DoubleWritable", "null", "0", str(1 << 20), "0"]
        assert cons.command(bridge.form_cmd(bridge.INIT, init)) == -1       # exception in init => Java falls back
        assert cons.failed and "unsupported type" in cons.last_error
        assert cons.command(bridge.form_cmd(bridge.FETCH, ["h", "j", "m", "0"])) == 0   # ignored after the failure
    finally:
        cons.close()


def test_missing_provider_raises_failure_upcall(built):
    cons = bridge.NativeRole(True, bridge.build_argv(port=1))   # nothing listens on port 1
    try:
        init = ["1", "job_1", "attempt_r_0", "0", str(1 << 20), "4096", "org.apache.hadoop.io.LongWritable", "null", "0", str(64 << 20), "0"]
        assert cons.command(bridge.form_cmd(bridge.INIT, init)) == 0
        cons.command(bridge.form_cmd(bridge.FETCH, ["127.0.0.1", "job_1", "m0", "0"]))
        t0 = time.time()
        while cons.failures == 0 and time.time() - t0 < 20:
            time.sleep(0.02)
        assert cons.failures == 1                                # failureInUda: native-thread failure => fallback
    finally:
        cons.close()


def test_log_line_contract(built, tmp_path):
    prov = bridge.NativeRole(False, bridge.build_argv(port=0, provider=True), log_level=4)
    cons = bridge.NativeRole(True, bridge.build_argv(port=prov.port), log_level=4)
    try:
        ver = bridge.lib().uda_rt_version().decode()
        assert any(m.startswith(f"The version is {ver}") for _, m in prov.logs)
        assert any(m.startswith(f"UDA version is {ver}") for _, m in cons.logs)
        init = ["1", "job_1", "attempt_r_0", "0", str(1 << 20), "4096", "org.apache.hadoop.io.LongWritable", "null", "0", str(64 << 20), "0"]
        cons.command(bridge.form_cmd(bridge.INIT, init))
        assert any("After RDMA buffers registration: buffer1 = 1048576 bytes , buffer2 = 1048576 bytes , buffers count = 11" in m
                   for _, m in cons.logs)
    finally:
        cons.close(); prov.close()


def test_reader_bench_selftest(built):
    """N19: the native reader benchmark (counterpart of src/tests/AIOHandler_test.cc) agrees between back-ends."""
    import subprocess
    p = subprocess.run([str(built["reader_bench"])], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    assert p.stdout.count("BW:") == 3 and "checksums agree" in p.stdout


def test_cpu_merge_bench_native_plumbing_over_tcp(built):
    """BASELINE config 1 without Python in the loop: a provider role serving file.out files over TCP loop-back and two
    consumer roles merging at the same time, all inside tests/native/cpu_merge_bench."""
    import subprocess
    from pathlib import Path
    rt = os.environ.get("UDA_B200_RUNTIME_LIB")
    exe = (Path(rt).parent if rt else Path(__file__).resolve().parent.parent / "build" / "bin") / "cpu_merge_bench"
    if not exe.exists():
        pytest.skip(f"{exe} not built")
    p = subprocess.run([str(exe), "libuda.so", "--tcp", "--maps", "4", "--reducers", "2", "--records", "400000", "--steps", "2", "--warmup", "1",
                        "--check"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    kv = dict(x.split("=", 1) for x in next(l for l in p.stdout.splitlines() if l.startswith("OURS ")).split()[1:])
    assert kv["steps"] == "2" and kv["check"] == "ok" and int(kv["bytes"]) == 400000 * 102 + 2 * 2


@pytest.mark.parametrize("approach,maps", [(1, 8), (2, 16)])
def test_cpu_merge_bench_memory_fed_reducer(built, approach, maps):
    """The host runtime's reducer fed from memory through the fetcher factory hook (tests/native/cpu_merge_bench.cc, the
    counterpart of baseline/ref_harness.cc): the merged stream is complete, ordered, and every record is delivered."""
    import subprocess
    from pathlib import Path
    rt = os.environ.get("UDA_B200_RUNTIME_LIB")
    exe = (Path(rt).parent if rt else Path(__file__).resolve().parent.parent / "build" / "bin") / "cpu_merge_bench"
    if not exe.exists():
        pytest.skip(f"{exe} not built")
    p = subprocess.run([str(exe), "libuda.so", "--maps", str(maps), "--records", "300000", "--steps", "2", "--warmup", "1", "--approach", str(approach),
                        "--check"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    line = next(l for l in p.stdout.splitlines() if l.startswith("OURS "))
    kv = dict(x.split("=", 1) for x in line.split()[1:])
    assert kv["steps"] == "2" and kv["check"] == "ok" and int(kv["bytes"]) == 300000 * 102 + 2
    assert all(float(x) > 0 for x in kv["seconds"].split(","))


def test_regression_driver_parses_matrix(tmp_path):
    """The single-node regression driver reads its matrix and produces a report without executing anything."""
    import json, subprocess, sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    p = subprocess.run([sys.executable, str(root / "scripts" / "regression" / "auto_tester.py"), "--out", str(tmp_path),
                        "--phases", "input,control,configure,analyze,view,exit"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    rep = json.loads((tmp_path / "report.json").read_text())
    names = [t["name"] for t in rep["tests"]]
    assert "terasort-8gpu" in names and "plumbing-online" in names and len(names) >= 10


def test_plumbing_device_merge_falls_back_without_a_gpu(built, tmp_path):
    """Merge approach 3 (device merge) on a box without a usable GPU: the consumer keeps the segments it fetched and
    merges them on the CPU; the stream is the same as the online merge's."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible: the device path is covered by tests/test_device_merge_gpu.py")
    streams, truth, consumers, _ = _run_job(str(tmp_path), 5, 2, 300, approach=3, seed=3)
    for r, s in enumerate(streams):
        want = list(kway_merge(truth[r]))
        assert parse_stream(s) == want and s == merged_stream(want)
    assert all(c.final_stats["device_merges"] == 0 for c in consumers)
    assert any("device merge not possible" in m for c in consumers for _, m in c.logs)


# ------------------------------------------------------------------------------------------ built-in LZO1X
def _lzo_dec(L, stream, n):
    out = _buf(max(n, 1))
    rc = L.uda_rt_lzo1x_decompress(stream, len(stream), out, n)
    # the device kernel's cooperative copy (csrc/kernels/lzo.cu), emulated lane by lane: same verdict, same bytes
    for lanes in (32, 3):
        out2 = _buf(max(n, 1))
        rc2 = L.uda_rt_lzo1x_decompress_lanes(stream, C.c_longlong(len(stream)), out2, C.c_longlong(n), lanes)
        assert rc2 == rc and (rc != 0 or out2.raw[:n] == out.raw[:n]), (lanes, rc, rc2)
    return rc, out.raw[:n]


def test_lzo1x_decoder_hand_made_streams(L):
    """Streams written by hand from the format description (not by this repo's compressor), one per instruction form."""
    end = bytes([0x11, 0, 0])
    # first-byte literal (17+3), then 001L LLLL match: len 2+7=9, distance 3, no trailing literals
    assert _lzo_dec(L, bytes([20]) + b"abc" + bytes([0x27, (3 - 1) << 2, 0]) + end, 12) == (0, b"abcabcabcabc")
    # first-byte literal of 5 (state 4), then 0000 DDSS H = 3-byte match at distance 2049 + D + 4H: not reachable with 5
    # bytes of history -> the decoder must refuse instead of reading before the buffer
    assert _lzo_dec(L, bytes([22]) + b"hello" + bytes([0x00, 0x00]) + end, 8)[0] == -1
    # 01LD DDSS H: len 3+L, distance = 8H + D + 1; "abcdef" + copy 4 bytes from distance 6 + 1 trailing literal 'X'
    m2 = bytes([0x40 | (1 << 5) | ((6 - 1) << 2) | 1, 0])
    assert _lzo_dec(L, bytes([17 + 6]) + b"abcdef" + m2 + b"X" + end, 11) == (0, b"abcdefabcdX")
    # 1LLD DDSS H: len 5+LL=8, distance 2 (overlapping copy = repeating pattern)
    m2b = bytes([0x80 | (3 << 5) | ((2 - 1) << 2), 0])
    assert _lzo_dec(L, bytes([17 + 2]) + b"xy" + m2b + end, 10) == (0, b"xyxyxyxyxy")
    # 0000 LLLL literal run in the middle (state 0): "ab" + match len 4 dist 2 + run of 3+1=4 literals "WXYZ"
    m3 = bytes([0x20 | (4 - 2), (2 - 1) << 2, 0])
    assert _lzo_dec(L, bytes([19]) + b"ab" + m3 + bytes([1]) + b"WXYZ" + end, 10) == (0, b"abababWXYZ")
    # state 1..3 then 0000 DDSS H = 2-byte match at distance 4H + D + 1: "abcd" + match(len 3, dist 4, 1 trailing 'Q') + 2-byte match dist 1
    m3s = bytes([0x20 | (3 - 2), ((4 - 1) << 2) | 1, 0])
    assert _lzo_dec(L, bytes([21]) + b"abcd" + m3s + b"Q" + bytes([0x00, 0x00]) + end, 10) == (0, b"abcdabcQQQ")
    # 0001 HLLL far match (distance 16384 + D): 16400 bytes of history, copy 5 from 16390 back
    hist = bytes(i % 251 for i in range(16400))
    t = 16400 - 18
    run = bytes([0]) + bytes(t // 255 if t % 255 else t // 255 - 1) + bytes([t % 255 or 255])
    far = bytes([0x10 | (5 - 2), ((16390 - 16384) << 2) & 0xFF, ((16390 - 16384) << 2) >> 8])
    rc, out = _lzo_dec(L, run + hist + far + end, 16405)
    assert rc == 0 and out == hist + hist[16400 - 16390:16400 - 16390 + 5]
    # truncated stream and wrong output length are errors
    good = bytes([20]) + b"abc" + bytes([0x27, 8, 0]) + end
    assert _lzo_dec(L, good[:-1], 12)[0] == -1 and _lzo_dec(L, good, 11)[0] == -1 and _lzo_dec(L, good, 13)[0] == -1


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 18, 19, 238, 239, 240, 1000, 70000, 300000])
def test_lzo1x_round_trip(L, n):
    L.uda_rt_lzo1x_compress.restype = C.c_longlong
    rng = random.Random(n)
    for kind in range(4):
        if kind == 0:
            data = bytes(rng.randrange(256) for _ in range(n))
        elif kind == 1:
            data = (b"abcabcabd" * (n // 9 + 1))[:n]
        elif kind == 2:
            data = bytes(rng.choice(b"ab") for _ in range(n))
        else:   # short literal runs between matches: every state value appears
            data = b"".join(b"0123456789ABCDEF" + bytes(rng.randrange(256) for _ in range(rng.randint(0, 5))) for _ in range(n // 16 + 1))[:n]
        cap = n + n // 16 + 64 + 3
        out = _buf(cap)
        m = L.uda_rt_lzo1x_compress(data, C.c_longlong(n), out, C.c_longlong(cap))
        assert m >= 3, (n, kind, m)
        rc, back = _lzo_dec(L, out.raw[:m], n)
        assert rc == 0 and back == data, (n, kind)
        if kind == 1 and n >= 1000:
            assert m < n // 4


@pytest.mark.parametrize("ahead", ["true", "false"])
def test_corrupt_compressed_block_fails_the_task(built, tmp_path, L, ahead, monkeypatch):
    """A compressed block that does not decode (here: its Snappy length prefix is wrong) fails the fetch of that map output
    and through it the task -- with the decompress thread (the error is parked in the source and rethrown on the merge
    thread) and with inline decoding alike."""
    L.uda_rt_snappy_compress.restype = C.c_longlong

    def comp(b):
        out = _buf(len(b) + len(b) // 6 + 64)
        n = L.uda_rt_snappy_compress(bytes(b), C.c_longlong(len(b)), out, C.c_longlong(len(out)))
        return out.raw[:n]

    mofs, truth = _make_mofs(str(tmp_path), 3, 1, 400, seed=8, compress=comp)
    path, idx = mofs["attempt_m_000001_0"]
    data = bytearray(open(path, "rb").read())
    second_block = idx[0].start_offset + 8 + struct.unpack(">I", data[idx[0].start_offset + 4:idx[0].start_offset + 8])[0]
    data[second_block + 8] ^= 0x7F                      # first byte of the second block's Snappy stream: the varint length
    open(path, "wb").write(bytes(data))
    real = bridge.NativeRole.__init__

    def init_with_conf(self, is_net_merger, argv, *a, **kw):
        if is_net_merger:
            kw["conf"] = {**(kw.get("conf") or {}), "mapred.uda.decompress.ahead": ahead}
        real(self, is_net_merger, argv, *a, **kw)

    monkeypatch.setattr(bridge.NativeRole, "__init__", init_with_conf)
    _run_job(str(tmp_path), 3, 1, 0, prebuilt=(mofs, truth), codec="org.apache.hadoop.io.compress.SnappyCodec", expect_failure=True)


def test_plumbing_lzo_compressed_segments(built, tmp_path, L):
    """LZO map outputs through provider -> consumer without liblzo2 on the box: the built-in LZO1X decoder serves."""
    L.uda_rt_lzo1x_compress.restype = C.c_longlong

    def comp(b):
        out = _buf(len(b) + len(b) // 16 + 80)
        n = L.uda_rt_lzo1x_compress(bytes(b), C.c_longlong(len(b)), out, C.c_longlong(len(out)))
        assert n > 0
        return out.raw[:n]

    streams, truth, _, _ = _run_job(str(tmp_path), 3, 2, 250, compress=comp, codec="com.hadoop.compression.lzo.LzoCodec", seed=6)
    for r, s in enumerate(streams):
        assert parse_stream(s) == list(kway_merge(truth[r]))


@pytest.mark.parametrize("approach", [1, 2])
def test_example_fake_hadoop_plumbing(built, approach):
    import subprocess, sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    p = subprocess.run([sys.executable, str(root / "examples" / "fake_hadoop_plumbing.py"), "--approach", str(approach)],
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip().endswith("OK"), p.stdout + p.stderr


# ------------------------------------------------------------------------------------------ hostile fetch requests
def test_provider_rejects_malformed_fetch_requests(built, tmp_path):
    """The request text comes off a socket: a negative chunk size / offset or an offset past the partition must be
    answered with a rejecting ack (nothing read, nothing sent), not turned into a wrapped read length; an unknown map
    is rejected too instead of leaving the reducer waiting (the reference never answers such a request)."""
    import socket

    mofs, _truth = _make_mofs(str(tmp_path), 1, 1, 50, seed=3)
    (mapid, (path, idx)), = mofs.items()

    def get_path(job, m, reduce):
        if m != mapid:
            return None
        r = idx[reduce]
        return r.start_offset, r.raw_length, r.part_length, path

    prov = bridge.NativeRole(False, bridge.build_argv(port=0, buf_kb=8, provider=True), get_path=get_path)
    try:
        assert not prov.failed and prov.port > 0
        part_len = idx[0].part_length

        def ask(text):
            with socket.create_connection(("127.0.0.1", prov.port), timeout=10) as s:
                s.sendall(struct.pack("<BBHIQQ", 0x02, 0, 0, len(text), 0, 0) + text.encode())
                hdr = b""
                while len(hdr) < 24:
                    chunk = s.recv(24 - len(hdr))
                    assert chunk, "provider closed the connection without an ack"
                    hdr += chunk
                _t, _c, _p, tot, _cookie, plen = struct.unpack("<BBHIQQ", hdr)
                body = b""
                while len(body) < tot + plen:
                    chunk = s.recv(tot + plen - len(body))
                    assert chunk
                    body += chunk
                return body[:tot].decode(), plen

        # job:map:map_offset:reduce:addr:cookie:chunk_size:file_offset:path:raw:part
        good, plen = ask(f"job_1:{mapid}:0:0:0:7:4096:0::0:0")
        assert "UDA_FETCH_REJECTED" not in good and plen == min(4096, part_len)
        for bad in (f"job_1:{mapid}:0:0:0:7:-5:0::0:0",                       # negative chunk size
                    f"job_1:{mapid}:-4096:0:0:7:4096:0::0:0",                 # negative offset
                    f"job_1:{mapid}:{part_len + 1}:0:0:7:4096:0::0:0",        # offset past the partition
                    f"job_1:{mapid}:0:0:0:7:4096:-8:{path}:10:10",            # negative file offset in a follow-up request
                    "job_1:attempt_unknown:0:0:0:7:4096:0::0:0"):             # no such map output
            text, plen = ask(bad)
            assert "UDA_FETCH_REJECTED" in text and plen == 0, (bad, text)
        # an offset exactly at the end of the partition is legal: empty chunk, normal ack
        text, plen = ask(f"job_1:{mapid}:{part_len}:0:0:7:4096:0::0:0")
        assert "UDA_FETCH_REJECTED" not in text and plen == 0
        # follow-up requests carry the index record themselves: a file that does not exist, a file shorter than the record
        # claims (short read), a path above the 600-character limit (answered with the marker the reducer logs about)
        text, plen = ask(f"job_1:{mapid}:0:0:0:7:4096:0:{path}.missing:{part_len}:{part_len}")
        assert "UDA_FETCH_REJECTED" in text and plen == 0
        text, plen = ask(f"job_1:{mapid}:{(part_len + 4095) // 4096 * 4096 + 8192}:0:0:7:4096:0:{path}:{part_len + 100000}:{part_len + 100000}")
        assert "UDA_FETCH_REJECTED" in text and plen == 0                         # a chunk past the end of the file: short read
        long_path = "/" + "d" * 700
        text, plen = ask(f"job_1:{mapid}:0:0:0:7:4096:0:{long_path}:10:10")
        assert plen == 0 and "UDA_FETCH_REJECTED" not in text and long_path not in text
        good, plen = ask(f"job_1:{mapid}:0:0:0:7:4096:0::0:0")              # the provider is still serving
        assert plen == min(4096, part_len)
        assert any("short" in m or "failed" in m for lvl, m in prov.logs if lvl <= 2)
    finally:
        prov.command(bridge.form_cmd(bridge.EXIT, []))
        prov.close()
