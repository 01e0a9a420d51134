"""CPU tests: VInt, IFile segment/index, comparators, partitioners, merge oracle, teragen restatement."""
import random
import struct
import zlib

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from uda_b200.models import teragen
from uda_b200.oracle import (UnsupportedKeyClass, get_compare_key, hash_partition, kway_merge, make_splitters,
                             merged_stream, parse_stream, range_partition, record_checksum, teravalidate)
from uda_b200.oracle.partition import java_hash_bytes
from uda_b200.utils import ifile, vint


# ---------------------------------------------------------------------------------------- VInt
GOLDEN_VINT = {
    0: b"\x00", 1: b"\x01", 127: b"\x7f", -1: b"\xff", -112: b"\x90",
    128: b"\x8f\x80", 255: b"\x8f\xff", 256: b"\x8e\x01\x00", 65535: b"\x8e\xff\xff", 65536: b"\x8d\x01\x00\x00",
    -113: b"\x87\x70", -129: b"\x87\x80", -257: b"\x86\x01\x00",
    (1 << 31) - 1: b"\x8c\x7f\xff\xff\xff", (1 << 63) - 1: b"\x88\x7f" + b"\xff" * 7, -(1 << 63): b"\x80\x7f" + b"\xff" * 7,
}


@pytest.mark.parametrize("v,enc", sorted(GOLDEN_VINT.items()))
def test_vint_golden(v, enc):
    assert vint.encode_vlong(v) == enc
    assert vint.decode_vlong(enc) == (v, len(enc))
    assert vint.vint_size(v) == len(enc)
    assert vint.decode_vint_size(enc[0]) == len(enc)


@given(st.integers(min_value=-(1 << 63), max_value=(1 << 63) - 1))
@settings(max_examples=300, deadline=None)
def test_vint_roundtrip(v):
    e = vint.encode_vlong(v)
    assert vint.decode_vlong(e) == (v, len(e))
    assert vint.vint_size(v) == len(e) == vint.decode_vint_size(e[0])


def test_vint_truncated():
    with pytest.raises(IndexError):
        vint.decode_vlong(b"\x8e\x01")


def test_eof_marker():
    assert vint.EOF_MARKER == b"\xff\xff"


# ---------------------------------------------------------------------------------------- IFile
def _rand_records(rng, n, kmax=40, vmax=300):
    return [(bytes(rng.getrandbits(8) for _ in range(rng.randint(0, kmax))), bytes(rng.getrandbits(8) for _ in range(rng.randint(0, vmax))))
            for _ in range(n)]


def test_segment_roundtrip_and_crc():
    rng = random.Random(1)
    recs = _rand_records(rng, 200)
    seg = ifile.encode_segment(recs)
    assert seg[-6:-4] == b"\xff\xff"
    assert struct.unpack(">I", seg[-4:])[0] == zlib.crc32(seg[:-4]) & 0xFFFFFFFF
    assert ifile.decode_segment(seg, verify_crc=True) == recs
    bad = bytearray(seg)
    bad[3] ^= 1
    with pytest.raises(ValueError):
        ifile.decode_segment(bytes(bad), verify_crc=True)


def test_mof_index_layout():
    rng = random.Random(2)
    parts = [_rand_records(rng, n) for n in (5, 0, 17)]
    mof, idx = ifile.write_mof(parts)
    assert len(idx) == 3 * 24 + 8
    recs = ifile.read_index(idx)
    assert recs[0].start_offset == 0
    for p, r in enumerate(recs):
        assert r.part_length == r.raw_length + 4
        assert ifile.read_partition(mof, r) == parts[p]
    assert recs[1].raw_length == 2  # empty partition = EOF marker only
    assert recs[2].start_offset == recs[1].start_offset + recs[1].part_length
    with pytest.raises(ValueError):
        ifile.read_index(idx[:-1] + bytes([idx[-1] ^ 1]))


def test_terasort_record_is_102_bytes():
    assert len(ifile.encode_record(b"k" * 10, b"v" * 90)) == 102
    assert ifile.encode_record(b"k" * 10, b"v" * 90)[:2] == b"\x0a\x5a"


# ---------------------------------------------------------------------------------------- comparators
def test_compare_keys():
    text = get_compare_key("org.apache.hadoop.io.Text")
    # Text key = VInt(len) + bytes ; ordering ignores the length prefix
    a, b = vint.encode_vlong(3) + b"abc", vint.encode_vlong(2) + b"ab"
    assert text(b) < text(a)
    long_key = vint.encode_vlong(200) + b"x" * 200
    assert text(long_key) == b"x" * 200
    byt = get_compare_key("org.apache.hadoop.io.BytesWritable")
    assert byt(b"\x00\x00\x00\x02zz") == b"zz"
    raw = get_compare_key("org.apache.hadoop.io.LongWritable")
    assert raw(b"\x00\x01") == b"\x00\x01"
    with pytest.raises(UnsupportedKeyClass):
        get_compare_key("org.apache.hadoop.io.DoubleWritable")


# ---------------------------------------------------------------------------------------- partitioners
def test_partitioners():
    sp = make_splitters(4)
    assert len(sp) == 3 and sp[0] == b"\x40" + b"\x00" * 9
    assert range_partition(b"\x00" * 10, sp) == 0
    assert range_partition(sp[0], sp) == 1          # key == splitter goes right
    assert range_partition(b"\xff" * 10, sp) == 3
    assert java_hash_bytes(b"") == 1
    assert java_hash_bytes(b"a") == 31 + 97
    assert java_hash_bytes(b"\xff") == (31 - 1) & 0xFFFFFFFF
    assert 0 <= hash_partition(b"hello", 7) < 7


# ---------------------------------------------------------------------------------------- merge oracle
def test_kway_merge_and_stream():
    rng = random.Random(3)
    runs = []
    for _ in range(7):
        r = sorted(_rand_records(rng, rng.randint(0, 50), kmax=3), key=lambda kv: kv[0])
        runs.append(r)
    merged = list(kway_merge(runs))
    allr = [kv for r in runs for kv in r]
    assert sorted(k for k, _ in merged) == [k for k, _ in merged]
    assert record_checksum(merged) == record_checksum(allr)
    # stability by run index among equal keys
    pos = {}
    for ri, r in enumerate(runs):
        for kv in r:
            pos.setdefault(kv[0], []).append(ri)
    seen = {}
    for k, v in merged:
        seen.setdefault(k, []).append(v)
    stream = merged_stream(merged)
    assert stream.endswith(b"\xff\xff")
    assert parse_stream(stream) == merged
    teravalidate([merged], record_checksum(allr))
    with pytest.raises(AssertionError):
        teravalidate([list(reversed(merged))] if len(merged) > 1 and merged[0][0] != merged[-1][0] else [[(b"b", b""), (b"a", b"")]])


# ---------------------------------------------------------------------------------------- teragen
def test_teragen_numpy_properties():
    a = teragen.teragen_framed(1000, seed=7)
    assert a.shape == (1000, 102) and (a[:, 0] == 10).all() and (a[:, 1] == 90).all()
    b = teragen.teragen_framed(500, seed=7, first_row=500)
    assert (a[500:] == b).all()                       # deterministic in (seed, row)
    rows = np.array([int.from_bytes(bytes(r[12:20]), "big") for r in a])
    assert (rows == np.arange(1000)).all()
    d = teragen.teragen_raw(4096, seed=7, key_bits=4)
    assert len({bytes(r[:10]) for r in d}) <= 16      # reduced entropy keys
    fp1 = teragen.fingerprint_sum(teragen.teragen_raw(100, 9))
    perm = np.random.RandomState(0).permutation(100)
    assert teragen.fingerprint_sum(teragen.teragen_raw(100, 9)[perm]) == fp1
