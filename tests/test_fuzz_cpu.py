"""Property / fuzz tests of the host runtime's parsers and decoders (hypothesis): arbitrary and mutated inputs must be
rejected or decoded, never crash. `scripts/ci/host_sanitizers.sh address` runs the same tests against the ASAN build,
where an out-of-bounds access is a failure rather than luck."""
import ctypes as C
import random

import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from uda_b200 import bridge

import os

# UDA_FUZZ_EXAMPLES=20000 turns the quick pass into a long session (run against the ASAN build: scripts/ci/host_sanitizers.sh)
FUZZ = settings(max_examples=int(os.environ.get("UDA_FUZZ_EXAMPLES", "300")), deadline=None,
                suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])


@pytest.fixture(scope="module")
def L(built):
    """A private handle of the runtime library: its functions carry no argtypes, so raw ctypes arrays can be passed."""
    import os
    from uda_b200 import _lib
    bridge.lib()   # makes sure the library loads the regular way first
    path = os.environ.get("UDA_B200_RUNTIME_LIB") or str(_lib.lib_path("libuda.so"))
    lib = C.CDLL(path)
    lib.uda_rt_snappy_compress.restype = C.c_longlong
    lib.uda_rt_lzo1x_compress.restype = C.c_longlong
    return lib


def _exact(data: bytes):
    """A heap buffer of exactly len(data) bytes, so that a read past the end is visible to a sanitizer."""
    return (C.c_ubyte * max(len(data), 1)).from_buffer_copy(data if data else b"\0"), len(data)


@FUZZ
@given(data=st.binary(max_size=600), out_len=st.integers(0, 2000))
def test_snappy_decoder_never_crashes(L, data, out_len):
    buf, n = _exact(data)
    out = (C.c_ubyte * max(out_len, 1))()
    assert L.uda_rt_snappy_uncompress(buf, C.c_longlong(n), out, C.c_longlong(out_len)) in (0, -1)


@FUZZ
@given(data=st.binary(max_size=600), out_len=st.integers(0, 2000))
def test_lzo_decoder_never_crashes(L, data, out_len):
    buf, n = _exact(data)
    out = (C.c_ubyte * max(out_len, 1))()
    rc = L.uda_rt_lzo1x_decompress(buf, C.c_longlong(n), out, C.c_longlong(out_len))
    assert rc in (0, -1)
    # the device kernel's copy discipline (lane-strided, self-overlapping matches as a repeating pattern) agrees
    out2 = (C.c_ubyte * max(out_len, 1))()
    assert L.uda_rt_lzo1x_decompress_lanes(buf, C.c_longlong(n), out2, C.c_longlong(out_len), 32) == rc
    assert rc != 0 or bytes(out2)[:out_len] == bytes(out)[:out_len]


@FUZZ
@given(seed=st.integers(0, 10**6), flips=st.integers(1, 6), codec=st.sampled_from(["snappy", "lzo"]))
def test_mutated_compressed_streams(L, seed, flips, codec):
    rng = random.Random(seed)
    plain = b"".join(rng.choice([b"shuffle", b"merge", b"segment", bytes([rng.randrange(256)])]) for _ in range(rng.randint(1, 300)))
    cap = len(plain) + len(plain) // 6 + 80
    comp = (C.c_ubyte * cap)()
    fn = L.uda_rt_snappy_compress if codec == "snappy" else L.uda_rt_lzo1x_compress
    m = fn(plain, C.c_longlong(len(plain)), comp, C.c_longlong(cap))
    assert m > 0
    stream = bytearray(bytes(comp)[:m])
    dec = L.uda_rt_snappy_uncompress if codec == "snappy" else L.uda_rt_lzo1x_decompress
    out = (C.c_ubyte * max(len(plain), 1))()
    buf, n = _exact(bytes(stream))
    assert dec(buf, C.c_longlong(n), out, C.c_longlong(len(plain))) == 0 and bytes(out)[:len(plain)] == plain
    for _ in range(flips):
        stream[rng.randrange(len(stream))] ^= 1 << rng.randrange(8)
    cut = rng.randint(0, len(stream))
    for s in (bytes(stream), bytes(stream[:cut])):
        buf, n = _exact(s)
        assert dec(buf, C.c_longlong(n), out, C.c_longlong(len(plain))) in (0, -1)


@FUZZ
@given(text=st.text(alphabet=st.characters(min_codepoint=32, max_codepoint=126), max_size=200))
def test_text_protocol_parsers_never_crash(L, text):
    out = C.create_string_buffer(4096)
    s = text.encode()
    L.uda_rt_parse_cmd(s, out, len(out))
    L.uda_rt_fetch_request_roundtrip(s, out, len(out))
    L.uda_rt_fetch_ack_roundtrip(s, out, len(out))


@FUZZ
@given(data=st.binary(min_size=0, max_size=12))
def test_vint_decoder_bounds(L, data):
    buf, n = _exact(data)
    got = C.c_longlong()
    r = L.uda_rt_vint_decode(buf, n, C.byref(got))
    assert 0 <= r <= max(n, 0) and r <= 9


# ---------------------------------------------------------------------------------------- the merge's record parser
def _fnv_records(recs):
    import struct
    h = 1469598103934665603
    for k, v in recs:
        for b in struct.pack("<II", len(k), len(v)) + k + v:
            h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def _parse_segment(L, data: bytes, chunk: int):
    L.uda_rt_selftest_parse_segment.restype = C.c_longlong
    buf, n = _exact(data)
    digest = C.c_ulonglong(0)
    rc = L.uda_rt_selftest_parse_segment(buf, C.c_longlong(n), C.c_int(chunk), C.byref(digest))
    return rc, digest.value


@FUZZ
@given(recs=st.lists(st.tuples(st.binary(max_size=40), st.binary(max_size=300)), max_size=30), chunk=st.integers(1, 400),
       big=st.booleans())
def test_segment_parser_is_independent_of_the_block_size(L, recs, chunk, big):
    """Records straddling block boundaries anywhere (inside a VInt, a key, a value) parse to the same records."""
    from uda_b200.utils import ifile
    if big and recs:
        recs = recs + [(recs[0][0], recs[0][1] * 40)]             # one record much larger than a block
    data = ifile.encode_segment(recs, with_crc=False)
    rc, digest = _parse_segment(L, data, chunk)
    assert rc == len(recs) and digest == _fnv_records(recs)
    rc2, digest2 = _parse_segment(L, data + b"\\0\\0\\0\\0", max(1, chunk // 2))     # trailing checksum bytes are ignored
    assert rc2 == len(recs) and digest2 == digest


@FUZZ
@given(data=st.binary(max_size=500), chunk=st.integers(1, 64))
def test_segment_parser_never_crashes_on_garbage(L, data, chunk):
    rc, _ = _parse_segment(L, data, chunk)
    assert rc >= -1


@FUZZ
@given(seed=st.integers(0, 10**6), flips=st.integers(1, 5), chunk=st.integers(1, 128))
def test_segment_parser_on_mutated_segments(L, seed, flips, chunk):
    from uda_b200.utils import ifile
    rng = random.Random(seed)
    recs = [(bytes(rng.randrange(256) for _ in range(rng.randint(0, 20))), bytes(rng.randrange(256) for _ in range(rng.randint(0, 200))))
            for _ in range(rng.randint(0, 20))]
    data = bytearray(ifile.encode_segment(recs, with_crc=False))
    for _ in range(flips):
        data[rng.randrange(len(data))] ^= 1 << rng.randrange(8)
    cut = rng.randint(0, len(data))
    for s in (bytes(data), bytes(data[:cut])):
        rc, _ = _parse_segment(L, s, chunk)
        assert rc >= -1


def test_segment_parser_rejects_lengths_no_record_can_have(L):
    """A header whose key or value length exceeds an int (Hadoop's own limit), chosen so that naive size arithmetic wraps:
    rejected, never turned into a negative length or a small wrapped total. A segment that stops without its EOF marker,
    even at a record boundary, is rejected too (silently delivering the shorter stream would lose records)."""
    from uda_b200.utils import ifile, vint
    good = ifile.encode_segment([(b"k1", b"v1"), (b"k2", b"v2")], with_crc=False)
    assert _parse_segment(L, good, 3)[0] == 2
    for kl, vl in ((2**63 - 1, 2**63 - 1), (2**62, 5), (5, 2**31), (2**32 + 2, 2), (2**64 - 2**32 - 0 if False else 2**40, 2**40)):
        bad = ifile.encode_record(b"k1", b"v1") + vint.encode_vlong(kl) + vint.encode_vlong(vl) + b"xxxxxxxx" + ifile.EOF_MARKER
        for chunk in (1, 7, 4096):
            assert _parse_segment(L, bad, chunk)[0] == -1, (kl, vl, chunk)
    assert _parse_segment(L, good[:-2], 4096)[0] == -1            # both records intact, EOF marker missing
    assert _parse_segment(L, good[:-1], 4096)[0] == -1            # half an EOF marker
    assert _parse_segment(L, b"", 16)[0] == 0                     # a partition of zero bytes is empty, not truncated
    assert _parse_segment(L, ifile.EOF_MARKER, 1)[0] == 0         # an empty partition is just the marker
