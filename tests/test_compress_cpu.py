"""Compressed-segment framing and the LZO1X block decoder's device copy discipline, checked on the host.

The device kernel (csrc/kernels/lzo.cu) runs the parser of csrc/runtime/lzo1x_parse.h with a warp-cooperative copier; the
native runtime exposes the same parser with that copier emulated lane by lane (uda_rt_lzo1x_decompress_lanes). These
tests decode the very segments tests/test_kernels_gpu.py feeds to the GPU, block by block, through that emulation.
"""
import ctypes as C
import random
import struct

import pytest

from uda_b200._lib import runtime
from uda_b200.ops.compress import frame_compress_segment


def _decode_framed(comp: bytes, lanes: int):
    L = runtime()
    pos, out, ok, blocks = 0, [], True, 0
    while pos + 8 <= len(comp):
        raw, clen = struct.unpack(">II", comp[pos:pos + 8])
        blk = comp[pos + 8:pos + 8 + clen]
        assert len(blk) == clen
        buf = C.create_string_buffer(max(raw, 1))
        rc = L.uda_rt_lzo1x_decompress_lanes(blk, C.c_longlong(clen), buf, C.c_longlong(raw), lanes)
        ok = ok and rc == 0
        out.append(buf.raw[:raw])
        pos += 8 + clen
        blocks += 1
    assert pos == len(comp)
    return ok, b"".join(out), blocks


def _plain(kind: str, n: int) -> bytes:
    rng = random.Random(n + 1)
    if kind == "random":
        return bytes(rng.randrange(256) for _ in range(n))
    if kind == "text":
        words = [bytes(rng.choice(b"etaoinshr") for _ in range(rng.randint(1, 9))) for _ in range(200)]
        return b" ".join(rng.choice(words) for _ in range(n // 5))[:n]
    if kind == "runs":
        return b"".join(bytes([rng.randrange(256)]) * rng.randint(1, 300) for _ in range(n // 100))[:n]
    if kind == "states":
        return b"".join(b"0123456789ABCDEF" + bytes(rng.randrange(256) for _ in range(rng.randint(0, 5))) for _ in range(n // 16 + 1))[:n]
    return b"abcde"[:n]


@pytest.mark.parametrize("kind,n,block", [("random", 50_000, 65536), ("text", 300_000, 262144), ("runs", 200_000, 4096),
                                          ("states", 100_000, 16384), ("tiny", 5, 65536), ("empty", 0, 65536)])
@pytest.mark.parametrize("lanes", [32, 1, 7])
def test_lzo_framed_segment_decodes_with_the_device_copy_discipline(kind, n, block, lanes):
    plain = _plain(kind, n)
    comp = frame_compress_segment(plain, block, codec="lzo")
    ok, back, blocks = _decode_framed(comp, lanes)
    assert ok and back == plain and blocks == -(-len(plain) // block)
    if kind in ("text", "runs", "states"):
        assert len(comp) < 0.8 * len(plain)


def test_lzo_corruptions_used_by_the_gpu_test_are_detected():
    plain = b"the quick brown fox jumps over the lazy dog " * 2000
    good = frame_compress_segment(plain, 16384, codec="lzo")
    first = struct.unpack(">I", good[4:8])[0]
    a = bytearray(good); a[8 + first - 3] = 0x19
    b = bytearray(good); b[8] = 17 + 4; b[9:13] = b"\x00\x00\x00\x00"
    for bad in (a, b):
        ok, back, _ = _decode_framed(bytes(bad), 32)
        assert not ok
        # only the damaged block is lost: the blocks after it decode to their plain bytes
        assert back[16384:] == plain[16384:]
    assert _decode_framed(good, 32)[:2] == (True, plain)


def test_frame_compress_segment_snappy_and_lzo_share_the_framing():
    plain = _plain("text", 100_000)
    for codec in ("snappy", "lzo"):
        comp = frame_compress_segment(plain, 32768, codec=codec)
        pos, raws = 0, []
        while pos < len(comp):
            raw, clen = struct.unpack(">II", comp[pos:pos + 8])
            raws.append(raw)
            pos += 8 + clen
        assert pos == len(comp) and sum(raws) == len(plain) and all(r == 32768 for r in raws[:-1])
    with pytest.raises(KeyError):
        frame_compress_segment(plain, 32768, codec="zstd")


@pytest.mark.parametrize("codec", ["snappy", "lzo"])
def test_write_mof_compressed_layout(codec):
    """Index of a compressed MOF: rawLength = plain bytes incl. the EOF marker, partLength = codec blocks + 4 checksum bytes;
    the blocks of every segment decode back to the plain segment (what DeviceMOF.from_host_bytes(codec=...) does on the GPU)."""
    import zlib
    from uda_b200.utils import ifile
    L = runtime()
    rng = random.Random(3)
    parts = [sorted((bytes(rng.randrange(97, 100) for _ in range(rng.randint(1, 12))), bytes(rng.randrange(256) for _ in range(rng.randint(0, 40))))
                    for _ in range(n)) for n in (300, 0, 1200)]
    file_out, index = ifile.write_mof_compressed(parts, lambda plain: frame_compress_segment(plain, 4096, codec=codec))
    recs = ifile.read_index(index)
    assert len(recs) == 3 and recs[0].start_offset == 0 and recs[2].start_offset + recs[2].part_length == len(file_out)
    for p, r in enumerate(recs):
        seg = file_out[r.start_offset:r.start_offset + r.part_length]
        assert struct.unpack(">I", seg[-4:])[0] == zlib.crc32(seg[:-4]) & 0xFFFFFFFF
        pos, plain = 0, b""
        while pos + 8 <= len(seg) - 4:
            raw, clen = struct.unpack(">II", seg[pos:pos + 8])
            buf = C.create_string_buffer(max(raw, 1))
            blk = seg[pos + 8:pos + 8 + clen]
            fn = L.uda_rt_lzo1x_decompress if codec == "lzo" else L.uda_rt_snappy_uncompress
            assert fn(blk, C.c_longlong(clen), buf, C.c_longlong(raw)) == 0
            plain += buf.raw[:raw]
            pos += 8 + clen
        assert pos == len(seg) - 4 and len(plain) == r.raw_length
        assert plain == ifile.encode_segment(parts[p], with_crc=False)
