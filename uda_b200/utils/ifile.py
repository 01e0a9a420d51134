"""Hadoop IFile segment + spill index (``file.out`` / ``file.out.index``) reader/writer.

CPU oracle for the device codec and the host-side MOF file format.

Layout (the reference never parses the index itself — it receives
``{startOffset, rawLength, partLength}`` from Hadoop's IndexCache
(plugins/shared/org/apache/hadoop/mapred/IndexCacheBridge.java:33-38) and reads
``partLength`` bytes at ``startOffset`` (src/MOFServer/IndexInfo.cc:308-310)):

* segment  = ``[VInt klen][VInt vlen][key][val]`` ... ``VInt(-1) VInt(-1)`` ``[u32 BE CRC32 of all preceding bytes]``
* ``rawLength``  = bytes up to and including the EOF marker (decompressed length)
* ``partLength`` = on-disk length (= rawLength + 4 when uncompressed)
* index    = per partition three big-endian int64 ``(startOffset, rawLength, partLength)``
  followed by one big-endian int64 holding the CRC32 of the index body.
* compressed segments use Hadoop block framing ``[u32 BE rawLen][u32 BE compLen][bytes]``
  (src/Merger/LzoDecompressor.cc:151-167, src/Merger/SnappyDecompressor.cc:89-105).
"""
from __future__ import annotations

import struct
import zlib
from dataclasses import dataclass
from typing import Iterable, Iterator, List, Sequence, Tuple

from .vint import EOF_MARKER, decode_vlong, encode_vlong

__all__ = [
    "IndexRecord",
    "encode_record",
    "encode_segment",
    "iter_segment",
    "decode_segment",
    "write_mof",
    "read_index",
    "write_index",
    "read_partition",
    "CRC_LEN",
]

CRC_LEN = 4
INDEX_RECORD_LEN = 24


@dataclass(frozen=True)
class IndexRecord:
    """One ``file.out.index`` entry (reference: src/MOFServer/IndexInfo.h:98-104)."""

    start_offset: int
    raw_length: int
    part_length: int


def encode_record(key: bytes, val: bytes) -> bytes:
    return encode_vlong(len(key)) + encode_vlong(len(val)) + key + val


def encode_segment(records: Iterable[Tuple[bytes, bytes]], with_crc: bool = True) -> bytes:
    """Serialise ``records`` as one IFile segment (EOF marker + CRC included)."""
    body = b"".join(encode_record(k, v) for k, v in records) + EOF_MARKER
    if with_crc:
        body += struct.pack(">I", zlib.crc32(body) & 0xFFFFFFFF)
    return body


def iter_segment(buf, pos: int = 0, end: int | None = None) -> Iterator[Tuple[bytes, bytes]]:
    """Yield ``(key, val)`` until the EOF marker (mirrors BaseSegment::nextKVInternal,
    src/Merger/StreamRW.cc:334-404: two VInts, EOF = (-1,-1), bounds check)."""
    end = len(buf) if end is None else end
    while pos < end:
        klen, n1 = decode_vlong(buf, pos)
        vlen, n2 = decode_vlong(buf, pos + n1)
        if klen == -1 and vlen == -1:
            return
        if klen < 0 or vlen < 0:
            raise ValueError(f"corrupt IFile record header at {pos}: klen={klen} vlen={vlen}")
        pos += n1 + n2
        if pos + klen + vlen > end:
            raise ValueError("record overruns segment")
        yield bytes(buf[pos : pos + klen]), bytes(buf[pos + klen : pos + klen + vlen])
        pos += klen + vlen
    raise ValueError("segment ended without EOF marker")


def decode_segment(buf, verify_crc: bool = False) -> List[Tuple[bytes, bytes]]:
    if verify_crc:
        body, crc = bytes(buf[:-CRC_LEN]), struct.unpack(">I", bytes(buf[-CRC_LEN:]))[0]
        if zlib.crc32(body) & 0xFFFFFFFF != crc:
            raise ValueError("IFile CRC mismatch")
    return list(iter_segment(buf))


def write_index(records: Sequence[IndexRecord]) -> bytes:
    body = b"".join(struct.pack(">qqq", r.start_offset, r.raw_length, r.part_length) for r in records)
    return body + struct.pack(">q", zlib.crc32(body) & 0xFFFFFFFF)


def read_index(buf: bytes, verify_crc: bool = True) -> List[IndexRecord]:
    if (len(buf) - 8) % INDEX_RECORD_LEN:
        raise ValueError("index length is not 24*P + 8")
    body = buf[:-8]
    if verify_crc and (zlib.crc32(body) & 0xFFFFFFFF) != struct.unpack(">q", buf[-8:])[0]:
        raise ValueError("index CRC mismatch")
    return [IndexRecord(*struct.unpack_from(">qqq", body, i)) for i in range(0, len(body), INDEX_RECORD_LEN)]


def write_mof(partitions: Sequence[Iterable[Tuple[bytes, bytes]]]) -> Tuple[bytes, bytes]:
    """Build ``(file.out bytes, file.out.index bytes)`` from per-partition record lists."""
    out = bytearray()
    idx: List[IndexRecord] = []
    for recs in partitions:
        seg = encode_segment(recs)
        idx.append(IndexRecord(len(out), len(seg) - CRC_LEN, len(seg)))
        out += seg
    return bytes(out), write_index(idx)


def write_mof_compressed(partitions: Sequence[Iterable[Tuple[bytes, bytes]]], frame) -> Tuple[bytes, bytes]:
    """Like :func:`write_mof` for a job with compressed map outputs. ``frame(plain) -> bytes`` turns a plain segment
    body (records + EOF marker) into codec blocks ``[u32 BE rawLen][u32 BE compLen][bytes]...`` (e.g.
    ``uda_b200.ops.compress.frame_compress_segment``). Hadoop's IFile.Writer puts the 4-byte checksum after the
    compressed stream: ``rawLength`` = plain bytes incl. the EOF marker, ``partLength`` = compressed bytes + 4."""
    out = bytearray()
    idx: List[IndexRecord] = []
    for recs in partitions:
        plain = encode_segment(recs, with_crc=False)
        seg = frame(plain)
        seg += struct.pack(">I", zlib.crc32(seg) & 0xFFFFFFFF)
        idx.append(IndexRecord(len(out), len(plain), len(seg)))
        out += seg
    return bytes(out), write_index(idx)


def read_partition(mof: bytes, rec: IndexRecord, verify_crc: bool = True) -> List[Tuple[bytes, bytes]]:
    seg = mof[rec.start_offset : rec.start_offset + rec.part_length]
    return decode_segment(seg, verify_crc=verify_crc)
