"""Hadoop zero-compressed VInt / VLong codec (CPU oracle + host-side use).

Wire format (reference behaviour: src/CommUtils/IOUtility.cc:167-196
encode, :287-333 decode, :367-382 size, :389-396 size-from-first-byte):

* ``-112 <= v <= 127``  -> one byte holding the value itself.
* otherwise a tag byte followed by the big-endian magnitude:
  tag ``-113..-120`` = positive, ``n = -(tag + 112)`` magnitude bytes;
  tag ``-121..-128`` = negative (magnitude is the one's complement), ``n = -(tag + 120)``.

The device codec (csrc/kernels/ifile_codec.cu) implements the same rules and is
tested against this module.
"""
from __future__ import annotations

from typing import Tuple

__all__ = [
    "encode_vlong",
    "decode_vlong",
    "vint_size",
    "decode_vint_size",
    "EOF_MARKER",
]


def encode_vlong(v: int) -> bytes:
    """Encode ``v`` (signed 64-bit range) as a Hadoop VLong."""
    if not -(1 << 63) <= v < (1 << 63):
        raise ValueError("VLong out of int64 range")
    if -112 <= v <= 127:
        return bytes([v & 0xFF])
    tag = -112
    if v < 0:
        v = ~v
        tag = -120
    n = (v.bit_length() + 7) // 8
    tag -= n
    return bytes([tag & 0xFF]) + v.to_bytes(n, "big")


def decode_vint_size(first_byte: int) -> int:
    """Total encoded size (1..9) given the first byte (signed or unsigned)."""
    b = first_byte - 256 if first_byte > 127 else first_byte
    if b >= -112:
        return 1
    if b < -120:
        return -119 - b
    return -111 - b


def decode_vlong(buf, pos: int = 0) -> Tuple[int, int]:
    """Decode a VLong at ``buf[pos:]``. Returns ``(value, bytes_consumed)``.

    Raises ``IndexError`` when the buffer ends inside the value (the reference
    returns ``false`` and rewinds; callers here treat that as "need more data").
    """
    b0 = buf[pos]
    b = b0 - 256 if b0 > 127 else b0
    if b >= -112:
        return b, 1
    negative = b < -120
    n = (-120 - b) if negative else (-112 - b)
    if pos + 1 + n > len(buf):
        raise IndexError("truncated VLong")
    mag = int.from_bytes(bytes(buf[pos + 1 : pos + 1 + n]), "big")
    if negative:
        mag = ~mag
        # interpret as signed 64-bit (one's complement of up-to-8-byte magnitude)
        mag &= (1 << 64) - 1
        if mag >= 1 << 63:
            mag -= 1 << 64
    return mag, n + 1


def vint_size(v: int) -> int:
    """Encoded size of ``v`` without encoding it."""
    if -112 <= v <= 127:
        return 1
    if v < 0:
        v = ~v
    return (v.bit_length() + 7) // 8 + 1


#: IFile end-of-segment marker: VInt(-1) VInt(-1)
EOF_MARKER = encode_vlong(-1) + encode_vlong(-1)
