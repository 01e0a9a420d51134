"""Key comparators keyed by the Java map-output key class name.

Reference behaviour (src/Merger/CompareFunc.cc:29-42 tables, :70-91 compare bodies,
:95-113 selection): ``Text`` skips its leading VInt length, ``BytesWritable`` /
``ImmutableBytesWritable`` skip a 4-byte length, the fixed-width Writables compare raw
bytes; all then ``memcmp`` with length tie-break. Anything else raises, which the Java
side turns into a fallback to the vanilla shuffle.
"""
from __future__ import annotations

from typing import Callable

from ..utils.vint import decode_vint_size

TEXT = "text"
BYTE = "byte"
BYTES = "bytes"

_KINDS = {
    "org.apache.hadoop.io.Text": TEXT,
    "org.apache.hadoop.io.BooleanWritable": BYTE,
    "org.apache.hadoop.io.ByteWritable": BYTE,
    "org.apache.hadoop.io.ShortWritable": BYTE,
    "org.apache.hadoop.io.IntWritable": BYTE,
    "org.apache.hadoop.io.LongWritable": BYTE,
    "org.apache.hadoop.io.BytesWritable": BYTES,
    "org.apache.hadoop.hbase.io.ImmutableBytesWritable": BYTES,
}
SUPPORTED_KEY_CLASSES = tuple(_KINDS)


class UnsupportedKeyClass(ValueError):
    """Raised for key classes the engine cannot order (⇒ fallback contract)."""


def key_class_kind(java_class: str) -> str:
    try:
        return _KINDS[java_class]
    except KeyError:
        raise UnsupportedKeyClass(f"using compare function for unsupported type: '{java_class}'") from None


def get_compare_key(java_class: str) -> Callable[[bytes], bytes]:
    """Return ``f(serialized_key) -> bytes`` such that Python's ``bytes`` ordering of
    ``f(k)`` equals the reference comparator's ordering (memcmp + shorter-first)."""
    kind = key_class_kind(java_class)
    if kind == TEXT:
        return lambda k: k[decode_vint_size(k[0]) :] if k else k
    if kind == BYTES:
        return lambda k: k[4:]
    return lambda k: k
