"""TeraValidate-equivalent checker (the reference's terasort oracle is Hadoop's
``teravalidate`` with empty error output and input bytes == output bytes,
scripts/regression/mr-dstatExcel.sh:246-340)."""
from __future__ import annotations

import zlib
from typing import Callable, Iterable, Sequence, Tuple

KV = Tuple[bytes, bytes]


def record_checksum(records: Iterable[KV]) -> Tuple[int, int, int]:
    """Order-independent multiset fingerprint: ``(count, total_bytes, sum of CRC32(key+val) mod 2^64)``."""
    n = tot = acc = 0
    for k, v in records:
        n += 1
        tot += len(k) + len(v)
        acc = (acc + zlib.crc32(k + v)) & 0xFFFFFFFFFFFFFFFF
    return n, tot, acc


def teravalidate(
    outputs: Sequence[Sequence[KV]],
    inputs_checksum: Tuple[int, int, int] | None = None,
    sort_key: Callable[[bytes], bytes] = lambda k: k,
) -> None:
    """Raise ``AssertionError`` unless every partition is sorted, partitions are globally
    ordered (last key of r <= first key of r+1) and the record multiset matches."""
    prev = None
    for r, part in enumerate(outputs):
        for i, (k, _) in enumerate(part):
            sk = sort_key(k)
            if prev is not None and sk < prev:
                raise AssertionError(f"misordered record {i} in partition {r}")
            prev = sk
    if inputs_checksum is not None:
        got = record_checksum(kv for part in outputs for kv in part)
        if got != inputs_checksum:
            raise AssertionError(f"record multiset mismatch: in={inputs_checksum} out={got}")
