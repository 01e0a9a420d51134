"""k-way merge oracle and merged-stream framing.

Reference behaviour: the heap merge (src/Merger/MergeQueue.h:126-270,299-321) pops the
segment with the smallest current key; ``write_kv_to_stream``
(src/Merger/StreamRW.cc:151-225) serialises ``VInt(klen) VInt(vlen) key val`` and ends the
whole stream with ``VInt(-1) VInt(-1)``. Order among equal keys is heap dependent in the
reference; this oracle (and the GPU merge) are *stable by run index*, which is one of the
orders the reference can legally produce.
"""
from __future__ import annotations

import heapq
from typing import Callable, Iterable, Iterator, List, Sequence, Tuple

from ..utils.ifile import encode_record
from ..utils.vint import EOF_MARKER, decode_vlong

KV = Tuple[bytes, bytes]


def kway_merge(runs: Sequence[Iterable[KV]], sort_key: Callable[[bytes], bytes] = lambda k: k) -> Iterator[KV]:
    """Stable k-way merge of sorted runs of ``(key, val)``."""
    iters = [iter(r) for r in runs]
    heap: List[Tuple[bytes, int, int, KV]] = []
    for ri, it in enumerate(iters):
        for kv in it:
            heap.append((sort_key(kv[0]), ri, 0, kv))
            break
    heapq.heapify(heap)
    while heap:
        _, ri, seq, kv = heap[0]
        yield kv
        nxt = next(iters[ri], None)
        if nxt is None:
            heapq.heappop(heap)
        else:
            heapq.heapreplace(heap, (sort_key(nxt[0]), ri, seq + 1, nxt))


def merged_stream(records: Iterable[KV]) -> bytes:
    """The byte stream handed to the Java reducer."""
    return b"".join(encode_record(k, v) for k, v in records) + EOF_MARKER


def parse_stream(buf) -> List[KV]:
    """Inverse of :func:`merged_stream`; mirrors J2CQueue.next
    (plugins/shared/com/mellanox/hadoop/mapred/UdaPlugin.java:498-538): stop at negative lengths."""
    out: List[KV] = []
    pos = 0
    while pos < len(buf):
        klen, n1 = decode_vlong(buf, pos)
        vlen, n2 = decode_vlong(buf, pos + n1)
        if klen < 0 or vlen < 0:
            break
        pos += n1 + n2
        out.append((bytes(buf[pos : pos + klen]), bytes(buf[pos + klen : pos + klen + vlen])))
        pos += klen + vlen
    return out
