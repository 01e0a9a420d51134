"""Partitioners (key -> reducer).

The reference does not partition: Hadoop's MapTask does, and UDA only moves partition
``r`` of every map output to reducer ``r`` (plugins/shared/com/mellanox/hadoop/mapred/UdaPlugin.java:331
adds ``reduceTask.getPartition()``; the provider serves ``(job,map,reduce)``,
src/MOFServer/IndexInfo.cc:243-251). The new engine owns the map side too, so it needs the
two Hadoop partitioners used by the reference's regression workloads
(scripts/regression/sortcountRunner.sh: terasort = total-order range, sort/wordcount = hash).
"""
from __future__ import annotations

import bisect
from typing import List, Sequence


def make_splitters(num_partitions: int, key_len: int = 10) -> List[bytes]:
    """``P-1`` equally spaced big-endian splitters over the ``key_len``-byte key space."""
    space = 1 << (8 * key_len)
    return [((space * i) // num_partitions).to_bytes(key_len, "big") for i in range(1, num_partitions)]


def range_partition(key: bytes, splitters: Sequence[bytes]) -> int:
    """TotalOrderPartitioner: index of the first splitter strictly greater than ``key``."""
    return bisect.bisect_right(splitters, key)


def java_hash_bytes(b: bytes) -> int:
    """``WritableComparator.hashBytes``: h = 1; h = 31*h + (signed byte), int32 wraparound."""
    h = 1
    for x in b:
        h = (31 * h + (x - 256 if x > 127 else x)) & 0xFFFFFFFF
    return h


def hash_partition(key: bytes, num_partitions: int) -> int:
    """Hadoop HashPartitioner: ``(hash & Integer.MAX_VALUE) % P``."""
    return (java_hash_bytes(key) & 0x7FFFFFFF) % num_partitions
