"""Fault-injection hooks for tests of the failure -> fallback contract.

The reference has no fault-injection framework: its regression only has configuration-driven negative
tests (fallback expected, zero-size fetch VUDA-35, scripts/regression/testStatusAnalyzer.sh:294-319). The GPU data
path has three failure classes of its own, each with an injector here and a detector in the engine:

=====================  ==================================================  ===========================================
fault                  injector                                            detector
=====================  ==================================================  ===========================================
lost / short push      :func:`drop_push_segment` (zero one descriptor)     wait_flags spin budget -> status word ->
                                                                           ``runtime.failure.check_wait_status``
corrupted segment      :func:`corrupt_segment` (flip a byte in a MOF)      :func:`verify_landed_crc` (uda_crc32 over the
                                                                           landed bytes vs the IFile trailer)
arena allocation       :func:`arena_oom` (context manager)                 ``SymmetricArena`` raises -> ArenaAllocationFailed
=====================  ==================================================  ===========================================
"""
from __future__ import annotations

import contextlib
from typing import List, Sequence, Tuple

import numpy as np
import torch

from ..ops import kernels as K
from ..runtime.failure import ArenaAllocationFailed, ChecksumMismatch


def drop_push_segment(descs_dev: torch.Tensor, seg: int, keep_bytes: int = 0) -> torch.Tensor:
    """Returns a copy of a push descriptor table (uint8 view of ``UdaPushDesc[]``) in which segment ``seg`` only
    sends ``keep_bytes``: the destination's landed-byte counter never reaches the expected value."""
    d = descs_dev.clone()
    host = d.cpu().numpy().view(K.PUSH_DESC_DTYPE).copy()
    host["bytes"][seg] = keep_bytes
    d.copy_(torch.from_numpy(host.view(np.uint8).reshape(-1)))
    return d


def corrupt_segment(mof, partition: int, byte: int = 7, xor: int = 0x40) -> None:
    """Flips bits of one payload byte of ``partition``'s segment in a device MOF (in place)."""
    start, raw, _part = (int(x) for x in mof.index[partition])
    if raw <= 2:
        raise ValueError("empty segment: nothing to corrupt")
    pos = start + min(byte, raw - 3)
    mof.data[pos] ^= xor


def verify_landed_crc(segments: Sequence[Tuple[torch.Tensor, int]]) -> None:
    """``segments[i] = (uint8 device tensor at the segment start, rawLength)``; the 4 bytes after ``rawLength`` are
    the big-endian CRC32 of the preceding bytes (Hadoop IFile trailer; the reference transfers and ignores it,
    src/MOFServer/IndexInfo.cc:308-310). One kernel pair per segment, one host sync for all. Raises
    :class:`ChecksumMismatch` naming the bad segments."""
    if not segments:
        return
    dev = segments[0][0].device
    got = torch.empty((len(segments), 4), dtype=torch.uint8, device=dev)
    want = torch.empty_like(got)
    for i, (seg, raw) in enumerate(segments):
        K.crc32(seg, raw, out_be4=got[i])
        want[i].copy_(seg[raw:raw + 4])
    bad: List[int] = torch.nonzero((got != want).any(dim=1)).flatten().cpu().tolist()
    if bad:
        raise ChecksumMismatch(f"CRC32 mismatch in landed segment(s) {bad[:8]}{'…' if len(bad) > 8 else ''}")


@contextlib.contextmanager
def arena_oom():
    """Inside the block every symmetric-arena allocation fails the way a cudaMalloc failure does."""
    from ..parallel import arena as A

    orig = A.dev_malloc

    def failing(nbytes: int):
        raise ArenaAllocationFailed(f"injected: cudaMalloc of {nbytes} bytes for the receive arena failed (error 2: out of memory)")

    A.dev_malloc = failing
    try:
        yield
    finally:
        A.dev_malloc = orig
