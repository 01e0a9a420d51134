"""Shuffle plan: which bytes go where, derived from the all-gathered MOF index tables.

Reference behaviour being replaced: the reducer learns a segment's size only from the first ack
(``rawLen:partLen:sent:offset:path`` parsed in src/Merger/MergeManager.cc:367-409) and the provider
learns the destination address from each fetch request (src/MOFServer/MOFServlet.cc:28-96). With the
push model both sides need the full picture up front, so the per-map index tables
(``[P,3] = startOffset, rawLength, partLength`` — the ``file.out.index`` content) are
all-gathered once per job. Every rank then derives, with the same deterministic function,

* the layout of every reducer's receive arena (run ``(src_rank, map)`` -> offset), and
* its own push descriptor list (segment -> peer pointer + flag pointer),

so no addresses are ever exchanged at fetch time.

Peer rotation: rank ``s`` orders its descriptors by destination ``(s+1+i) % P`` — at any moment the
ranks target a permutation of the peers, which replaces the reference's randomised fetch order
(``list_shuffle_in_vector``, src/include/UdaUtil.h:97-101) as the hot-spot avoidance mechanism.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence

import numpy as np

from ..ops.kernels import GATHER_SLACK, PUSH_DESC_DTYPE, RUN_DESC_DTYPE

ALIGN = 16


def _align(x: int, a: int = ALIGN) -> int:
    return (x + a - 1) // a * a


@dataclass
class ReducerLayout:
    """Receive-side view for one reducer rank."""

    run_offsets: np.ndarray  # [nruns] byte offset of each run's segment in the arena data region
    run_bytes: np.ndarray    # [nruns] partLength
    run_nrec: np.ndarray     # [nruns] records (fixed-stride jobs) or -1
    total_bytes: int


class ShufflePlan:
    """``index[s][m][p] = (startOffset, rawLength, partLength)`` for every rank ``s``, local map ``m``,
    partition ``p``; all ranks hold the same array. ``record_stride`` > 0 marks a fixed-stride job."""

    def __init__(self, index: np.ndarray, record_stride: int = 0, subparts: int = 1, slots: int = 0):
        index = np.asarray(index, dtype=np.int64)
        assert index.ndim == 4 and index.shape[3] == 3, "index must be [world, maps, P*subparts, 3]"
        self.index = index
        self.world, self.maps = index.shape[:2]
        self.subparts = subparts
        self.P = index.shape[2] // subparts
        assert self.P == self.world and self.P * subparts == index.shape[2], "one reducer per rank (reference: src/Merger/reducer.h:137)"
        self.stride = record_stride
        self.nruns = self.world * self.maps
        #: receive slots per reducer. ``slots == subparts`` (default): the arena holds the whole partition, every
        #: sub-partition has a place of its own and a push never waits. ``slots < subparts``: a ring — sub-partition j
        #: lands in slot ``j % slots`` and may only be pushed once the reducer has merged sub-partition ``j - slots``
        #: (credits in peer memory, see uda_credit_publish), so receive memory is ``slots / subparts`` of the partition.
        #: This is the counterpart of the reference's two 1 MiB buffers per map with credit-based flow control
        #: (src/Merger/StreamRW.cc:462-486, src/DataNet/RDMAServer.cc:44-136).
        self.slots = subparts if slots <= 0 or slots > subparts else slots
        #: flat list over (reducer r, sub-partition j): self.layouts[r * subparts + j]
        self.layouts: List[ReducerLayout] = []
        self.arena_bytes: List[int] = []
        for r in range(self.P):
            base = _align(GATHER_SLACK)
            if self.slots == subparts:
                cur = base
                for j in range(subparts):   # one after the other
                    lay, cur = self._layout(r, j, cur)
                    self.layouts.append(lay)
                self.arena_bytes.append(_align(cur + GATHER_SLACK))
            else:
                # every slot is as large as the largest sub-partition of this reducer
                slot_bytes = _align(max(self._layout(r, j, 0)[1] for j in range(subparts)), 256)
                for j in range(subparts):
                    lay, _ = self._layout(r, j, base + (j % self.slots) * slot_bytes)
                    self.layouts.append(lay)
                self.arena_bytes.append(_align(base + self.slots * slot_bytes + GATHER_SLACK))

    def run_id(self, src_rank: int, m: int) -> int:
        return src_rank * self.maps + m

    def layout(self, r: int, j: int = 0) -> ReducerLayout:
        return self.layouts[r * self.subparts + j]

    def _layout(self, r: int, j: int, cur: int):
        offs = np.zeros(self.nruns, dtype=np.int64)
        nbytes = np.zeros(self.nruns, dtype=np.int64)
        nrec = np.full(self.nruns, -1, dtype=np.int64)
        lp = r * self.subparts + j  # logical partition
        for s in range(self.world):
            for m in range(self.maps):
                start, raw, part = (int(x) for x in self.index[s, m, lp])
                rid = self.run_id(s, m)
                # keep the source's 16-byte phase so the copy is vectorised end to end
                cur = _align(cur) + (start % ALIGN)
                offs[rid] = cur
                nbytes[rid] = part
                if self.stride:
                    body = raw - 2  # rawLength includes the 2-byte EOF marker
                    assert body % self.stride == 0, "fixed-stride job with a ragged segment"
                    nrec[rid] = body // self.stride
                cur += part
        cur = _align(cur + GATHER_SLACK)
        return ReducerLayout(offs, nbytes, nrec, cur), cur

    def max_arena_bytes(self) -> int:
        return max(self.arena_bytes)

    # ---- sender side ------------------------------------------------------------------------
    def push_descs(self, rank: int, mof_ptrs: Sequence[int], peer_data_ptrs: Sequence[int],
                   peer_flag_base: Sequence[int], sub: int = 0) -> np.ndarray:
        """Descriptor table of sub-partition ``sub`` for ``rank``: ``mof_ptrs[m]`` = device address of local map
        ``m``'s file.out bytes; ``peer_data_ptrs[p]`` / ``peer_flag_base[p]`` = rank ``p``'s arena data / flag
        region as mapped here. Flag slot of (sub, run) is ``sub * nruns + run``."""
        d = np.zeros(self.maps * self.P, dtype=PUSH_DESC_DTYPE)
        i = 0
        for step in range(self.P):
            p = (rank + 1 + step) % self.P
            lay = self.layout(p, sub)
            for m in range(self.maps):
                start, _raw, part = (int(x) for x in self.index[rank, m, p * self.subparts + sub])
                rid = self.run_id(rank, m)
                d[i] = (mof_ptrs[m] + start, peer_data_ptrs[p] + int(lay.run_offsets[rid]), part,
                        peer_flag_base[p] + 8 * (sub * self.nruns + rid))
                i += 1
        return d

    # ---- receiver side ----------------------------------------------------------------------
    def run_descs(self, rank: int, arena_data_ptr: int, sub: int = 0):
        """``(UdaRunDesc table, run_off[nruns+1], run_base[nruns], expected_flag_bytes[nruns])`` for sub-partition
        ``sub`` of reducer ``rank``."""
        lay = self.layout(rank, sub)
        d = np.zeros(self.nruns, dtype=RUN_DESC_DTYPE)
        base = arena_data_ptr + lay.run_offsets
        d["base"] = base.astype(np.uint64)
        d["nrec"] = np.maximum(lay.run_nrec, 0)
        off = np.zeros(self.nruns + 1, dtype=np.int64)
        np.cumsum(d["nrec"], out=off[1:])
        d["out_off"] = off[:-1]
        d["run_id"] = np.arange(self.nruns, dtype=np.uint32)
        d["idx0"] = 0
        return d, off, base.astype(np.uint64), lay.run_bytes.copy()
