"""Hybrid merge for data that does not fit the HBM budget: bounded groups of runs are merged on the device and
spilled to pinned host DRAM on a side stream, then the spilled runs are cut into key ranges that fit the budget
and merged range by range while the next range is already being copied in (BASELINE.json config 4).

Reference behaviour being replaced: the hybrid merge (src/Merger/MergeManager.cc:202-288) builds sqrt(N) groups
("LPQs", bounded by ``mapred.rdma.num.parallel.lpqs``), merges each to a local spill file
``<dir>/uda.<task>.lpq-NNN`` with buffered stdio (src/Merger/StreamRW.cc:863-887) and runs a final merge ("RPQ")
over the spill files (src/Merger/StreamRW.cc:813-861); AIOHandler (src/CommUtils/AIOHandler.cc:80-150) is the
reference's asynchronous disk reader. Here the spill medium is pinned host memory and the asynchronous engine is
``cudaMemcpyAsync`` on dedicated copy streams, so H2D of group i+1, the merge of group i and the D2H of group
i-1 overlap.

Final-merge ranges: equi-depth splitters are sampled from the first merged group
(``uda_sample_splitters``); every merged group is cut at those keys (``uda_lower_bound_records``) while it is
still on the device, so the final phase needs no search in host memory unless the keys are skewed: a range that the
sampled splitters left larger than the staging buffer is halved at median keys on the host (``split_range``; binary
searches in the pinned spill buffers), down to pieces of a single key value, which are cut by position.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import kernels as K


Slice = Tuple[int, int, int]   # (spilled group, first record, one past the last record)


def _key_at(buf: np.ndarray, i: int, S: int, key_off: int, key_len: int) -> bytes:
    o = i * S + key_off
    return buf[o:o + key_len].tobytes()


def _bisect(buf: np.ndarray, lo: int, hi: int, S: int, key_off: int, key_len: int, key: bytes, upper: bool) -> int:
    """First record in [lo, hi) whose key is >= ``key`` (``upper``: > ``key``); keys compare as unsigned bytes."""
    while lo < hi:
        mid = (lo + hi) // 2
        k = _key_at(buf, mid, S, key_off, key_len)
        if k < key or (upper and k == key):
            lo = mid + 1
        else:
            hi = mid
    return lo


def split_range(spilled: Sequence[np.ndarray], slices: Sequence[Slice], S: int, key_off: int, key_len: int, cap_records: int) -> List[List[Slice]]:
    """Cuts one final-merge key range that is larger than the staging buffer into consecutive sub-ranges of at most
    ``cap_records`` records each (skewed keys: the sampled splitters are equi-depth only for the group they were sampled
    from). ``slices`` names the range's sorted piece in every spilled group; the groups live in host memory, so the
    search runs on the host: the range is halved at the median key of its largest piece until every part fits. Records
    that share one key cannot be separated by a key, and need not be: any order among them is a valid merge order
    (src/Merger/MergeQueue.h:250-269 breaks ties arbitrarily as well), so such a part is cut by position, group by
    group. Returns the sub-ranges in key order; merging each of them and concatenating the results is the merged range."""
    assert cap_records >= 1
    slices = [(g, a, b) for g, a, b in slices if b > a]
    total = sum(b - a for _, a, b in slices)
    if total <= cap_records:
        return [slices] if slices else []
    kmin = min(_key_at(spilled[g], a, S, key_off, key_len) for g, a, _ in slices)
    kmax = max(_key_at(spilled[g], b - 1, S, key_off, key_len) for g, _, b in slices)
    if kmin == kmax:
        # one key value: pack whole pieces greedily, cut a piece that alone is too large by position
        tasks, cur, room = [], [], cap_records
        for g, a, b in slices:
            while a < b:
                take = min(b - a, room)
                cur.append((g, a, a + take))
                a += take
                room -= take
                if room == 0:
                    tasks.append(cur)
                    cur, room = [], cap_records
        if cur:
            tasks.append(cur)
        return tasks
    g, a, b = max(slices, key=lambda t: t[2] - t[1])
    pivot = _key_at(spilled[g], (a + b) // 2, S, key_off, key_len)
    if pivot == kmin:
        # everything below the median of the largest piece is the minimum key: peel that key off (left part = one key
        # value, handled above), continue with the rest. kmax > kmin guarantees a next key exists.
        nxt = []
        for g2, a2, b2 in slices:
            u = _bisect(spilled[g2], a2, b2, S, key_off, key_len, kmin, upper=True)
            if u < b2:
                nxt.append(_key_at(spilled[g2], u, S, key_off, key_len))
        pivot = min(nxt)
    left, right = [], []
    for g2, a2, b2 in slices:
        m = _bisect(spilled[g2], a2, b2, S, key_off, key_len, pivot, upper=False)
        left.append((g2, a2, m))
        right.append((g2, m, b2))
    return (split_range(spilled, left, S, key_off, key_len, cap_records)
            + split_range(spilled, right, S, key_off, key_len, cap_records))


class SpillMerger:
    """Fixed-stride runs in pinned host memory -> merged stream delivered block by block to ``sink``."""

    def __init__(self, stride: int, key_off: int, key_len: int, hbm_budget_bytes: int, device, max_group_runs: int = 64):
        self.stride, self.key_off, self.key_len = stride, key_off, key_len
        self.budget, self.device, self.max_group_runs = hbm_budget_bytes, device, max_group_runs
        self.s_in, self.s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
        self.stats = {}
        #: pinned spill buffers, kept across merges: pinning host memory costs ~0.1 s per GB, far more than the copy it
        #: serves, so the buffers of the first merge are reused by the following ones (the reference's LPQ files are
        #: likewise created once per reduce task, src/Merger/MergeManager.cc:202-288)
        self._pinned: List[torch.Tensor] = []

    def _spill_buffer(self, i: int, nbytes: int) -> torch.Tensor:
        """Pinned host buffer number ``i`` of this merger with room for ``nbytes`` (grown when a later merge needs more)."""
        while len(self._pinned) <= i:
            self._pinned.append(torch.empty(0, dtype=torch.uint8))
        if self._pinned[i].numel() < nbytes:
            self._pinned[i] = torch.empty(nbytes + nbytes // 16, dtype=torch.uint8, pin_memory=True)
        return self._pinned[i][:nbytes]

    # ------------------------------------------------------------------------------------------------
    def _plan_groups(self, nrecs: Sequence[int]) -> List[List[int]]:
        """Consecutive runs per group so that a group's input + output fit a third of the budget each
        (two input buffers in flight + one output)."""
        cap = max(self.stride, self.budget // 4)
        groups, cur, size = [], [], 0
        for i, n in enumerate(nrecs):
            b = n * self.stride
            if b > cap:
                raise ValueError(f"run {i} ({b} bytes) alone exceeds a quarter of the HBM budget ({cap} bytes)")
            if cur and (size + b > cap or len(cur) >= self.max_group_runs):
                groups.append(cur)
                cur, size = [], 0
            cur.append(i)
            size += b
        if cur:
            groups.append(cur)
        return groups

    def merge(self, host_runs: Sequence[torch.Tensor], nrecs: Sequence[int], sink: Callable[[torch.Tensor, int], None]) -> int:
        """``host_runs[i]``: pinned uint8 tensor holding ``nrecs[i]`` sorted records. ``sink(device_block, nbytes)``
        is called in stream order on the current stream for consecutive pieces of the merged stream (the last piece
        ends with the EOF marker). Returns the number of records."""
        dev, S = self.device, self.stride
        cur = torch.cuda.current_stream(dev)
        total = int(sum(nrecs))
        groups = self._plan_groups(nrecs)
        cap = max(sum(nrecs[i] for i in g) for g in groups) * S
        single = len(groups) == 1
        slack = K.GATHER_SLACK
        d_in = [torch.empty(cap + 64 * len(host_runs) + 2 * slack + 64, dtype=torch.uint8, device=dev) for _ in range(2)]
        d_out = [torch.empty(cap + 64, dtype=torch.uint8, device=dev) for _ in range(1 if single else 2)]
        max_n = cap // S
        scratch = torch.empty(K.merge_records_scratch_bytes(max_n + 64 * len(groups), max(max(len(g) for g in groups), len(groups))),
                              dtype=torch.uint8, device=dev)
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_merged = [torch.cuda.Event() for _ in range(2)]
        ev_spilled = [torch.cuda.Event() for _ in range(2)]
        spilled: List[torch.Tensor] = []
        cuts: List[torch.Tensor] = []
        splitters = None

        def stage(gi: int, slot: int):
            """H2D of group gi into d_in[slot]; returns (bases, counts)."""
            bases, off = [], slack
            with torch.cuda.stream(self.s_in):
                if gi >= 2:
                    self.s_in.wait_event(ev_merged[slot])  # the merge that read this buffer two groups ago is done
                for r in groups[gi]:
                    nb = nrecs[r] * S
                    off = (off + 15) // 16 * 16
                    d_in[slot][off:off + nb].copy_(host_runs[r][:nb], non_blocking=True)
                    bases.append(d_in[slot].data_ptr() + off)
                    off += nb
                ev_in[slot].record(self.s_in)
            return bases, [nrecs[r] for r in groups[gi]]

        def device_merge(bases, counts, out, write_eof):
            desc, _ = K.make_run_descs(bases, counts)
            n = int(sum(counts))
            K.merge_records_fixed(K.to_device_struct(desc, dev), len(bases), n, max(counts) if counts else 0, S, self.key_off,
                                  self.key_len, out, scratch, write_eof=write_eof, align2=(S % 2 == 0))
            return n

        # ---------------- phase 1: merge every group, spill it (the reference's LPQ phase)
        staged = stage(0, 0)
        for gi in range(len(groups)):
            slot = gi % 2
            nxt = stage(gi + 1, (gi + 1) % 2) if gi + 1 < len(groups) else None
            cur.wait_event(ev_in[slot])
            oslot = 0 if single else slot
            if gi >= 2:
                cur.wait_event(ev_spilled[oslot])              # its previous content has left for host memory
            n = device_merge(staged[0], staged[1], d_out[oslot], write_eof=single)
            ev_merged[slot].record(cur)
            if single:
                sink(d_out[0], n * S + 2)
                self.stats = {"groups": 1, "ranges": 0, "spilled_bytes": 0}
                return total
            if splitters is None:
                # ranges are equi-depth only in expectation (splitters come from one group): aim at 60 % of the staging buffer
                nranges = max(2, -(-total * S // max(S, int(cap * 0.6))))
                splitters = K.sample_splitters(d_out[oslot], n, S, self.key_off, self.key_len, nranges - 1)
            cuts.append(K.lower_bound_records(d_out[oslot], n, S, self.key_off, self.key_len, splitters))
            h = self._spill_buffer(gi, n * S)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(ev_merged[slot])
                h.copy_(d_out[oslot][: n * S], non_blocking=True)   # the spill: pinned DRAM instead of uda.<task>.lpq-NNN
                ev_spilled[oslot].record(self.s_out)
            spilled.append(h)
            staged = nxt

        # ---------------- phase 2: final merge of the spilled runs, key range by key range (the RPQ phase)
        self.s_out.synchronize()                                   # spill buffers complete; cut tables are needed on the host
        G = len(spilled)
        gn = [h.numel() // S for h in spilled]
        cut = np.stack([np.concatenate([[0], c.cpu().numpy(), [gn[i]]]) for i, c in enumerate(cuts)])  # [G, nranges+1]
        nranges = cut.shape[1] - 1
        # a range the sampled splitters left larger than the staging buffer (skewed keys) is cut further on the host
        cap_records = cap // S
        views = [h.numpy() for h in spilled]
        tasks: List[List[Slice]] = []
        refined = 0
        for c in range(nranges):
            parts = split_range(views, [(g, int(cut[g, c]), int(cut[g, c + 1])) for g in range(G)], S, self.key_off, self.key_len, cap_records)
            refined += len(parts) > 1
            tasks.extend(parts)
        if not tasks:
            tasks = [[]]                                           # no records at all: one empty merge writes the EOF marker
        ev_in2 = [torch.cuda.Event() for _ in range(2)]
        ev_done2 = [torch.cuda.Event() for _ in range(2)]

        def stage_range(c: int, slot: int):
            piece = {g: (a, b) for g, a, b in tasks[c]}
            bases, counts, off = [], [], slack
            with torch.cuda.stream(self.s_in):
                if c >= 2:
                    self.s_in.wait_event(ev_done2[slot])
                for g in range(G):
                    a, b = piece.get(g, (0, 0))
                    off = (off + 15) // 16 * 16
                    if b > a:
                        d_in[slot][off:off + (b - a) * S].copy_(spilled[g][a * S:b * S], non_blocking=True)
                    bases.append(d_in[slot].data_ptr() + off)
                    counts.append(b - a)
                    off += (b - a) * S
                ev_in2[slot].record(self.s_in)
            return bases, counts

        staged = stage_range(0, 0)
        for c in range(len(tasks)):
            slot = c % 2
            nxt = stage_range(c + 1, (c + 1) % 2) if c + 1 < len(tasks) else None
            cur.wait_event(ev_in2[slot])
            last = c == len(tasks) - 1
            n = device_merge(staged[0], staged[1], d_out[slot], write_eof=last)
            sink(d_out[slot], n * S + (2 if last else 0))
            ev_done2[slot].record(cur)
            staged = nxt
        self.stats = {"groups": G, "ranges": len(tasks), "refined_ranges": int(refined),
                      "spilled_bytes": int(sum(h.numel() for h in spilled))}
        return total
