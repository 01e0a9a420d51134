"""Hadoop "sort" example on the engine: variable-length records, BytesWritable or Text keys.

The reference's regression runs this workload next to TeraSort (RandomWriter / RandomTextWriter input, Hadoop
``testmapredsort`` as the oracle, scripts/regression/mr-dstatExcel.sh:341-348). It is the job that exercises every
piece the TeraSort fast path skips: record boundary discovery in landed segments, the per-key-class comparator
(src/Merger/CompareFunc.cc:70-113) including keys that tie on the sort prefix, and the variable-length gather.

Phases::

    setup   random records generated + IFile-encoded on the device -> VarMapTask per map (sort, range partition,
            segments + CRC) -> all-gather index and record-count tables -> symmetric arena, push descriptors
    step    [timed] fused shuffle push -> landed-flag acquire -> VarMerger (decode, normalise, k-way merge,
            tie-break, gather)
    check   the merged stream of this rank equals, byte for byte, a stable CPU sort (exact comparator) of the
            records that were pushed to it, in (source rank, map) segment order
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..mof import DeviceMOF, uniform_splitters_w0
from ..ops import kernels as K
from ..ops.varlen import KEY_BYTES_WRITABLE, KEY_TEXT, VarMapTask, VarMerger
from ..parallel.arena import SymmetricArena
from ..parallel.shuffle import ShufflePlan
from ..utils import ifile


def generate_segment(n: int, seed: int, device, key_kind: int = KEY_BYTES_WRITABLE, key_len: Tuple[int, int] = (10, 100),
                     val_len: Tuple[int, int] = (0, 200), tie_fraction: float = 0.0, tie_pool: int = 4,
                     pool_seed: Optional[int] = None) -> Tuple[torch.Tensor, int]:
    """One unsorted IFile segment of ``n`` random records, built on the device (RandomWriter's shape: random
    key/value sizes in a range, random bytes). ``tie_fraction`` of the keys start with one of ``tie_pool`` fixed
    10-byte strings (drawn from ``pool_seed``, so that all maps of a job share them), so that they tie on the sort
    prefix across maps and only the exact comparator can order them.
    Returns ``(segment incl. EOF marker, its length)``."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    kmin, kmax = key_len
    vmin, vmax = val_len
    assert key_kind in (KEY_BYTES_WRITABLE, KEY_TEXT) and 0 <= kmin <= kmax and 0 <= vmin <= vmax
    if key_kind == KEY_TEXT:
        assert kmax <= 127, "Text keys are generated with a one-byte VInt length"
    hdr = 4 if key_kind == KEY_BYTES_WRITABLE else 1
    klen = torch.randint(kmin, kmax + 1, (n,), generator=g, device=device, dtype=torch.int64)
    vlen = torch.randint(vmin, vmax + 1, (n,), generator=g, device=device, dtype=torch.int64)
    kser = klen + hdr
    key_off = torch.zeros(n + 1, dtype=torch.int64, device=device)
    val_off = torch.zeros(n + 1, dtype=torch.int64, device=device)
    torch.cumsum(kser, 0, out=key_off[1:])
    torch.cumsum(vlen, 0, out=val_off[1:])
    tk, tv = (int(x) for x in torch.stack([key_off[-1], val_off[-1]]).cpu().tolist())
    keys = torch.randint(0, 256, (tk + 16,), generator=g, device=device, dtype=torch.uint8)
    vals = torch.randint(0, 256, (tv + 16,), generator=g, device=device, dtype=torch.uint8)
    ko = key_off[:-1]
    if key_kind == KEY_BYTES_WRITABLE:
        for b in range(4):
            keys[ko + b] = ((klen >> (8 * (3 - b))) & 0xFF).to(torch.uint8)
    else:
        keys[ko] = klen.to(torch.uint8)
    if tie_fraction > 0 and n:
        gp = torch.Generator(device=device)
        gp.manual_seed(seed if pool_seed is None else pool_seed)
        pool = torch.randint(0, 256, (tie_pool, 10), generator=gp, device=device, dtype=torch.uint8)
        sel = torch.rand(n, generator=g, device=device) < tie_fraction
        pid = torch.randint(0, tie_pool, (n,), generator=g, device=device)
        for j in range(10):
            m = sel & (klen > j)
            keys[(ko + hdr + j)[m]] = pool[pid[m], j]
    seg = K.ifile_encode_var(keys, ko.contiguous(), kser.to(torch.int32), vals, val_off[:-1].contiguous(), vlen.to(torch.int32))
    return seg, int(seg.numel())


def _sort_key(key_kind: int):
    if key_kind == KEY_BYTES_WRITABLE:
        return lambda kv: kv[0][4:]
    if key_kind == KEY_TEXT:
        from ..utils import vint
        return lambda kv: kv[0][vint.decode_vint_size(kv[0][0]) if kv[0] else 0:]
    return lambda kv: kv[0]


class SortJob:
    def __init__(self, records_per_rank: int, maps_per_rank: int, device: Optional[torch.device] = None, seed: int = 99,
                 key_kind: int = KEY_BYTES_WRITABLE, key_len: Tuple[int, int] = (10, 100), val_len: Tuple[int, int] = (0, 200),
                 tie_fraction: float = 0.0, group=None, push_ctas: int = 0, side_index: bool = True):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        self.n_local, self.maps, self.seed = records_per_rank, maps_per_rank, seed
        self.key_kind, self.key_len, self.val_len, self.tie_fraction = key_kind, key_len, val_len, tie_fraction
        self.push_ctas = push_ctas
        #: ship the maps' record-offset side index with the segments (8 bytes per record) so that the reducer does not
        #: rediscover record boundaries; False = segments only, as a MOF written by Hadoop would arrive
        self.side_index = side_index
        self.mofs: List[DeviceMOF] = []
        self.steps_run = 0

    # ------------------------------------------------------------------------------------ setup
    def setup(self) -> "SortJob":
        dev, P = self.device, self.world
        per_map = [self.n_local // self.maps + (1 if i < self.n_local % self.maps else 0) for i in range(self.maps)]
        splitters = uniform_splitters_w0(P)
        task = VarMapTask(dev, self.key_kind, P)
        self.input_bytes = 0
        for m, nm in enumerate(per_map):
            seg, seg_len = generate_segment(nm, self.seed + 1000 * self.rank + m, dev, self.key_kind, self.key_len, self.val_len,
                                            self.tie_fraction, pool_seed=self.seed)
            self.input_bytes += seg_len - 2
            self.mofs.append(task.run(seg, seg_len, nm, splitters))
            del seg
        torch.cuda.synchronize(dev)
        local_idx = np.stack([m.index for m in self.mofs])              # [maps, P, 3]
        local_nrec = np.stack([m.nrec for m in self.mofs])              # [maps, P]
        if P > 1:
            t = torch.from_numpy(np.concatenate([local_idx, local_nrec[..., None]], axis=2)).to(dev)
            allt = torch.empty((P,) + tuple(t.shape), dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(allt, t, group=self.group)
            both = allt.cpu().numpy()
            index, nrec = both[..., :3], both[..., 3]
        else:
            index, nrec = local_idx[None], local_nrec[None]
        self.plan = ShufflePlan(index, record_stride=0)
        nruns = self.plan.nruns
        self.arena = SymmetricArena(self.plan.max_arena_bytes(), nruns, dev, self.group)
        ptrs = [m.data.data_ptr() for m in self.mofs]
        self.push_descs_dev = K.to_device_struct(
            self.plan.push_descs(self.rank, ptrs, [self.arena.peer_data_ptr(p) for p in range(P)],
                                 [self.arena.peer_flag_ptr(p, 0) for p in range(P)], 0), dev)
        self.nseg = self.maps * P
        self.push_scratch = torch.empty(K.push_scratch_bytes(self.nseg), dtype=torch.uint8, device=dev)
        lay = self.plan.layout(self.rank)
        data = self.arena.data()
        #: landed segments of this reducer in run order: (tensor view, raw length incl. EOF, exact record count)
        self.segments = []
        for s in range(P):
            for m in range(self.maps):
                rid = self.plan.run_id(s, m)
                raw = int(index[s, m, self.rank, 1])
                self.segments.append((data[int(lay.run_offsets[rid]):], raw, int(nrec[s, m, self.rank])))
        self.n_out = int(sum(c for _, _, c in self.segments))
        self.bytes_out = int(sum(r - 2 for _, r, _ in self.segments)) + 2
        self.expect_step = torch.from_numpy(lay.run_bytes.astype(np.int64)).to(dev)
        self.expect = torch.zeros_like(self.expect_step)
        self.rec_offs = None
        if self.side_index:
            # second, small shuffle of the side indexes: own plan (sizes follow from the record counts every rank
            # already holds), own arena, same push kernel
            sidx = np.zeros_like(index)
            for s in range(P):
                for m in range(self.maps):
                    cur = 0
                    for p in range(P):
                        nb = (int(nrec[s, m, p]) + 1) * 8
                        sidx[s, m, p] = (cur, nb, nb)
                        cur = (cur + nb + 15) // 16 * 16
            self.side_plan = ShufflePlan(sidx, record_stride=0)
            self.side_arena = SymmetricArena(self.side_plan.max_arena_bytes(), nruns, dev, self.group)
            self.side_descs_dev = K.to_device_struct(
                self.side_plan.push_descs(self.rank, [m.meta["side"].data_ptr() for m in self.mofs],
                                          [self.side_arena.peer_data_ptr(p) for p in range(P)],
                                          [self.side_arena.peer_flag_ptr(p, 0) for p in range(P)], 0), dev)
            self.side_scratch = torch.empty(K.push_scratch_bytes(self.nseg), dtype=torch.uint8, device=dev)
            slay = self.side_plan.layout(self.rank)
            sdata = self.side_arena.data()
            self.rec_offs = []
            for s in range(P):
                for m in range(self.maps):
                    rid = self.side_plan.run_id(s, m)
                    o, nb = int(slay.run_offsets[rid]), (int(nrec[s, m, self.rank]) + 1) * 8
                    assert o % 8 == 0
                    self.rec_offs.append(sdata[o:o + nb].view(torch.int64))
            self.side_expect_step = torch.from_numpy(slay.run_bytes.astype(np.int64)).to(dev)
            self.side_expect = torch.zeros_like(self.side_expect_step)
        self.merger = VarMerger(dev, self.key_kind, self.n_out, nruns, self.bytes_out)
        torch.cuda.synchronize(dev)
        if P > 1:
            dist.barrier(group=self.group)
        return self

    # ------------------------------------------------------------------------------------ step
    def step(self) -> torch.Tensor:
        self.steps_run += 1
        if self.world > 1:
            self.arena.barrier()
        self.expect.add_(self.expect_step)
        K.shuffle_push(self.push_descs_dev, self.nseg, self.push_scratch, self.push_ctas, self.world)
        if self.side_index:
            if self.world > 1:
                self.side_arena.barrier()
            self.side_expect.add_(self.side_expect_step)
            K.shuffle_push(self.side_descs_dev, self.nseg, self.side_scratch, self.push_ctas, self.world)
            K.wait_flags(self.side_arena.flags()[: self.plan.nruns], self.side_expect, self.plan.nruns, self.side_arena.status)
        K.wait_flags(self.arena.flags()[: self.plan.nruns], self.expect, self.plan.nruns, self.arena.status)
        out, _ = self.merger.merge(self.segments, exact_counts=True, rec_offs=self.rec_offs)
        return out

    @property
    def launches_per_step(self) -> int:
        return (2 + 1) * (2 if self.side_index else 1) + VarMerger.launches(len(self.segments), self.side_index) + (2 if self.world > 1 else 0)

    # ------------------------------------------------------------------------------------ check
    def check(self, out: torch.Tensor) -> dict:
        """Byte-exact comparison against the CPU oracle (stable sort of the pushed records with the exact comparator)."""
        torch.cuda.synchronize(self.device)
        status = int(self.arena.status.cpu()[0]) if hasattr(self.arena, "status") else 0
        got = bytes(out.cpu().numpy())
        recs = []
        for seg, raw, cnt in self.segments:
            part = ifile.decode_segment(bytes(seg[:raw].cpu().numpy()))
            assert len(part) == cnt
            recs.extend(part)
        keyf = _sort_key(self.key_kind)
        want = ifile.encode_segment(sorted(recs, key=keyf), with_crc=False)     # sorted() is stable
        ok = got == want
        moved = self.merger.tiebreak_stats()
        res = {"ok": bool(ok and status == 0), "records": len(recs), "bytes": len(got), "tiebreak_groups": moved[0],
               "tiebreak_largest": moved[1], "wait_status": status}
        if self.world > 1:
            t = torch.tensor([1 if res["ok"] else 0, len(recs), len(got)], dtype=torch.int64, device=self.device)
            mn = t.clone()
            dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=self.group)
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            res.update(ok_all=bool(mn[0].item()), records_all=int(t[1].item()), bytes_all=int(t[2].item()))
        else:
            res.update(ok_all=res["ok"], records_all=len(recs), bytes_all=len(got))
        return res

    def close(self) -> None:
        self.arena.close()
        if self.side_index and hasattr(self, "side_arena"):
            self.side_arena.close()
