"""Reducer-side merge of landed runs into the merged KV stream (device resident).

Replaces the reference's MergeManager online merge: fetch phase inserts one Segment per map into
a priority queue, merge phase pops records into a 1 MiB staging buffer handed to Java
(src/Merger/MergeManager.cc:47-182, src/Merger/StreamRW.cc:151-238). Here the stages are kernels on
one stream: key normalise (K3) -> k-way merge of entries (K4) -> record gather into the output
stream (``VInt VInt key val`` ... ``VInt(-1) VInt(-1)``).

All buffers are allocated once (sized from the shuffle plan) and reused every step.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import kernels as K


class DeviceMerger:
    def __init__(self, max_records: int, max_runs: int, stride: int, device, out_bytes: Optional[int] = None,
                 fused: Optional[bool] = None, key_len: int = 10, ctas_per_sm: int = 0):
        self.device, self.stride, self.max_records, self.max_runs = device, stride, max_records, max_runs
        #: cap on resident merge CTAs per SM (0 = occupancy limit); lowered when a push kernel runs concurrently
        self.ctas_per_sm = ctas_per_sm
        # fused record merge (one read + one write per record) when the shape allows it, else the
        # entry pipeline (key normalise -> entry merge -> gather)
        self.fused = K.merge_records_supported(max_runs, max_records, stride, key_len) if fused is None else fused
        if self.fused:
            self.scratch = torch.empty(K.merge_records_scratch_bytes(max_records, max_runs), dtype=torch.uint8, device=device)
        else:
            self.entries = torch.empty((max(max_records, 1), 2), dtype=torch.int64, device=device)
            self.sorted = torch.empty_like(self.entries)
            self.scratch = torch.empty(K.merge_scratch_bytes(max_records, max_runs), dtype=torch.uint8, device=device)
        ob = (max_records * stride + 2 if out_bytes is None else out_bytes)
        self.out = torch.empty(ob + 64, dtype=torch.uint8, device=device)
        self.n = 0

    def snapshot_runs(self) -> dict:
        """The currently installed run tables (device tensors), for jobs that alternate between several."""
        return {k: getattr(self, k) for k in ("nruns", "n", "max_nrec", "run_descs_dev", "run_off_dev", "run_base_dev", "align2")}

    def use_runs(self, snap: dict) -> None:
        for k, v in snap.items():
            setattr(self, k, v)

    def set_runs(self, run_descs: np.ndarray, run_off: np.ndarray, run_base: np.ndarray) -> None:
        """Upload the (static for the job) run tables."""
        self.nruns = len(run_descs)
        assert self.nruns <= self.max_runs
        self.n = int(run_off[-1])
        assert self.n <= self.max_records
        self.max_nrec = int(run_descs["nrec"].max()) if self.nruns else 0
        self.run_descs_dev = K.to_device_struct(run_descs, self.device)
        self.run_off_dev = torch.from_numpy(run_off.astype(np.int64)).to(self.device)
        self.run_base_dev = torch.from_numpy(run_base.astype(np.uint64).view(np.int64)).to(self.device)
        self.align2 = bool(self.stride % 2 == 0 and not (np.asarray(run_base, dtype=np.uint64) & np.uint64(1)).any())

    def merge_fixed(self, key_off: int = 2, key_len: int = 10, write_eof: bool = True,
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Merge the runs described by :meth:`set_runs`; returns the merged stream (view of ``out`` /
        ``self.out``)."""
        n = self.n
        out = self.out if out is None else out
        if self.fused:
            K.merge_records_fixed(self.run_descs_dev, self.nruns, n, self.max_nrec, self.stride, key_off, key_len, out,
                                  self.scratch, write_eof=write_eof, align2=self.align2, ctas_per_sm=self.ctas_per_sm)
            return out[: n * self.stride + (2 if write_eof else 0)]
        if n:
            K.merge_runs_fixed(self.run_descs_dev, self.run_off_dev, self.nruns, n, self.max_nrec, self.stride, key_off, key_len,
                               self.entries, self.sorted, self.scratch)
        K.gather_fixed(self.sorted, n, self.run_base_dev, self.stride, out, write_eof)
        return out[: n * self.stride + (2 if write_eof else 0)]

    #: kernels launched by one fused merge_fixed() call (range, sample-hist, 3 scan, bkeys, partition, sizes, 3 scan, merge, eof)
    LAUNCHES_PER_MERGE = 13
