"""Baseline arm named in BASELINE.json: NCCL ``all_to_all_single`` + ``torch.sort``.

Same data, same metric and same timed region as the engine's TeraSort step, built only from stock
PyTorch / NCCL calls (none of this repository's kernels are on the timed path; the set-up reuses the
engine's map side so that both arms start from identical sorted map outputs):

    step = all_to_all_single(partition-ordered map-output records)        # shuffle
         + key extraction + two stable torch.sort passes (2-byte tail, then 8-byte prefix)
         + index_select of the 102-byte records into the merged stream    # "merge"

This is "the baseline, not the product" (BASELINE.json north_star)."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from uda_b200.models.terasort import TeraSortJob
from uda_b200.mof import TERA_STRIDE
from uda_b200.ops import kernels as K


def pack_keys(recs: torch.Tensor):
    """10-byte keys of framed records ``[n, 102]`` -> (8-byte prefix as a sortable int64, 2-byte tail as int32).

    One byte-reversed copy of the key bytes viewed as integers (little-endian hosts): 10 bytes written per record,
    no per-byte widening. (The first version of this arm shifted and summed an ``[n, 8]`` int64 temporary, 64 bytes per
    record; the numbers in BASELINE.md were measured with that version.)"""
    w0 = recs[:, 2:10].flip(1).contiguous().view(torch.int64).view(-1) ^ (-(1 << 63))   # unsigned order as signed order
    tail = recs[:, 10:12].flip(1).contiguous().view(torch.int16).view(-1).to(torch.int32) & 0xFFFF
    return w0, tail


class NcclSortJob:
    launches_per_step = 0  # no kernels of this repository inside the timed region

    def __init__(self, records_per_rank: int, maps_per_rank: int, device: Optional[torch.device] = None, seed: int = 1234):
        self.inner = TeraSortJob(records_per_rank, maps_per_rank, device=device, seed=seed)
        self.device = self.inner.device
        self.world, self.rank = self.inner.world, self.inner.rank

    def setup(self) -> "NcclSortJob":
        j = self.inner.setup()
        dev, P = self.device, self.world
        # send buffer: records grouped by destination (partition-major, then map) — what a NCCL user would hand over
        chunks, counts = [], []
        for p in range(P):
            c = 0
            for m in j.mofs:
                start, raw, _ = (int(x) for x in m.index[p])
                n = (raw - 2) // TERA_STRIDE
                chunks.append(m.data[start:start + n * TERA_STRIDE])
                c += n
            counts.append(c)
        self.send = torch.cat(chunks).view(-1, TERA_STRIDE)
        self.send_counts = counts
        if P > 1:
            t = torch.tensor(counts, dtype=torch.int64, device=dev)
            r = torch.empty_like(t)
            dist.all_to_all_single(r, t)
            self.recv_counts = [int(x) for x in r.cpu()]
        else:
            self.recv_counts = counts
        self.n_out = sum(self.recv_counts)
        assert self.n_out == j.n_out
        self.recv = torch.empty((self.n_out, TERA_STRIDE), dtype=torch.uint8, device=dev)
        self.out = torch.empty(self.n_out * TERA_STRIDE + 2 + 64, dtype=torch.uint8, device=dev)
        # free what the baseline does not use
        j.mofs = []
        j.merger = None
        torch.cuda.synchronize(dev)
        return self

    def step(self) -> torch.Tensor:
        if self.world > 1:
            dist.all_to_all_single(self.recv, self.send, output_split_sizes=self.recv_counts, input_split_sizes=self.send_counts)
        else:
            self.recv.copy_(self.send)
        w0, tail = pack_keys(self.recv)
        i1 = torch.sort(tail, stable=True).indices
        i2 = torch.sort(w0[i1], stable=True).indices
        perm = i1[i2]
        n = self.n_out
        torch.index_select(self.recv, 0, perm, out=self.out[: n * TERA_STRIDE].view(n, TERA_STRIDE))
        self.out[n * TERA_STRIDE: n * TERA_STRIDE + 2] = 0xFF
        return self.out[: n * TERA_STRIDE + 2]

    def check(self, out: Optional[torch.Tensor] = None) -> dict:
        j = self.inner
        j.merger = type("M", (), {"out": self.out})()
        res = j.check(self.out)
        return res

    def close(self) -> None:
        self.inner.close()
