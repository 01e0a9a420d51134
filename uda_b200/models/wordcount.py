"""Zipf-skewed WordCount (BASELINE.json config 3): exercises partition imbalance and the combiner.

Record = IFile framing of (Text word, IntWritable count): ``0A 04 | 09 d1..d9 | c3 c2 c1 c0`` — a 9-digit word
makes the serialized Text key 10 bytes, so the whole record is 16 bytes (fixed stride) and the fixed-stride
kernels apply; the ordering is the reference's Text comparator (skip the VInt length, memcmp:
src/Merger/CompareFunc.cc:80-84), which for equal-length words is a memcmp of the digits.

Step (timed)::

    count      uda_partition_count        exact records per destination (Hadoop HashPartitioner on the word)
    exchange   all_gather of the P counts (plumbing) -> per-source regions in every arena, on the device
    scatter    uda_partition_scatter      fused hash partition + peer stores (shared-memory staging per destination)
    barrier    uda_barrier_all            every rank's records have landed
    sort       key normalise + radix sort + gather of the received records
    combine    uda_combine_*              one (word, total) record per distinct word

The map outputs are *not* pre-partitioned here, which is what distinguishes this path from TeraSort's segment
push; the heaviest words make some reducers receive about twice the average.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from ..ops import kernels as K
from ..parallel.arena import SymmetricArena

STRIDE = 16
KEY_OFF, KEY_LEN, KEY_SKIP, VAL_OFF = 2, 10, 1, 12


def zipf_word_ids(n: int, vocab: int, s: float, seed: int, device) -> torch.Tensor:
    """``n`` word ids in [0, vocab) with P(id = k) proportional to (k+1)^-s (inverse-CDF sampling on the device)."""
    w = torch.arange(1, vocab + 1, dtype=torch.float64, device=device).pow_(-s)
    cdf = torch.cumsum(w, 0)
    cdf /= cdf[-1].clone()
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    u = torch.rand(n, generator=g, device=device, dtype=torch.float64)
    return torch.searchsorted(cdf, u).clamp_(max=vocab - 1)


def encode_records(ids: torch.Tensor, counts: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8[n,16] IFile records for word ids (9 ASCII digits) with big-endian int counts (default 1)."""
    n = ids.numel()
    out = torch.zeros((n, STRIDE), dtype=torch.uint8, device=ids.device)
    out[:, 0], out[:, 1], out[:, 2] = 10, 4, 9
    x = ids.clone()
    for d in range(9):
        out[:, 11 - d] = (x % 10 + 48).to(torch.uint8)
        x //= 10
    c = torch.ones(n, dtype=torch.int64, device=ids.device) if counts is None else counts.to(torch.int64)
    for b in range(4):
        out[:, 15 - b] = ((c >> (8 * b)) & 0xFF).to(torch.uint8)
    return out


def java_hash_of_word_ids(ids: np.ndarray) -> np.ndarray:
    """``WritableComparator.hashBytes`` of the 9 ASCII digits (int32 wrap-around), vectorised."""
    h = np.ones(ids.shape, dtype=np.int64)
    digits = [(ids // 10 ** (8 - d)) % 10 + 48 for d in range(9)]
    for dg in digits:
        h = (31 * h + dg) & 0xFFFFFFFF
    return h


class WordCountJob:
    def __init__(self, records_per_rank: int, vocab: int = 1_000_000, zipf_s: float = 1.1, device=None, seed: int = 99,
                 capacity_factor: float = 3.0):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.n, self.vocab, self.s, self.seed = records_per_rank, vocab, zipf_s, seed
        self.cap = int(records_per_rank * capacity_factor) + 4096

    def setup(self) -> "WordCountJob":
        dev, P = self.device, self.world
        self.ids = zipf_word_ids(self.n, self.vocab, self.s, self.seed + self.rank, dev)
        self.recs = encode_records(self.ids).reshape(-1)
        self.arena = SymmetricArena(self.cap * STRIDE + 4096, 1, dev)
        self.cursor = torch.zeros(P, dtype=torch.int64, device=dev)
        self.peer_base = torch.tensor([self.arena.peer_data_ptr(p) + K.GATHER_SLACK for p in range(P)], dtype=torch.int64, device=dev)
        self.entries = torch.empty((self.cap, 2), dtype=torch.int64, device=dev)
        self.tmp = torch.empty_like(self.entries)
        self.sort_scratch = torch.empty(K.radix_sort_scratch_bytes(self.cap), dtype=torch.uint8, device=dev)
        self.sorted = torch.empty(self.cap * STRIDE + 64, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize(dev)
        return self

    def step(self):
        dev, P = self.device, self.world
        counts = K.partition_count(self.recs, self.n, STRIDE, KEY_OFF, KEY_LEN, KEY_SKIP, 0, P)
        if P > 1:
            allc = torch.empty((P, P), dtype=torch.int64, device=dev)  # allc[s][p] = records rank s sends to p
            dist.all_gather_into_tensor(allc, counts)
        else:
            allc = counts.view(1, 1)
        before = torch.cumsum(allc, 0) - allc            # records placed in p's arena by lower ranks
        dst = self.peer_base + before[self.rank] * STRIDE
        n_recv_dev = allc[:, self.rank].sum()
        self.cursor.zero_()
        if P > 1:
            self.arena.barrier()                          # nobody is still reading the previous step's arena
        K.partition_scatter(self.recs, self.n, STRIDE, KEY_OFF, KEY_LEN, KEY_SKIP, 0, P, dst, self.cursor)
        if P > 1:
            self.arena.barrier()                          # everything addressed to this rank has landed
        n_recv = int(n_recv_dev.item())                   # the one host sync of the step (sizes the sort)
        if n_recv > self.cap:
            raise RuntimeError(f"receive arena too small for this skew: {n_recv} > {self.cap} records")
        data = self.arena.data()[K.GATHER_SLACK:]
        desc, _ = K.make_run_descs([data.data_ptr()], [n_recv])
        ent = self.entries[:n_recv]
        K.extract_keys_fixed(K.to_device_struct(desc, dev), 1, n_recv, STRIDE, KEY_OFF, KEY_LEN, ent)
        K.radix_sort_entries(ent, KEY_LEN, self.tmp[:n_recv], self.sort_scratch)
        run_base = torch.tensor([data.data_ptr()], dtype=torch.int64, device=dev)
        K.gather_fixed(ent, n_recv, run_base, STRIDE, self.sorted, write_eof=False)
        self.out, self.groups = K.combine_sum_fixed(self.sorted, n_recv, STRIDE, KEY_OFF, KEY_LEN, VAL_OFF)
        self.n_recv = n_recv
        return self.out

    def check(self) -> dict:
        """Every (word, total) of this reducer against a bincount of all ranks' generated ids."""
        dev, P = self.device, self.world
        hist = torch.bincount(self.ids, minlength=self.vocab)
        if P > 1:
            dist.all_reduce(hist)
        hist = hist.cpu().numpy()
        words = np.nonzero(hist)[0]
        mine = words[(java_hash_of_word_ids(words) & 0x7FFFFFFF) % P == self.rank]
        got = self.out.cpu().numpy().reshape(-1, STRIDE)
        ok_shape = got.shape[0] == len(mine)
        got_ids = np.zeros(got.shape[0], dtype=np.int64)
        for d in range(9):
            got_ids = got_ids * 10 + (got[:, 3 + d].astype(np.int64) - 48)
        got_cnt = (got[:, 12].astype(np.int64) << 24) | (got[:, 13].astype(np.int64) << 16) | (got[:, 14].astype(np.int64) << 8) | got[:, 15]
        # 9 equal-width digits: numeric order == Text order
        ok = ok_shape and bool((got_ids == np.sort(mine)).all()) and bool((got_cnt == hist[np.sort(mine)]).all())
        imbalance = float(self.n_recv) * P / max(1, self.n * P)
        return {"ok": ok, "distinct_words": int(len(mine)), "received": int(self.n_recv), "imbalance_vs_mean": imbalance}

    def close(self):
        self.arena.close()
