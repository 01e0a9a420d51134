"""General path for IFile segments with variable-length records and any supported key class.

Map side (:class:`VarMapTask`): record boundary discovery -> key normalise -> radix sort on the 80-bit prefix ->
exact-comparator tie-break -> per-partition gather into a MOF (segments, EOF markers, CRC32, index).
Reduce side (:class:`VarMerger`): per landed segment boundary discovery + key normalise, k-way merge of the
entries, tie-break, 64-bit output offsets, record gather, EOF marker.

Reference behaviour being replaced: the consumer's segment reader (``nextKVInternal`` with its buffer-straddle
join, src/Merger/StreamRW.cc:334-404,592-662), the comparator selected from the Java key class
(src/Merger/CompareFunc.cc:95-113), the heap merge (src/Merger/MergeQueue.h:250-269,390-419) and the serialiser of
the merged stream (src/Merger/StreamRW.cc:151-225). The TeraSort fast path (fixed stride, 10-byte keys) lives in
:mod:`uda_b200.ops.merger`; this module is what every other job goes through.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import kernels as K
from ..mof import DeviceMOF, SEG_TRAILER

#: key_kind values of the kernels (csrc/kernels/keys.cu); the class-name table is csrc/runtime/compare.h
KEY_RAW, KEY_TEXT, KEY_BYTES_WRITABLE = 0, 1, 2

_KEY_CLASS_KIND = {
    "org.apache.hadoop.io.Text": KEY_TEXT,
    "org.apache.hadoop.io.BytesWritable": KEY_BYTES_WRITABLE,
    "org.apache.hadoop.hbase.io.ImmutableBytesWritable": KEY_BYTES_WRITABLE,
    "org.apache.hadoop.io.BooleanWritable": KEY_RAW,
    "org.apache.hadoop.io.ByteWritable": KEY_RAW,
    "org.apache.hadoop.io.ShortWritable": KEY_RAW,
    "org.apache.hadoop.io.IntWritable": KEY_RAW,
    "org.apache.hadoop.io.LongWritable": KEY_RAW,
}


def key_kind_of(java_class: str) -> int:
    """Comparator selection by Java key class (src/Merger/CompareFunc.cc:95-113). Unknown classes raise, which the
    plug-in turns into a fallback to the vanilla shuffle."""
    try:
        return _KEY_CLASS_KIND[java_class]
    except KeyError:
        raise ValueError(f"unsupported key class {java_class!r}: no comparator") from None


def _ptr_table(ptrs: Sequence[int], device) -> torch.Tensor:
    return torch.from_numpy(np.asarray(ptrs, dtype=np.uint64).view(np.int64).copy()).to(device)


class VarMerger:
    """Reduce-side merge of ``k`` sorted variable-length segments resident on the device.

    All buffers are allocated once for ``max_records`` / ``max_runs`` / ``max_bytes`` and reused. With exact record
    counts the call never synchronises with the host."""

    def __init__(self, device, key_kind: int, max_records: int, max_runs: int, max_bytes: int):
        self.device, self.key_kind = device, key_kind
        self.max_records, self.max_runs, self.max_bytes = max_records, max_runs, max_bytes
        n = max(max_records, 1)
        self.entries = torch.empty((n, 2), dtype=torch.int64, device=device)
        self.sorted = torch.empty_like(self.entries)
        self.merge_scratch = torch.empty(K.merge_scratch_bytes(n, max(max_runs, 1)), dtype=torch.uint8, device=device)
        self.tb_scratch = torch.empty(K.tiebreak_scratch_bytes(n), dtype=torch.uint8, device=device)
        self.rec_len = torch.empty(n, dtype=torch.int32, device=device)
        self.dst_off = torch.empty(n + 1, dtype=torch.int64, device=device)
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)
        self.out = torch.empty(max_bytes + 64, dtype=torch.uint8, device=device)
        self.n = 0
        self._keep: List[torch.Tensor] = []

    def merge(self, segments: Sequence[Tuple[torch.Tensor, int, int]], exact_counts: bool = True,
              rec_offs: Optional[Sequence[torch.Tensor]] = None) -> Tuple[torch.Tensor, int]:
        """``segments[r] = (uint8 device tensor starting at the segment, seg_len incl. the EOF marker, records)``.
        ``records`` is the exact count (``exact_counts``) or an upper bound (one host sync to learn the counts).
        ``rec_offs[r]`` (int64, records + 1 entries) is the record-offset side index of segment r when the map side
        shipped one; without it the boundaries are rediscovered from the bytes (K5).
        Returns ``(merged stream incl. EOF marker, its length in bytes)``."""
        k, dev = len(segments), self.device
        assert k <= self.max_runs
        nrec_dev = []
        if rec_offs is None:
            rec_offs = []
            for seg, seg_len, cap in segments:
                ro, nr = K.ifile_decode_offsets(seg, seg_len, cap)
                rec_offs.append(ro)
                nrec_dev.append(nr)
        else:
            assert exact_counts and len(rec_offs) == k, "a side index implies exact record counts"
            rec_offs = list(rec_offs)
        if exact_counts:
            counts = [int(c) for _, _, c in segments]
        else:
            counts = [int(x) for x in torch.cat(nrec_dev).cpu().tolist()] if k else []
        run_off = np.zeros(k + 1, dtype=np.int64)
        np.cumsum(counts, out=run_off[1:])
        n = int(run_off[-1])
        assert n <= self.max_records, f"{n} records exceed the merger's capacity {self.max_records}"
        total = int(sum(sl - 2 for _, sl, _ in segments)) + 2
        assert total <= self.max_bytes + 2
        self.n, self.nrec_dev = n, nrec_dev
        self.flag.zero_()
        for r, (seg, _sl, _c) in enumerate(segments):
            if counts[r]:
                K.extract_keys_var(seg, rec_offs[r], counts[r], self.key_kind, r, 0, self.entries[int(run_off[r]):], self.flag)
        run_base = _ptr_table([s.data_ptr() for s, _, _ in segments], dev)
        run_recoff = _ptr_table([ro.data_ptr() for ro in rec_offs], dev)
        self._keep = [run_base, run_recoff, *rec_offs]        # the kernels below read them asynchronously
        if n:
            K.kway_merge_entries(self.entries, torch.from_numpy(run_off).to(dev), k, n, self.sorted, self.merge_scratch)
            K.tiebreak_var(self.sorted, n, run_base, run_recoff, self.key_kind, self.entries, self.tb_scratch)
            K.gather_var_lengths(self.sorted, n, run_recoff, self.rec_len)
            K.scan_u32_to_u64(self.rec_len, n, out=self.dst_off)
            K.gather_var(self.sorted, n, run_base, run_recoff, self.dst_off, self.out)
            K.write_eof_at(self.out, self.dst_off[n:])
        else:
            K.write_eof_at(self.out, None)
        return self.out[:total], total

    def tiebreak_stats(self) -> Tuple[int, int, int]:
        return K.tiebreak_stats(self.tb_scratch, self.n)

    #: kernels per merge of k runs: k x (decode 5 [absent with a side index] + extract 1) + merge 9 + tie-break 7 +
    #: lengths 1 + scan 3 + gather 1 + eof 1
    @staticmethod
    def launches(k: int, side_index: bool = False) -> int:
        return (1 if side_index else 6) * k + 22


class VarMapTask:
    """Sort one map's variable-length records on the device and write its MOF (K5 -> K3 -> K2 -> tie-break -> K6/K7).

    Stands in for Hadoop's MapTask sort + spill for jobs other than TeraSort. Partitioning is by key range on the
    first 8 compared key bytes (TotalOrderPartitioner with binary splitters), so that the concatenation of the
    reducers' outputs is globally sorted."""

    def __init__(self, device, key_kind: int, num_partitions: int):
        self.device, self.key_kind, self.P = device, key_kind, num_partitions

    def run(self, seg: torch.Tensor, seg_len: int, n: int, splitters_w0: np.ndarray) -> DeviceMOF:
        """``seg``: one unsorted IFile segment (records + EOF marker) of exactly ``n`` records."""
        dev, P = self.device, self.P
        rec_off, _ = K.ifile_decode_offsets(seg, seg_len, n)
        ent = torch.empty((max(n, 1), 2), dtype=torch.int64, device=dev)
        tmp = torch.empty_like(ent)
        flag = torch.zeros(1, dtype=torch.int32, device=dev)
        run_base, run_recoff = _ptr_table([seg.data_ptr()], dev), _ptr_table([rec_off.data_ptr()], dev)
        if n:
            K.extract_keys_var(seg, rec_off, n, self.key_kind, 0, 0, ent, flag)
            K.radix_sort_entries(ent[:n], 10, tmp[:n])
            K.tiebreak_var(ent, n, run_base, run_recoff, self.key_kind, tmp, torch.empty(K.tiebreak_scratch_bytes(n), dtype=torch.uint8, device=dev))
        w0_signed = (ent[:n, 0] ^ torch.tensor(-(1 << 63), dtype=torch.int64, device=dev)).contiguous()
        sp = torch.from_numpy((np.asarray(splitters_w0, dtype=np.uint64) ^ np.uint64(1 << 63)).view(np.int64)).to(dev)
        bounds = torch.searchsorted(w0_signed, sp, right=False).cpu().numpy() if len(splitters_w0) else np.zeros(0, np.int64)
        bounds = np.concatenate([[0], bounds, [n]]).astype(np.int64)
        nrec = np.diff(bounds)
        rec_len = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        if n:
            K.gather_var_lengths(ent, n, run_recoff, rec_len)
        off64 = K.scan_u32_to_u64(rec_len, n)
        cuts = off64[torch.from_numpy(bounds).to(dev)].cpu().numpy()      # byte position of every partition boundary
        index = np.zeros((P, 3), dtype=np.int64)
        cur = 0
        for p in range(P):
            body = int(cuts[p + 1] - cuts[p])
            index[p] = (cur, body + 2, body + SEG_TRAILER)
            cur = (cur + body + SEG_TRAILER + 15) // 16 * 16
        buf = torch.empty(K.GATHER_SLACK + cur + K.GATHER_SLACK + 16, dtype=torch.uint8, device=dev)
        data = buf[K.GATHER_SLACK:]
        # record-offset side index: per partition the n_p + 1 offsets of its records relative to the segment start
        # (the last one is the EOF marker), int64, partitions 16-byte aligned. A reducer that receives it next to
        # the segment does not have to rediscover the record boundaries (K5).
        side_index = np.zeros((P, 3), dtype=np.int64)
        scur = 0
        for p in range(P):
            nb = (int(nrec[p]) + 1) * 8
            side_index[p] = (scur, nb, nb)
            scur = (scur + nb + 15) // 16 * 16
        side = torch.zeros(scur + 16, dtype=torch.uint8, device=dev)
        for p in range(P):
            b0, b1 = int(bounds[p]), int(bounds[p + 1])
            out = data[int(index[p, 0]):]
            rel = (off64[b0:b1 + 1] - off64[b0]).contiguous()
            if b1 > b0:
                K.gather_var(ent[b0:b1], b1 - b0, run_base, run_recoff, rel, out)
            side[int(side_index[p, 0]):int(side_index[p, 0]) + (b1 - b0 + 1) * 8].view(torch.int64).copy_(rel)
            raw = int(index[p, 1])
            K.write_eof_at(out[raw - 2:], None)
            K.crc32(out, raw, out_be4=out[raw:raw + 4])
        torch.cuda.current_stream(dev).synchronize()                      # the temporaries above die with this frame
        return DeviceMOF(data, index, nrec, meta={"sorted": True, "key_kind": self.key_kind, "side": side, "side_index": side_index})
