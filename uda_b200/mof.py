"""Map-output files (MOFs) resident in HBM, and the map-side task that produces them.

A MOF is exactly Hadoop's ``file.out`` + ``file.out.index``: IFile segments
(records, EOF marker, 4-byte CRC32) one per partition, and an index of
``(startOffset, rawLength, partLength)`` per partition. In the reference the provider keeps MOFs on
local disks and reads 1 MiB chunks with O_DIRECT AIO on demand
(src/MOFServer/IndexInfo.cc:237-335, src/CommUtils/AIOHandler.cc:80-150); here a MOF lives in HBM as a
``uint8`` tensor plus a host copy of its index, and "serving" a partition is a peer store of the
segment bytes (uda_b200/parallel/shuffle.py).

GPU-produced MOFs start every segment on a 16-byte boundary (the index makes gaps legal); MOFs
loaded from real Hadoop files keep their byte-exact layout.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np
import torch

from .ops import kernels as K
from .utils.ifile import IndexRecord, read_index, write_index

TERA_KEY = 10
TERA_VAL = 90
TERA_STRIDE = 102  # VInt(10) VInt(90) key val
SEG_TRAILER = 6    # EOF marker (2) + CRC32 (4)


@dataclass
class DeviceMOF:
    data: torch.Tensor                 # uint8, file.out bytes (+ slack)
    index: np.ndarray                  # int64 [P,3] startOffset, rawLength, partLength
    nrec: Optional[np.ndarray] = None  # int64 [P] records per partition when known
    path: str = ""                     # logical path (jobId/mapId), used by the text protocol
    meta: dict = field(default_factory=dict)

    @property
    def num_partitions(self) -> int:
        return int(self.index.shape[0])

    def index_records(self) -> List[IndexRecord]:
        return [IndexRecord(*(int(x) for x in row)) for row in self.index]

    def index_bytes(self) -> bytes:
        return write_index(self.index_records())

    def file_bytes(self) -> bytes:
        end = int(self.index[-1, 0] + self.index[-1, 2]) if len(self.index) else 0
        return bytes(self.data[:end].cpu().numpy())

    @staticmethod
    def from_host_bytes(file_out: bytes, file_out_index: bytes, device, path: str = "", codec: Optional[str] = None,
                        codec_block: int = 256 * 1024) -> "DeviceMOF":
        """A MOF written by Hadoop (byte-exact ``file.out`` + ``file.out.index``) into HBM.

        ``codec``: the job's ``mapred.map.output.compression.codec`` class name when the map outputs are compressed
        (Snappy or LZO). The segments are then expanded on the device (``ops/compress.py``) as they are loaded, and the
        resident MOF is the plain one the push and merge kernels expect: every segment on a 16-byte boundary,
        ``rawLength`` plain bytes (records + EOF marker) followed by a CRC32 computed here, ``partLength = rawLength + 4``.
        The reference expands compressed segments on the reducer, one block at a time on a decompress thread
        (src/Merger/DecompressorWrapper.cc:168-197). ``codec_block`` is the smallest block size the map side may have
        used (``io.compression.codec.{lzo,snappy}.buffersize``, 256 KiB by default): it sizes the block table.
        """
        idx = np.array([[r.start_offset, r.raw_length, r.part_length] for r in read_index(file_out_index)], dtype=np.int64)
        buf = torch.empty(K.GATHER_SLACK + len(file_out) + K.GATHER_SLACK + 16, dtype=torch.uint8, device=device)
        buf[K.GATHER_SLACK:K.GATHER_SLACK + len(file_out)] = torch.frombuffer(bytearray(file_out), dtype=torch.uint8).to(device)
        if codec is None or codec in ("", "null"):
            return DeviceMOF(buf[K.GATHER_SLACK:], idx, None, path)
        from .ops.compress import decoder_for

        comp = buf[K.GATHER_SLACK:]
        starts, off = [], 0
        for _, raw, _ in idx:
            off = (off + 15) // 16 * 16
            starts.append(off)
            off += int(raw) + 4
        out = torch.zeros(K.GATHER_SLACK + off + K.GATHER_SLACK + 16, dtype=torch.uint8, device=device)
        plain = out[K.GATHER_SLACK:]
        max_raw = int(idx[:, 1].max()) if len(idx) else 0
        dec = decoder_for(codec, device, max(1, -(-max_raw // max(1, codec_block)) + 1))
        for p, (start, raw, part) in enumerate(idx):
            start, raw, part = int(start), int(raw), int(part)
            seg = plain[starts[p]:]
            dec.decode(comp[start:], part, out=seg, raw_len=raw)
            dec.check(expect_raw=raw)                      # corrupt block, more blocks than planned, length mismatch
            K.crc32(seg, raw, out_be4=seg[raw:raw + 4])
        new_idx = np.array([[starts[p], int(idx[p, 1]), int(idx[p, 1]) + 4] for p in range(len(idx))], dtype=np.int64).reshape(-1, 3)
        return DeviceMOF(plain, new_idx, None, path, {"expanded_from": codec})


def uniform_splitters_w0(num_partitions: int) -> np.ndarray:
    """TotalOrderPartitioner splitters equally spaced over the first 8 key bytes (uint64)."""
    return np.array([((1 << 64) * i) // num_partitions for i in range(1, num_partitions)], dtype=np.uint64)


class MapTask:
    """Sort one map's records and write its MOF, all on the device (K3 -> K2 -> gather -> K7).

    Stands in for Hadoop's MapTask sort+spill, which the reference relies on but does not contain.
    Buffers are reusable across maps of the same size."""

    def __init__(self, max_records: int, num_partitions: int, device, stride: int = TERA_STRIDE):
        self.device, self.P, self.stride, self.cap = device, num_partitions, stride, max_records
        self.entries = torch.empty((max_records, 2), dtype=torch.int64, device=device)
        self.tmp = torch.empty_like(self.entries)
        self.sort_scratch = torch.empty(K.radix_sort_scratch_bytes(max_records), dtype=torch.uint8, device=device)

    def run(self, records: torch.Tensor, n: int, splitters_w0: np.ndarray, key_bytes: int = TERA_KEY,
            verify: bool = False) -> DeviceMOF:
        """``records``: ``n`` IFile-framed fixed-stride records (unsorted). Returns the sorted, partitioned MOF."""
        assert n <= self.cap
        dev = self.device
        desc, _ = K.make_run_descs([records.data_ptr()], [n])
        desc_dev = K.to_device_struct(desc, dev)
        ent = self.entries[:n]
        K.extract_keys_fixed(desc_dev, 1, n, self.stride, 2, key_bytes, ent)
        K.radix_sort_entries(ent, key_bytes, self.tmp[:n], self.sort_scratch)
        # partition boundaries: first entry with w0 >= splitter (splitters have a zero 2-byte tail)
        w0_signed = ent[:, 0] ^ torch.tensor(-(1 << 63), dtype=torch.int64, device=dev)
        sp = torch.from_numpy((splitters_w0 ^ np.uint64(1 << 63)).view(np.int64)).to(dev)
        bounds = torch.searchsorted(w0_signed.contiguous(), sp, right=False).cpu().numpy() if len(splitters_w0) else np.zeros(0, np.int64)
        bounds = np.concatenate([[0], bounds, [n]]).astype(np.int64)
        nrec = np.diff(bounds)
        # file.out layout: 16-byte aligned segment starts
        index = np.zeros((self.P, 3), dtype=np.int64)
        cur = 0
        for p in range(self.P):
            body = int(nrec[p]) * self.stride
            index[p] = (cur, body + 2, body + SEG_TRAILER)
            cur = (cur + body + SEG_TRAILER + 15) // 16 * 16
        buf = torch.empty(K.GATHER_SLACK + cur + K.GATHER_SLACK + 16, dtype=torch.uint8, device=dev)
        data = buf[K.GATHER_SLACK:]
        run_base = torch.tensor([records.data_ptr()], dtype=torch.int64, device=dev)
        for p in range(self.P):
            seg = data[int(index[p, 0]):]
            K.gather_fixed(ent[int(bounds[p]):int(bounds[p + 1])], int(nrec[p]), run_base, self.stride, seg, write_eof=True)
            raw = int(index[p, 1])
            K.crc32(seg, raw, out_be4=seg[raw:raw + 4])
        return DeviceMOF(data, index, nrec, meta={"sorted": True})
