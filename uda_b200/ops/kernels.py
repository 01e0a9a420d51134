"""Thin torch-tensor wrappers over the C ABI in include/uda_b200.h.

All functions launch on the *current* torch CUDA stream and never synchronise unless their docstring
says so. Record buffers are ``torch.uint8`` device tensors; sort entries are ``int64[n, 2]`` tensors
holding the two unsigned words of ``UdaEntry``.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from .._lib import check, kernels

RUN_DESC_DTYPE = np.dtype([("base", "<u8"), ("nrec", "<i8"), ("out_off", "<i8"), ("run_id", "<u4"), ("idx0", "<u4")])
PUSH_DESC_DTYPE = np.dtype([("src", "<u8"), ("dst", "<u8"), ("bytes", "<i8"), ("flag", "<u8")])
ENTRY_BYTES = 16
#: readable slack the gather needs before/after every source buffer (see uda_gather_fixed)
GATHER_SLACK = 32


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def to_device_struct(arr: np.ndarray, device) -> torch.Tensor:
    """Upload a NumPy structured array as a uint8 device tensor."""
    raw = torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy())
    return raw.to(device, non_blocking=False)


# ------------------------------------------------------------------------------------------------ data
def teragen(n: int, seed: int, first_row: int = 0, framed: bool = True, key_bits: int = 0,
            out: Optional[torch.Tensor] = None, device="cuda") -> torch.Tensor:
    """Generate ``n`` teragen-shaped records (102 B IFile-framed or 100 B raw) on the device."""
    stride = 102 if framed else 100
    if out is None:
        out = torch.empty(n * stride + 64, dtype=torch.uint8, device=device)
    check(kernels().uda_teragen(out.data_ptr(), n, seed, first_row, int(framed), key_bits, _stream()), "uda_teragen")
    return out


def validate_fixed(recs: torch.Tensor, n: int, stride: int, key_off: int) -> Tuple[int, int, int]:
    """TeraValidate: returns ``(misordered_pairs, fingerprint_sum, count)``. Synchronises."""
    res = torch.zeros(3, dtype=torch.int64, device=recs.device)
    check(kernels().uda_validate_fixed(recs.data_ptr(), n, stride, key_off, res.data_ptr(), _stream()), "uda_validate_fixed")
    r = res.cpu().numpy().view(np.uint64)
    return int(r[0]), int(r[1]), int(r[2])


# ------------------------------------------------------------------------------------------------ keys
def make_run_descs(bases: Sequence[int], nrecs: Sequence[int], run_ids: Optional[Sequence[int]] = None,
                   idx0: Optional[Sequence[int]] = None) -> Tuple[np.ndarray, np.ndarray]:
    """Host-side run table + entry offsets (``run_off[nruns+1]``)."""
    k = len(bases)
    d = np.zeros(k, dtype=RUN_DESC_DTYPE)
    d["base"] = np.asarray(bases, dtype=np.uint64)
    d["nrec"] = np.asarray(nrecs, dtype=np.int64)
    off = np.zeros(k + 1, dtype=np.int64)
    np.cumsum(d["nrec"], out=off[1:])
    d["out_off"] = off[:-1]
    d["run_id"] = np.arange(k, dtype=np.uint32) if run_ids is None else np.asarray(run_ids, dtype=np.uint32)
    d["idx0"] = 0 if idx0 is None else np.asarray(idx0, dtype=np.uint32)
    return d, off


def extract_keys_fixed(run_descs_dev: torch.Tensor, nruns: int, max_nrec: int, stride: int, key_off: int, key_len: int,
                       entries_out: torch.Tensor) -> None:
    check(kernels().uda_extract_keys_fixed(run_descs_dev.data_ptr(), nruns, max_nrec, stride, key_off, key_len,
                                           entries_out.data_ptr(), _stream()), "uda_extract_keys_fixed")


def extract_keys_var(data: torch.Tensor, rec_off: torch.Tensor, n: int, key_kind: int, run_id: int, idx0: int,
                     entries_out: torch.Tensor, long_key_flag: torch.Tensor) -> None:
    check(kernels().uda_extract_keys_var(data.data_ptr(), rec_off.data_ptr(), n, key_kind, run_id, idx0,
                                         entries_out.data_ptr(), long_key_flag.data_ptr(), _stream()), "uda_extract_keys_var")


# ------------------------------------------------------------------------------------------------ sort / merge
def radix_sort_scratch_bytes(n: int) -> int:
    return int(kernels().uda_radix_sort_scratch_bytes(n))


def radix_sort_entries(entries: torch.Tensor, key_bytes: int = 10, tmp: Optional[torch.Tensor] = None,
                       scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """In-place stable LSD radix sort of ``int64[n,2]`` entries by the top ``key_bytes`` key bytes."""
    n = entries.shape[0]
    if tmp is None:
        tmp = torch.empty_like(entries)
    if scratch is None:
        scratch = torch.empty(radix_sort_scratch_bytes(n), dtype=torch.uint8, device=entries.device)
    check(kernels().uda_radix_sort_entries(entries.data_ptr(), tmp.data_ptr(), n, key_bytes, scratch.data_ptr(), _stream()),
          "uda_radix_sort_entries")
    return entries


def merge_scratch_bytes(n: int, nruns: int) -> int:
    r = int(kernels().uda_merge_scratch_bytes(n, nruns))
    if r < 0:
        raise ValueError("uda_merge_scratch_bytes: invalid argument")
    return r


def kway_merge_entries(entries: torch.Tensor, run_off_dev: torch.Tensor, nruns: int, n: int, sorted_out: torch.Tensor,
                       scratch: torch.Tensor) -> None:
    check(kernels().uda_kway_merge_entries(entries.data_ptr(), run_off_dev.data_ptr(), nruns, n, sorted_out.data_ptr(),
                                           scratch.data_ptr(), _stream()), "uda_kway_merge_entries")


def merge_runs_fixed(run_descs_dev: torch.Tensor, run_off_dev: torch.Tensor, nruns: int, n: int, max_nrec: int, stride: int,
                     key_off: int, key_len: int, entries_tmp: torch.Tensor, sorted_out: torch.Tensor, scratch: torch.Tensor) -> None:
    """Fused key normalise + histogram + k-way merge for fixed-stride runs."""
    check(kernels().uda_merge_runs_fixed(run_descs_dev.data_ptr(), run_off_dev.data_ptr(), nruns, n, max_nrec, stride, key_off,
                                         key_len, entries_tmp.data_ptr(), sorted_out.data_ptr(), scratch.data_ptr(), _stream()),
          "uda_merge_runs_fixed")


def merge_records_supported(nruns: int, n: int, stride: int, key_len: int) -> bool:
    return bool(kernels().uda_merge_records_supported(nruns, n, stride, key_len))


def merge_records_scratch_bytes(n: int, nruns: int) -> int:
    return int(kernels().uda_merge_records_scratch_bytes(n, nruns))


def merge_records_fixed(run_descs_dev: torch.Tensor, nruns: int, n: int, max_nrec: int, stride: int, key_off: int, key_len: int,
                        out: torch.Tensor, scratch: torch.Tensor, write_eof: bool = True, align2: bool = True,
                        ctas_per_sm: int = 0) -> None:
    """Fused fixed-stride k-way merge: every record is read once and written once. ``ctas_per_sm`` caps the resident
    merge CTAs per SM (0 = occupancy limit)."""
    check(kernels().uda_merge_records_fixed(run_descs_dev.data_ptr(), nruns, n, max_nrec, stride, key_off, key_len, int(align2),
                                            out.data_ptr(), int(write_eof), scratch.data_ptr(), ctas_per_sm, _stream()),
          "uda_merge_records_fixed")


def merge_records_stats(scratch: torch.Tensor) -> Tuple[int, int]:
    out = (C.c_ulonglong * 2)()
    check(kernels().uda_merge_records_stats(scratch.data_ptr(), out, _stream()), "uda_merge_records_stats")
    return int(out[0]), int(out[1])


def merge_records_stats3(scratch: torch.Tensor) -> Tuple[int, int, int]:
    """(oversize buckets, bisection iterations, tiles ordered by the merge tree) of the last fused merge."""
    out = (C.c_ulonglong * 3)()
    check(kernels().uda_merge_records_stats3(scratch.data_ptr(), out, _stream()), "uda_merge_records_stats3")
    return int(out[0]), int(out[1]), int(out[2])


def merge_records_tile(nruns: int, stride: int) -> Tuple[int, int]:
    """(tile capacity in records, shared memory per CTA) of the fused merge for this shape."""
    cap, smem = C.c_int(0), C.c_longlong(0)
    check(kernels().uda_merge_records_tile(nruns, stride, C.byref(cap), C.byref(smem)), "uda_merge_records_tile")
    return cap.value, smem.value


def merge_stats(scratch: torch.Tensor) -> Tuple[int, int]:
    out = (C.c_ulonglong * 2)()
    check(kernels().uda_merge_stats(scratch.data_ptr(), out, _stream()), "uda_merge_stats")
    return int(out[0]), int(out[1])


def gather_fixed(sorted_entries: torch.Tensor, n: int, run_base_dev: torch.Tensor, stride: int, dst: torch.Tensor,
                 write_eof: bool = True) -> None:
    check(kernels().uda_gather_fixed(sorted_entries.data_ptr(), n, run_base_dev.data_ptr(), stride, dst.data_ptr(),
                                     int(write_eof), _stream()), "uda_gather_fixed")


def gather_var_lengths(sorted_entries: torch.Tensor, n: int, run_recoff_dev: torch.Tensor, rec_len_out: torch.Tensor) -> None:
    check(kernels().uda_gather_var_lengths(sorted_entries.data_ptr(), n, run_recoff_dev.data_ptr(), rec_len_out.data_ptr(),
                                           _stream()), "uda_gather_var_lengths")


def gather_var(sorted_entries: torch.Tensor, n: int, run_base_dev: torch.Tensor, run_recoff_dev: torch.Tensor,
               dst_off_dev: torch.Tensor, dst: torch.Tensor) -> None:
    check(kernels().uda_gather_var(sorted_entries.data_ptr(), n, run_base_dev.data_ptr(), run_recoff_dev.data_ptr(),
                                   dst_off_dev.data_ptr(), dst.data_ptr(), _stream()), "uda_gather_var")


def tiebreak_scratch_bytes(n: int) -> int:
    return int(kernels().uda_tiebreak_scratch_bytes(n))


def tiebreak_var(sorted_entries: torch.Tensor, n: int, run_base_dev: torch.Tensor, run_recoff_dev: torch.Tensor, key_kind: int,
                 tmp_entries: torch.Tensor, scratch: torch.Tensor) -> None:
    """Exact-comparator re-ordering of equal-prefix groups (in place). No sync."""
    check(kernels().uda_tiebreak_var(sorted_entries.data_ptr(), n, run_base_dev.data_ptr(), run_recoff_dev.data_ptr(), key_kind,
                                     tmp_entries.data_ptr(), scratch.data_ptr(), _stream()), "uda_tiebreak_var")


def tiebreak_stats(scratch: torch.Tensor, n: int) -> Tuple[int, int, int]:
    """(groups re-ordered, largest such group, entries moved) of the last :func:`tiebreak_var`. Synchronises."""
    out = (C.c_ulonglong * 3)()
    check(kernels().uda_tiebreak_stats(scratch.data_ptr(), n, out, _stream()), "uda_tiebreak_stats")
    return int(out[0]), int(out[1]), int(out[2])


def scan_u32_to_u64(v: torch.Tensor, m: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Exclusive 64-bit prefix of an int32/uint32 tensor: ``m+1`` int64 entries, ``out[m]`` = total. No sync."""
    m = v.numel() if m is None else m
    if out is None:
        out = torch.empty(m + 1, dtype=torch.int64, device=v.device)
    scratch = torch.empty(int(kernels().uda_scan_u32_to_u64_scratch_bytes(m)), dtype=torch.uint8, device=v.device)
    check(kernels().uda_scan_u32_to_u64(v.data_ptr(), m, out.data_ptr(), scratch.data_ptr(), _stream()), "uda_scan_u32_to_u64")
    return out


def write_eof_at(dst: torch.Tensor, total_dev: Optional[torch.Tensor]) -> None:
    check(kernels().uda_write_eof_at(dst.data_ptr(), _p(total_dev), _stream()), "uda_write_eof_at")


# ------------------------------------------------------------------------------------------------ shuffle
def push_scratch_bytes(nseg: int) -> int:
    return int(kernels().uda_push_scratch_bytes(nseg))


def shuffle_push(descs_dev: torch.Tensor, nseg: int, scratch: torch.Tensor, max_ctas: int = 0, nway: int = 1) -> None:
    """LSU push. ``nway`` = number of destination ranks the chunk order is interleaved over (1 = descriptor order)."""
    check(kernels().uda_shuffle_push_nway(descs_dev.data_ptr(), nseg, scratch.data_ptr(), max_ctas, nway, _stream()),
          "uda_shuffle_push_nway")


def push_tma_scratch_bytes(nseg: int) -> int:
    return int(kernels().uda_push_tma_scratch_bytes(nseg))


def shuffle_push_tma(descs_dev: torch.Tensor, nseg: int, nsub: int, nway: int, scratch: torch.Tensor, num_ctas: int = 0,
                     flush_chunks: int = 0) -> None:
    """Copy-engine push: ``nsub`` equally long descriptor tables, concatenated, pushed in order by ``num_ctas``
    single-warp CTAs that each own an SM (TMA bulk load -> shared-memory ring -> TMA bulk store to the peer)."""
    check(kernels().uda_shuffle_push_tma(descs_dev.data_ptr(), nseg, nsub, nway, scratch.data_ptr(), num_ctas, flush_chunks,
                                         _stream()), "uda_shuffle_push_tma")


def shuffle_push_tma_ring(descs_dev: torch.Tensor, nseg: int, nsub: int, nway: int, scratch: torch.Tensor, num_ctas: int, flush_chunks: int,
                          credits: torch.Tensor, world: int, need0: int, status: torch.Tensor, max_spins: int = 50_000_000) -> None:
    """The copy-engine push into a receive RING: before a warp touches sub-partition ``j`` it waits until every
    ``credits[p] >= need0 + j`` (reducer p has merged the sub-partition that used the slot before)."""
    check(kernels().uda_shuffle_push_tma_ring(descs_dev.data_ptr(), nseg, nsub, nway, scratch.data_ptr(), num_ctas, flush_chunks,
                                              credits.data_ptr(), world, need0, max_spins, status.data_ptr(), _stream()),
          "uda_shuffle_push_tma_ring")


def credit_publish(credit_slots_dev: torch.Tensor, world: int, counter_dev: torch.Tensor) -> None:
    check(kernels().uda_credit_publish(credit_slots_dev.data_ptr(), world, counter_dev.data_ptr(), _stream()), "uda_credit_publish")


def wait_flags(flags: torch.Tensor, expect: torch.Tensor, n: int, status: torch.Tensor, max_spins: int = 50_000_000) -> None:
    check(kernels().uda_wait_flags(flags.data_ptr(), expect.data_ptr(), n, max_spins, status.data_ptr(), _stream()),
          "uda_wait_flags")


def barrier_all(pads_dev: torch.Tensor, rank: int, world: int, epoch: int, status: torch.Tensor,
                max_spins: int = 50_000_000) -> None:
    check(kernels().uda_barrier_all(pads_dev.data_ptr(), rank, world, epoch, max_spins, status.data_ptr(), _stream()),
          "uda_barrier_all")


def barrier_all_counted(pads_dev: torch.Tensor, rank: int, world: int, counter: torch.Tensor, status: torch.Tensor,
                        max_spins: int = 50_000_000) -> None:
    """All-rank barrier whose epoch lives in ``counter`` (int64[1], this rank's memory): graph-capturable."""
    check(kernels().uda_barrier_all_counted(pads_dev.data_ptr(), rank, world, counter.data_ptr(), max_spins, status.data_ptr(), _stream()),
          "uda_barrier_all_counted")


# ------------------------------------------------------------------------------------------------ crc / ifile
def crc32(data: torch.Tensor, n: Optional[int] = None, out_be4: Optional[torch.Tensor] = None) -> torch.Tensor:
    """CRC32 (zlib) of the first ``n`` bytes; returns a 1-element int32 device tensor (no sync)."""
    n = data.numel() if n is None else n
    acc = torch.empty(2, dtype=torch.int32, device=data.device)
    check(kernels().uda_crc32(data.data_ptr(), n, acc[0:1].data_ptr(), acc[1:2].data_ptr(), _p(out_be4), _stream()), "uda_crc32")
    return acc[1:2]


def ifile_frame_fixed(raw: torch.Tensor, n: int, klen: int, vlen: int, out: torch.Tensor) -> None:
    check(kernels().uda_ifile_frame_fixed(raw.data_ptr(), n, klen, vlen, out.data_ptr(), _stream()), "uda_ifile_frame_fixed")


def ifile_unframe_fixed(framed: torch.Tensor, n: int, klen: int, vlen: int, raw: torch.Tensor) -> None:
    check(kernels().uda_ifile_unframe_fixed(framed.data_ptr(), n, klen, vlen, raw.data_ptr(), _stream()), "uda_ifile_unframe_fixed")


def ifile_decode_offsets(seg: torch.Tensor, seg_len: int, max_records: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Record boundary discovery. Returns ``(rec_off int64[max_records+1], nrec int64[1])`` device tensors;
    ``rec_off[nrec]`` is the offset of the EOF marker."""
    scratch = torch.empty(int(kernels().uda_ifile_decode_scratch_bytes(seg_len)), dtype=torch.uint8, device=seg.device)
    rec_off = torch.empty(max_records + 1, dtype=torch.int64, device=seg.device)
    nrec = torch.zeros(1, dtype=torch.int64, device=seg.device)
    check(kernels().uda_ifile_decode_offsets(seg.data_ptr(), seg_len, rec_off.data_ptr(), nrec.data_ptr(), scratch.data_ptr(),
                                             _stream()), "uda_ifile_decode_offsets")
    return rec_off, nrec


def ifile_encode_var(keys: torch.Tensor, key_off: torch.Tensor, klen: torch.Tensor, vals: torch.Tensor, val_off: torch.Tensor,
                     vlen: torch.Tensor) -> torch.Tensor:
    """Encode ``n`` (key,val) pairs as IFile records + EOF marker. Synchronises once (output size)."""
    n = klen.numel()
    sizes = torch.empty(n, dtype=torch.int32, device=keys.device)
    check(kernels().uda_ifile_encoded_sizes(klen.data_ptr(), vlen.data_ptr(), n, sizes.data_ptr(), _stream()),
          "uda_ifile_encoded_sizes")
    off = scan_u32_to_u64(sizes, n)
    total = int(off[-1].item())
    out = torch.empty(total + 2 + 16, dtype=torch.uint8, device=keys.device)
    check(kernels().uda_ifile_encode_var(keys.data_ptr(), key_off.data_ptr(), klen.data_ptr(), vals.data_ptr(),
                                         val_off.data_ptr(), vlen.data_ptr(), n, off.data_ptr(), out.data_ptr(), _stream()),
          "uda_ifile_encode_var")
    return out[: total + 2]


# ------------------------------------------------------------------------------------------------ partition / combine
def scan_u32(v: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Exclusive scan of an int32 tensor; returns ``m+1`` entries (``out[m]`` = total). No sync."""
    m = v.numel()
    if out is None:
        out = torch.empty(m + 1, dtype=torch.int32, device=v.device)
    scratch = torch.empty(int(kernels().uda_scan_u32_scratch_bytes(m)), dtype=torch.uint8, device=v.device)
    check(kernels().uda_scan_u32(v.data_ptr(), m, out.data_ptr(), scratch.data_ptr(), _stream()), "uda_scan_u32")
    return out


def partition_count(recs: torch.Tensor, n: int, stride: int, key_off: int, key_len: int, key_skip: int, mode: int, P: int,
                    splitters: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Records per destination (int64[P] device tensor, no sync). mode 0 = Hadoop hash, 1 = range."""
    counts = torch.empty(P, dtype=torch.int64, device=recs.device)
    check(kernels().uda_partition_count(recs.data_ptr(), n, stride, key_off, key_len, key_skip, mode, P, _p(splitters),
                                        counts.data_ptr(), _stream()), "uda_partition_count")
    return counts


def partition_scatter(recs: torch.Tensor, n: int, stride: int, key_off: int, key_len: int, key_skip: int, mode: int, P: int,
                      dst_ptrs: torch.Tensor, cursor: torch.Tensor, splitters: Optional[torch.Tensor] = None) -> None:
    check(kernels().uda_partition_scatter(recs.data_ptr(), n, stride, key_off, key_len, key_skip, mode, P, _p(splitters),
                                          dst_ptrs.data_ptr(), cursor.data_ptr(), _stream()), "uda_partition_scatter")


def combine_sum_fixed(recs: torch.Tensor, n: int, stride: int, key_off: int, key_len: int, val_off: int) -> Tuple[torch.Tensor, int]:
    """WordCount combiner over a key-sorted fixed-stride stream. Returns ``(combined records, groups)``; one sync
    (the number of distinct keys sizes the output)."""
    dev = recs.device
    head = torch.empty(n + 1, dtype=torch.int32, device=dev)
    check(kernels().uda_combine_heads(recs.data_ptr(), n, stride, key_off, key_len, head.data_ptr(), _stream()), "uda_combine_heads")
    gid = scan_u32(head[:n], out=head)
    groups = int(gid[n].item()) if n else 0
    out = torch.empty(max(groups, 1) * stride + 16, dtype=torch.uint8, device=dev)
    sums = torch.empty(max(groups, 1), dtype=torch.int64, device=dev)
    if n:
        check(kernels().uda_combine_emit(recs.data_ptr(), n, groups, stride, key_off, key_len, val_off, gid.data_ptr(), out.data_ptr(),
                                         sums.data_ptr(), _stream()), "uda_combine_emit")
    return out[: groups * stride], groups


def sample_splitters(run: torch.Tensor, n: int, stride: int, key_off: int, key_len: int, nkeys: int) -> torch.Tensor:
    """``nkeys`` equi-depth key prefixes (int64 view of uint64) of a sorted fixed-stride run."""
    out = torch.empty(max(nkeys, 1), dtype=torch.int64, device=run.device)
    check(kernels().uda_sample_splitters(run.data_ptr(), n, stride, key_off, key_len, nkeys, out.data_ptr(), _stream()), "uda_sample_splitters")
    return out[:nkeys]


def lower_bound_records(run: torch.Tensor, n: int, stride: int, key_off: int, key_len: int, keys: torch.Tensor) -> torch.Tensor:
    """First record index with key >= each probe key (int64 device tensor, no sync)."""
    out = torch.empty(max(keys.numel(), 1), dtype=torch.int64, device=run.device)
    check(kernels().uda_lower_bound_records(run.data_ptr(), n, stride, key_off, key_len, keys.data_ptr(), keys.numel(), out.data_ptr(),
                                            _stream()), "uda_lower_bound_records")
    return out[: keys.numel()]
