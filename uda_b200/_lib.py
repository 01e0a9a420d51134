"""ctypes loader for the in-tree native libraries.

``libuda_kernels.so`` (sm_100a kernels, C ABI in include/uda_b200.h) and ``libuda.so`` (C++ runtime +
JNI bridge; same name as the reference's library, src/Makefile.am:32-59). Both are built in-tree by
tools/build.py so they travel to the GPU box with the repo snapshot.

There is deliberately no PyTorch/eager fallback for the kernels: if the library is missing the ops
raise, so a GPU test can never pass on a silent fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_DIR = Path(__file__).resolve().parent / "lib"
_kernels = None
_runtime = None

c_ll = C.c_longlong
c_ull = C.c_ulonglong
c_vp = C.c_void_p
c_int = C.c_int
c_uint = C.c_uint


class UdaRunDesc(C.Structure):
    _fields_ = [("base", c_vp), ("nrec", c_ll), ("out_off", c_ll), ("run_id", c_uint), ("idx0", c_uint)]


class UdaPushDesc(C.Structure):
    _fields_ = [("src", c_vp), ("dst", c_vp), ("bytes", c_ll), ("flag", c_vp)]


_PROTOS = {
    "uda_teragen": (c_int, [c_vp, c_ll, c_ull, c_ull, c_int, c_int, c_vp]),
    "uda_validate_fixed": (c_int, [c_vp, c_ll, c_int, c_int, c_vp, c_vp]),
    "uda_extract_keys_fixed": (c_int, [c_vp, c_int, c_ll, c_int, c_int, c_int, c_vp, c_vp]),
    "uda_extract_keys_var": (c_int, [c_vp, c_vp, c_ll, c_int, c_uint, c_uint, c_vp, c_vp, c_vp]),
    "uda_radix_sort_scratch_bytes": (c_ll, [c_ll]),
    "uda_radix_sort_entries": (c_int, [c_vp, c_vp, c_ll, c_int, c_vp, c_vp]),
    "uda_merge_scratch_bytes": (c_ll, [c_ll, c_int]),
    "uda_kway_merge_entries": (c_int, [c_vp, c_vp, c_int, c_ll, c_vp, c_vp, c_vp]),
    "uda_merge_runs_fixed": (c_int, [c_vp, c_vp, c_int, c_ll, c_ll, c_int, c_int, c_int, c_vp, c_vp, c_vp, c_vp]),
    "uda_merge_records_scratch_bytes": (c_ll, [c_ll, c_int]),
    "uda_merge_records_supported": (c_int, [c_int, c_ll, c_int, c_int]),
    "uda_merge_records_stats3": (c_int, [c_vp, c_vp, c_vp]),
    "uda_merge_records_tile": (c_int, [c_int, c_int, c_vp, c_vp]),
    "uda_merge_records_fixed": (c_int, [c_vp, c_int, c_ll, c_ll, c_int, c_int, c_int, c_int, c_vp, c_int, c_vp, c_int, c_vp]),
    "uda_sample_splitters": (c_int, [c_vp, c_ll, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    "uda_lower_bound_records": (c_int, [c_vp, c_ll, c_int, c_int, c_int, c_vp, c_int, c_vp, c_vp]),
    "uda_merge_records_stats": (c_int, [c_vp, C.POINTER(c_ull), c_vp]),
    "uda_merge_stats": (c_int, [c_vp, C.POINTER(c_ull), c_vp]),
    "uda_gather_fixed": (c_int, [c_vp, c_ll, c_vp, c_int, c_vp, c_int, c_vp]),
    "uda_gather_var_lengths": (c_int, [c_vp, c_ll, c_vp, c_vp, c_vp]),
    "uda_gather_var": (c_int, [c_vp, c_ll, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "uda_tiebreak_scratch_bytes": (c_ll, [c_ll]),
    "uda_tiebreak_var": (c_int, [c_vp, c_ll, c_vp, c_vp, c_int, c_vp, c_vp, c_vp]),
    "uda_tiebreak_stats": (c_int, [c_vp, c_ll, C.POINTER(c_ull), c_vp]),
    "uda_scan_u32_to_u64_scratch_bytes": (c_ll, [c_ll]),
    "uda_scan_u32_to_u64": (c_int, [c_vp, c_ll, c_vp, c_vp, c_vp]),
    "uda_write_eof_at": (c_int, [c_vp, c_vp, c_vp]),
    "uda_snappy_table_bytes": (c_ll, [c_int]),
    "uda_snappy_scan_blocks": (c_int, [c_vp, c_ll, c_int, c_vp, c_vp, c_vp, c_vp]),
    "uda_snappy_decode_blocks": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp]),
    "uda_lzo_decode_blocks": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp]),
    "uda_push_scratch_bytes": (c_ll, [c_int]),
    "uda_shuffle_push": (c_int, [c_vp, c_int, c_vp, c_int, c_vp]),
    "uda_shuffle_push_nway": (c_int, [c_vp, c_int, c_vp, c_int, c_int, c_vp]),
    "uda_push_tma_scratch_bytes": (c_ll, [c_int]),
    "uda_shuffle_push_tma": (c_int, [c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_vp]),
    "uda_shuffle_push_tma_ring": (c_int, [c_vp, c_int, c_int, c_int, c_vp, c_int, c_int, c_vp, c_int, c_ll, c_ull, c_vp, c_vp]),
    "uda_credit_publish": (c_int, [c_vp, c_int, c_vp, c_vp]),
    "uda_wait_flags": (c_int, [c_vp, c_vp, c_int, c_ull, c_vp, c_vp]),
    "uda_barrier_all": (c_int, [c_vp, c_int, c_int, c_ull, c_ull, c_vp, c_vp]),
    "uda_barrier_all_counted": (c_int, [c_vp, c_int, c_int, c_vp, c_ull, c_vp, c_vp]),
    "uda_partition_count": (c_int, [c_vp, c_ll, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp]),
    "uda_partition_scatter": (c_int, [c_vp, c_ll, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp, c_vp]),
    "uda_combine_heads": (c_int, [c_vp, c_ll, c_int, c_int, c_int, c_vp, c_vp]),
    "uda_combine_emit": (c_int, [c_vp, c_ll, c_ll, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp, c_vp]),
    "uda_scan_u32_scratch_bytes": (c_ll, [c_ll]),
    "uda_scan_u32": (c_int, [c_vp, c_ll, c_vp, c_vp, c_vp]),
    "uda_multicast_store": (c_int, [c_vp, c_ull, c_ll, c_vp]),
    "uda_unicast_store": (c_int, [c_vp, c_vp, c_int, c_ll, c_vp]),
    "uda_crc32": (c_int, [c_vp, c_ll, c_vp, c_vp, c_vp, c_vp]),
    "uda_ifile_frame_fixed": (c_int, [c_vp, c_ll, c_int, c_int, c_vp, c_vp]),
    "uda_ifile_unframe_fixed": (c_int, [c_vp, c_ll, c_int, c_int, c_vp, c_vp]),
    "uda_ifile_decode_scratch_bytes": (c_ll, [c_ll]),
    "uda_ifile_decode_offsets": (c_int, [c_vp, c_ll, c_vp, c_vp, c_vp, c_vp]),
    "uda_ifile_encoded_sizes": (c_int, [c_vp, c_vp, c_ll, c_vp, c_vp]),
    "uda_dev_malloc": (c_int, [C.POINTER(c_vp), c_ll]),
    "uda_dev_free": (c_int, [c_vp]),
    "uda_ipc_handle_bytes": (c_int, []),
    "uda_ipc_get_handle": (c_int, [c_vp, c_vp]),
    "uda_ipc_open_handle": (c_int, [c_vp, C.POINTER(c_vp)]),
    "uda_ipc_close_handle": (c_int, [c_vp]),
    "uda_can_access_peer": (c_int, [c_int, c_int, C.POINTER(c_int)]),
    "uda_memset_async": (c_int, [c_vp, c_int, c_ll, c_vp]),
    "uda_host_alloc_pinned": (c_int, [C.POINTER(c_vp), c_ll]),
    "uda_host_free_pinned": (c_int, [c_vp]),
    "uda_ifile_encode_var": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_ll, c_vp, c_vp, c_vp]),
}


class NativeLibraryMissing(RuntimeError):
    pass


def lib_path(name: str) -> Path:
    return _LIB_DIR / name


def _bind(lib, protos, strict=True):
    for name, (res, args) in protos.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if strict:
                raise
            continue
        fn.restype = res
        fn.argtypes = args


def kernels() -> C.CDLL:
    """The CUDA kernel library. Raises loudly when it has not been built."""
    global _kernels
    if _kernels is None:
        p = Path(os.environ["UDA_B200_KERNELS_LIB"]) if os.environ.get("UDA_B200_KERNELS_LIB") else lib_path("libuda_kernels.so")
        if not p.exists():
            raise NativeLibraryMissing(
                f"{p} not found: run `python tools/build.py` (or __graft_entry__.build()) first; "
                "uda_b200 has no non-native fallback for its kernels"
            )
        lib = C.CDLL(str(p), mode=os.RTLD_NOW | getattr(os, "RTLD_LOCAL", 0))
        _bind(lib, _PROTOS, strict=False)
        _kernels = lib
    return _kernels


def runtime() -> C.CDLL:
    """The C++ runtime / JNI bridge library (CPU only, no CUDA dependency)."""
    global _runtime
    if _runtime is None:
        p = Path(os.environ["UDA_B200_RUNTIME_LIB"]) if os.environ.get("UDA_B200_RUNTIME_LIB") else lib_path("libuda.so")
        if not p.exists():
            raise NativeLibraryMissing(f"{p} not found: run `python tools/build.py` first")
        _runtime = C.CDLL(str(p), mode=os.RTLD_NOW)
    return _runtime


def check(rc: int, what: str) -> None:
    if rc != 0:
        if rc > 0:
            raise RuntimeError(f"{what}: CUDA error {rc}")
        raise ValueError(f"{what}: invalid argument (uda error {rc})")
