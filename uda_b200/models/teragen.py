"""NumPy restatement of the device teragen / fingerprint (csrc/kernels/teragen.cu), bit identical.

Used as the oracle for the generator and as the CPU-side data source for plumbing tests (the
reference drives Hadoop's teragen instead: scripts/regression/sortcountRunner.sh:123-241)."""
from __future__ import annotations

import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def mix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
        return z ^ (z >> np.uint64(31))


def teragen_raw(n: int, seed: int, first_row: int = 0, key_bits: int = 0) -> np.ndarray:
    """``uint8[n,100]`` raw records (10 B key + 90 B value)."""
    with np.errstate(over="ignore"):
        r = np.arange(first_row, first_row + n, dtype=np.uint64)
        a = mix64(np.uint64(seed) + np.uint64(2) * r)
        b = mix64(np.uint64(seed) + np.uint64(2) * r + np.uint64(1))
    tail_zero = 0 < key_bits < 64
    if tail_zero:
        a = a & np.uint64(((1 << 64) - 1) ^ ((1 << (64 - key_bits)) - 1))
    out = np.zeros((n, 100), dtype=np.uint8)
    for j in range(8):
        out[:, j] = (a >> np.uint64(8 * (7 - j))).astype(np.uint8)
    for j in (8, 9):
        out[:, j] = 0 if tail_zero else (b >> np.uint64(8 * (7 - (j - 8)))).astype(np.uint8)
    for j in range(10, 18):
        out[:, j] = (r >> np.uint64(8 * (7 - (j - 10)))).astype(np.uint8)
    for j in range(18, 100):
        out[:, j] = ((b >> np.uint64(8 * (j & 7))).astype(np.uint64) + np.uint64(j)).astype(np.uint8)
    return out


def teragen_framed(n: int, seed: int, first_row: int = 0, key_bits: int = 0) -> np.ndarray:
    """``uint8[n,102]`` IFile-framed records: VInt(10)=0x0A VInt(90)=0x5A key val."""
    raw = teragen_raw(n, seed, first_row, key_bits)
    out = np.empty((n, 102), dtype=np.uint8)
    out[:, 0] = 10
    out[:, 1] = 90
    out[:, 2:] = raw
    return out


def _load_be(rec: np.ndarray, off: int, nbytes: int) -> np.ndarray:
    v = np.zeros(rec.shape[0], dtype=np.uint64)
    for i in range(nbytes):
        v = (v << np.uint64(8)) | rec[:, off + i].astype(np.uint64)
    return v


def fingerprint_sum(raw: np.ndarray) -> int:
    """Order-independent fingerprint of ``uint8[n,100]`` raw records (sum of per-record hashes mod 2^64)."""
    h = np.zeros(raw.shape[0], dtype=np.uint64)
    for j in range(0, 96, 8):
        h = mix64(h ^ _load_be(raw, j, 8))
    h = mix64(h ^ _load_be(raw, 96, 4))
    with np.errstate(over="ignore"):
        return int(h.sum(dtype=np.uint64))
