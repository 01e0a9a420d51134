"""Compressed map outputs: Hadoop block framing + Snappy or LZO1X, decoded on the device (K9).

Reference behaviour being replaced: ``DecompressorWrapper`` decorates the RDMA client with a decompress thread and a
cyclic plain-text buffer (src/Merger/DecompressorWrapper.cc:168-197); codecs are dlopen'ed (Snappy
src/Merger/SnappyDecompressor.cc:48-105, LZO src/Merger/LzoDecompressor.cc:66-167) and selected from
``mapred.map.output.compression.codec`` (plugins/shared/.../UdaPlugin.java:278-293). Here the compressed segment
lands in the receive arena unchanged (``partLength`` bytes) and is expanded to ``rawLength`` bytes by
``uda_snappy_scan_blocks`` (the framing does not depend on the codec) + ``uda_snappy_decode_blocks`` or
``uda_lzo_decode_blocks`` before the merge kernels see it. The LZO1X block decoder is the one the native runtime uses on
the CPU when liblzo2 cannot be loaded (csrc/runtime/lzo1x_parse.h, one source for both).
"""
from __future__ import annotations

import ctypes as C
import struct
from typing import Optional, Tuple

import torch

from .._lib import check, kernels, runtime
from . import kernels as K

DEFAULT_BLOCK = 256 * 1024   # io.compression.codec.snappy.buffersize default (UdaPlugin.java:285-293)


def snappy_compress_host(data: bytes) -> bytes:
    """Raw Snappy stream of ``data`` (greedy matcher of the native runtime; CPU, for tests and tools)."""
    L = runtime()
    L.uda_rt_snappy_compress.restype = C.c_longlong
    cap = 32 + len(data) + len(data) // 6
    out = C.create_string_buffer(cap)
    n = L.uda_rt_snappy_compress(data, C.c_longlong(len(data)), out, C.c_longlong(cap))
    if n < 0:
        raise RuntimeError("uda_rt_snappy_compress: output buffer too small")
    return out.raw[:n]


def lzo_compress_host(data: bytes) -> bytes:
    """LZO1X stream of ``data`` (greedy compressor of the native runtime; CPU, for tests and tools)."""
    L = runtime()
    L.uda_rt_lzo1x_compress.restype = C.c_longlong
    cap = len(data) + len(data) // 16 + 64 + 3
    out = C.create_string_buffer(cap)
    n = L.uda_rt_lzo1x_compress(data, C.c_longlong(len(data)), out, C.c_longlong(cap))
    if n < 0:
        raise RuntimeError("uda_rt_lzo1x_compress: output buffer too small")
    return out.raw[:n]


_HOST_COMPRESSORS = {"snappy": snappy_compress_host, "lzo": lzo_compress_host}


def frame_compress_segment(plain: bytes, block: int = DEFAULT_BLOCK, codec: str = "snappy") -> bytes:
    """``plain`` (an IFile segment body incl. EOF marker) as compressed blocks ``[rawLen][compLen][bytes]``."""
    compress = _HOST_COMPRESSORS[codec]
    out = []
    for i in range(0, len(plain), block):
        chunk = plain[i:i + block]
        comp = compress(chunk)
        out.append(struct.pack(">II", len(chunk), len(comp)) + comp)
    return b"".join(out)


class BlockDecoder:
    """Reusable device decoder for segments of up to ``max_blocks`` blocks of one codec (``"snappy"`` or ``"lzo"``)."""

    def __init__(self, device, max_blocks: int, codec: str = "snappy"):
        if codec not in ("snappy", "lzo"):
            raise ValueError(f"no device decoder for codec {codec!r}")
        self.device, self.max_blocks, self.codec = device, max_blocks, codec
        self.table = torch.empty(int(kernels().uda_snappy_table_bytes(max_blocks)), dtype=torch.uint8, device=device)
        self.out2 = torch.zeros(2, dtype=torch.int64, device=device)
        self.status = torch.zeros(1, dtype=torch.int32, device=device)

    def decode(self, seg: torch.Tensor, seg_len: int, out: Optional[torch.Tensor] = None, raw_len: Optional[int] = None) -> Tuple[torch.Tensor, int]:
        """Expands the compressed segment. ``raw_len`` (the index record's rawLength) avoids the host sync that is
        otherwise needed to size the output. Returns ``(plain bytes, length)``; raises on corrupt input when the
        status is checked (:meth:`check`)."""
        s = torch.cuda.current_stream(self.device).cuda_stream
        check(kernels().uda_snappy_scan_blocks(seg.data_ptr(), seg_len, self.max_blocks, self.table.data_ptr(), self.out2.data_ptr(),
                                               self.status.data_ptr(), s), "uda_snappy_scan_blocks")
        if raw_len is None:
            raw_len = int(self.out2[1].item())
        if out is None:
            out = torch.empty(raw_len + 64, dtype=torch.uint8, device=self.device)
        assert out.numel() >= raw_len
        name = "uda_snappy_decode_blocks" if self.codec == "snappy" else "uda_lzo_decode_blocks"
        check(getattr(kernels(), name)(seg.data_ptr(), self.table.data_ptr(), self.out2.data_ptr(), self.max_blocks,
                                       out.data_ptr(), self.status.data_ptr(), s), name)
        return out[:raw_len], raw_len

    def check(self, expect_raw: Optional[int] = None) -> None:
        """Synchronises; raises :class:`~uda_b200.runtime.failure.UdaFailure` when the last decode met corrupt data."""
        from ..runtime.failure import UdaFailure

        st = int(self.status.item())
        if st:
            self.status.zero_()
            raise UdaFailure({1: "compressed segment has more blocks than the decoder was sized for", 2: "corrupt compressed block header",
                              3: f"corrupt {self.codec} block"}.get(st, f"decoder status {st}"))
        if expect_raw is not None and int(self.out2[1].item()) != expect_raw:
            raise UdaFailure(f"decompressed length {int(self.out2[1].item())} != rawLength {expect_raw}")


class SnappyDecoder(BlockDecoder):
    def __init__(self, device, max_blocks: int):
        super().__init__(device, max_blocks, "snappy")


class LzoDecoder(BlockDecoder):
    """LZO1X blocks (reference: src/Merger/LzoDecompressor.cc:121-167, there through a dlopen'ed liblzo2)."""

    def __init__(self, device, max_blocks: int):
        super().__init__(device, max_blocks, "lzo")


def decoder_for(codec_class: str, device, max_blocks: int) -> BlockDecoder:
    """Device decoder for a ``mapred.map.output.compression.codec`` class name (plugins/shared/.../UdaPlugin.java:278-293)."""
    c = codec_class.lower()
    if "lzo" in c:
        return LzoDecoder(device, max_blocks)
    if "snappy" in c:
        return SnappyDecoder(device, max_blocks)
    raise ValueError(f"no device decoder for codec class {codec_class!r}")
