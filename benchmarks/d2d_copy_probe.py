#!/usr/bin/env python
"""Same-device copy rates: torch copy_ (cudaMemcpyAsync D2D = copy engines) for several piece sizes and alignments, next to
the library's LSU push kernel on all SMs. Decides whether a rank's own map outputs should reach its arena through the copy
engines or through SMs. One JSON line per case."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from uda_b200.ops import kernels as K


def timed(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = torch.device("cuda", 0)
    total = 6 << 30
    src = torch.empty(total + 4096, dtype=torch.uint8, device=dev)
    dst = torch.empty(total + 4096, dtype=torch.uint8, device=dev)
    for piece, off in ((64 << 20, 0), (64 << 20, 2), (8 << 20, 0), (1 << 20, 0), (total, 0)):
        n = total // piece

        def fn():
            for i in range(n):
                dst[off + i * piece: off + (i + 1) * piece].copy_(src[off + i * piece: off + (i + 1) * piece], non_blocking=True)
        ms = timed(fn)
        print(json.dumps({"case": "copy_engine", "piece_bytes": piece, "phase": off, "pieces": n, "ms": ms, "GBps_copied": total / ms / 1e6}))
    # the LSU push kernel on all SMs, one descriptor per 64 MiB
    piece = 64 << 20
    n = total // piece
    d = np.zeros(n, dtype=K.PUSH_DESC_DTYPE)
    flags = torch.zeros(n, dtype=torch.int64, device=dev)
    for i in range(n):
        d[i] = (src.data_ptr() + i * piece, dst.data_ptr() + i * piece, piece, flags.data_ptr() + 8 * i)
    dd = K.to_device_struct(d, dev)
    scratch = torch.empty(K.push_scratch_bytes(n), dtype=torch.uint8, device=dev)
    ms = timed(lambda: K.shuffle_push(dd, n, scratch, 0, 1))
    print(json.dumps({"case": "lsu_push_all_sms", "piece_bytes": piece, "pieces": n, "ms": ms, "GBps_copied": total / ms / 1e6}))


if __name__ == "__main__":
    main()
