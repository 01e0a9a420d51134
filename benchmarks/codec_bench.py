#!/usr/bin/env python
"""Device block decompression rate (K9): Snappy and LZO1X segments in Hadoop block framing, decoded by one warp per block.

    python benchmarks/codec_bench.py --mb 64 --block-kb 256

Prints one JSON line per codec: plain GB/s produced (device-timed, CUDA events, after warm-up), compression ratio, blocks.
The data is word-like text (the WordCount / Sort inputs the reference's install guide benchmarks with LZO on).
Reference: one decompress thread per reducer, one block at a time (src/Merger/DecompressorWrapper.cc:168-197).
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np


def make_text(nbytes: int, seed: int = 5) -> bytes:
    rng = np.random.default_rng(seed)
    vocab = [bytes(rng.integers(97, 123, size=int(rng.integers(2, 10)), dtype=np.uint8)) for _ in range(4096)]
    # Zipf-like word choice: frequent words repeat at short distances, as in natural text
    ids = np.minimum(rng.zipf(1.3, size=nbytes // 5), 4095)
    return b" ".join(vocab[i] for i in ids)[:nbytes]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=64)
    ap.add_argument("--block-kb", type=int, default=256)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--host-only", action="store_true", help="compress only (no GPU): prints ratios and host compressor rates")
    a = ap.parse_args()
    from uda_b200.ops.compress import frame_compress_segment
    plain = make_text(a.mb << 20)
    block = a.block_kb << 10
    segs = {}
    for codec in ("snappy", "lzo"):
        t0 = time.time()
        segs[codec] = frame_compress_segment(plain, block, codec=codec)
        dt = time.time() - t0
        if a.host_only:
            print(json.dumps({"codec": codec, "plain_mb": len(plain) / 2**20, "ratio": len(segs[codec]) / len(plain),
                              "host_compress_mb_s": len(plain) / 2**20 / dt}))
    if a.host_only:
        return 0
    import torch
    from uda_b200.ops.compress import BlockDecoder
    dev = torch.device("cuda", 0)
    nblocks = -(-len(plain) // block)
    want = torch.frombuffer(bytearray(plain), dtype=torch.uint8).to(dev)
    for codec, comp in segs.items():
        seg = torch.frombuffer(bytearray(comp + b"\0" * 16), dtype=torch.uint8).to(dev)
        dec = BlockDecoder(dev, nblocks, codec)
        out = torch.empty(len(plain) + 64, dtype=torch.uint8, device=dev)
        for _ in range(a.warmup):
            dec.decode(seg, len(comp), out=out, raw_len=len(plain))
        dec.check(expect_raw=len(plain))
        ok = bool(torch.equal(out[:len(plain)], want))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            dec.decode(seg, len(comp), out=out, raw_len=len(plain))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        print(json.dumps({"metric": "device block decompression, plain GB/s", "codec": codec, "value": len(plain) / ms / 1e6, "unit": "GB/s",
                          "ms": ms, "plain_mb": len(plain) / 2**20, "ratio": round(len(comp) / len(plain), 4), "blocks": nblocks,
                          "block_kb": a.block_kb, "correct": ok, "steps": a.steps, "warmup": a.warmup}), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
