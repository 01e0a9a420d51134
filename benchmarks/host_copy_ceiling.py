#!/usr/bin/env python
"""Ceiling of the end-to-end path: how fast can pinned host memory be copied to and from the GPUs of this box?

The end-to-end TeraSort step moves every step's map outputs host -> device and the merged stream device -> host
(12.75 GB each way per GPU). Whatever the kernels do, it cannot run faster than those copies; this program measures
them alone: one process per GPU (torchrun for several), every rank copies `--gib` GiB each way through pinned buffers
with one cudaMemcpyAsync per 256 MiB piece, H2D and D2H on separate streams — first one direction at a time, then both
at once, which is what the end-to-end step does. Prints one JSON line: per-GPU and aggregate GB/s (device-timed, max
over ranks). `e2e_ceiling_GBps` is the record throughput (100 of every 102 bytes) the step would reach if it were
nothing but these copies.

    python benchmarks/host_copy_ceiling.py
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/host_copy_ceiling.py
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=11.875, help="GiB copied per direction per GPU and repetition (12.75 GB)")
    ap.add_argument("--piece-mib", type=int, default=256)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    from uda_b200.utils.affinity import bind_to_gpu_numa
    aff = bind_to_gpu_numa(lr)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    piece = a.piece_mib << 20
    npieces = max(1, int(a.gib * (1 << 30)) // piece)
    ring = 4
    h_in = [torch.empty(piece, dtype=torch.uint8, pin_memory=True) for _ in range(ring)]
    h_out = [torch.empty(piece, dtype=torch.uint8, pin_memory=True) for _ in range(ring)]
    for t in h_in:
        t.fill_(1)
    d_in = torch.empty(piece * ring, dtype=torch.uint8, device=dev)
    d_out = torch.ones(piece * ring, dtype=torch.uint8, device=dev)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def run(h2d: bool, d2h: bool) -> float:
        best = 1e30
        for _ in range(a.reps + 1):
            barrier()
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            s_in.wait_event(e0)
            s_out.wait_event(e0)
            for i in range(npieces):
                k = i % ring
                if h2d:
                    with torch.cuda.stream(s_in):
                        d_in[k * piece:(k + 1) * piece].copy_(h_in[k], non_blocking=True)
                if d2h:
                    with torch.cuda.stream(s_out):
                        h_out[k].copy_(d_out[k * piece:(k + 1) * piece], non_blocking=True)
            e1.record(s_in)
            e2.record(s_out)
            torch.cuda.current_stream().wait_event(e1)
            torch.cuda.current_stream().wait_event(e2)
            e3 = torch.cuda.Event(enable_timing=True)
            e3.record()
            torch.cuda.synchronize(dev)
            ms = e0.elapsed_time(e3)
            if world > 1:
                t = torch.tensor([ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            best = min(best, ms)
        return best

    nbytes = npieces * piece
    ms_h2d, ms_d2h, ms_both = run(True, False), run(False, True), run(True, True)
    if rank == 0:
        gb = nbytes / 1e9
        out = {
            "n_gpus": world, "bytes_per_direction_per_gpu": nbytes, "piece_bytes": piece,
            "h2d_alone_GBps_per_gpu": gb / ms_h2d * 1e3, "d2h_alone_GBps_per_gpu": gb / ms_d2h * 1e3,
            "both_GBps_per_direction_per_gpu": gb / ms_both * 1e3,
            "both_GBps_per_direction_aggregate": gb * world / ms_both * 1e3,
            # the step copies 12.75 GB in and 12.75 GB out for 12.5 GB of records per GPU
            "e2e_ceiling_GBps": (gb * world / ms_both * 1e3) * (100.0 / 102.0),
            "cpu_affinity_rank0": aff, "host_cpus": os.cpu_count(),
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
