#!/usr/bin/env python
"""BASELINE.json config 5: all-to-all shuffle bandwidth sweep, 4 KB - 1 GB per peer, fused push kernel vs
NCCL ``all_to_all_single``. Launch with torchrun (one process per GPU); rank 0 prints one JSON line per size.

The reference's transport numbers (RDMA fetch of 1 MiB chunks, src/DataNet/RDMAServer.cc:537-631) cannot be
measured on this box (no verbs), so the comparison is against NCCL only.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch
import torch.distributed as dist

from uda_b200.ops import kernels as K
from uda_b200.parallel.arena import SymmetricArena


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="4096,65536,1048576,16777216,134217728,1073741824")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--ctas", default="0")
    a = ap.parse_args()
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    os.environ["NCCL_DEBUG"] = "WARN"
    dist.init_process_group("nccl", device_id=dev)
    P, rank = dist.get_world_size(), dist.get_rank()
    sizes = [int(s) for s in a.sizes.split(",")]
    smax = max(sizes)
    arena = SymmetricArena(P * smax + 4096, P, dev)
    send = torch.randint(0, 255, (P * smax,), dtype=torch.uint8, device=dev)
    recv_nccl = torch.empty_like(send)
    status = arena.status
    for S in sizes:
        d = np.zeros(P, dtype=K.PUSH_DESC_DTYPE)
        for i in range(P):
            p = (rank + 1 + i) % P
            d[i] = (send.data_ptr() + p * S, arena.peer_data_ptr(p) + rank * S, S, arena.peer_flag_ptr(p, rank))
        descs = K.to_device_struct(d, dev)
        scratch = torch.empty(K.push_scratch_bytes(P), dtype=torch.uint8, device=dev)
        expect = torch.zeros(P, dtype=torch.int64, device=dev)
        step = torch.full((P,), S, dtype=torch.int64, device=dev)
        out = {"bytes_per_peer": S, "world": P}
        for ctas in [int(c) for c in a.ctas.split(",")]:
            def push():
                arena.barrier()
                K.shuffle_push(descs, P, scratch, ctas, P)
                expect.add_(step)
                K.wait_flags(arena.flags(), expect, P, status)
            for _ in range(3):
                push()
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                push()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            out[f"push_ctas{ctas}_us"] = ms * 1e3
            out[f"push_ctas{ctas}_GBs_out_per_gpu"] = (P - 1) * S / (ms * 1e-3) / 1e9
        # verify once
        ok = bool(torch.equal(arena.data()[rank * S: rank * S + min(S, 4096)], send[rank * S: rank * S + min(S, 4096)]))
        out["self_slice_ok"] = ok
        sv, rv = send[: P * S], recv_nccl[: P * S]
        for _ in range(3):
            dist.all_to_all_single(rv, sv)
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            dist.all_to_all_single(rv, sv)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        out["nccl_us"] = ms * 1e3
        out["nccl_GBs_out_per_gpu"] = (P - 1) * S / (ms * 1e-3) / 1e9
        if rank == 0:
            print(json.dumps(out), flush=True)
    arena.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
