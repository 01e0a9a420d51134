#!/usr/bin/env python
"""Hadoop "sort" example shape on the general path: variable-length records, BytesWritable / Text keys, a share of
keys that tie on the 10-byte sort prefix. One process per GPU (torchrun for several). Prints one JSON line:
shuffle + merge GB/s of serialized record bytes, device-timed, max over ranks."""
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
    ap.add_argument("--records", type=int, default=20_000_000, help="records per GPU")
    ap.add_argument("--maps", type=int, default=8)
    ap.add_argument("--key-kind", type=int, default=2, help="1 Text, 2 BytesWritable")
    ap.add_argument("--key-len", default="10,100")
    ap.add_argument("--val-len", default="0,200")
    ap.add_argument("--tie-fraction", type=float, default=0.001)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--no-side-index", action="store_true", help="do not ship record offsets: the reducer rediscovers the boundaries (K5)")
    ap.add_argument("--check", action="store_true", help="byte-exact CPU oracle (small inputs only)")
    a = ap.parse_args()
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from uda_b200.models.sortjob import SortJob

    kl = tuple(int(x) for x in a.key_len.split(","))
    vl = tuple(int(x) for x in a.val_len.split(","))
    job = SortJob(a.records, a.maps, device=dev, key_kind=a.key_kind, key_len=kl, val_len=vl, tie_fraction=a.tie_fraction, side_index=not a.no_side_index).setup()
    for _ in range(a.warmup):
        out = job.step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = job.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    t = torch.tensor([ms, float(job.bytes_out), float(job.n_out)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms = float(mx[0])
    total_bytes, total_rec = float(t[1]), int(t[2])
    groups, largest, moved = job.merger.tiebreak_stats()
    res = {"metric": "variable-length sort shuffle+merge GB/s (device-timed, max over ranks)", "value": total_bytes / ms / 1e6, "unit": "GB/s",
           "ms_per_step": ms, "n_gpus": world, "records": total_rec, "bytes": int(total_bytes), "tiebreak_groups": groups,
           "tiebreak_largest_group": largest, "gpu_launches": job.launches_per_step * a.steps,
           "config": {"records_per_gpu": a.records, "maps_per_gpu": a.maps, "key_kind": a.key_kind, "key_len": kl, "val_len": vl,
                      "tie_fraction": a.tie_fraction, "side_index": not a.no_side_index}}
    if a.check:
        c = job.check(out)
        res["ok"] = res["validated"] = bool(c["ok_all"])
    if (dist.get_rank() if world > 1 else 0) == 0:
        print(json.dumps(res), flush=True)
    job.close()
    if world > 1:
        dist.destroy_process_group()
    return 0 if res.get("ok", True) else 1


if __name__ == "__main__":
    sys.exit(main())
