#!/usr/bin/env python
"""BASELINE.json config 3: Zipf-skewed (s = 1.1) key shuffle with a WordCount-style combiner.
`python benchmarks/wordcount_bench.py [--records-per-gpu N]` (torchrun for several GPUs). 64 GB over 8 GPUs is
500 M 16-byte records per GPU; the default is that size. Prints one JSON line on rank 0."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--records-per-gpu", type=int, default=500_000_000)
    ap.add_argument("--vocab", type=int, default=1_000_000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    from uda_b200.models.wordcount import STRIDE, WordCountJob

    job = WordCountJob(a.records_per_gpu, vocab=a.vocab, device=dev).setup()
    for _ in range(a.warmup):
        job.step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        job.step()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = job.check()
    imb = torch.tensor([res["imbalance_vs_mean"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(imb, op=dist.ReduceOp.MAX)
    if (dist.get_rank() if world > 1 else 0) == 0:
        ms = float(t.item())
        print(json.dumps({"metric": "Zipf WordCount shuffle+sort+combine GB/s (device-timed, max over ranks)",
                          "value": a.records_per_gpu * STRIDE * world / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms,
                          "n_gpus": world, "records_per_gpu": a.records_per_gpu, "vocab": a.vocab, "zipf_s": 1.1,
                          "max_reducer_load_vs_mean": float(imb.item()), "validated": res["ok"]}), flush=True)
    job.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
