#!/usr/bin/env python
"""BASELINE.json config 4 (scaled to this box): TeraSort merge of data that exceeds the HBM *budget*, with pinned
host-DRAM spill on side streams (the AsyncIO path). The full 1 TB configuration needs 125 GB of pinned staging
per GPU for the runs plus the same again for the spilled groups, which the 8-GPU box (1007 GB of RAM for 8 ranks)
cannot hold; the budget is therefore forced below the data size instead (default: 12.75 GB of runs, 6 GB budget),
which exercises exactly the same code path: grouped device merges, D2H spill overlapped with the next group's
merge, final key-range merge streaming back from host memory.  Prints one JSON line."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--records", type=int, default=125_000_000)
    ap.add_argument("--runs", type=int, default=32)
    ap.add_argument("--budget-gb", type=float, default=6.0)
    ap.add_argument("--steps", type=int, default=2)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    from uda_b200.mof import TERA_STRIDE, MapTask, uniform_splitters_w0
    from uda_b200.ops import kernels as K
    from uda_b200.ops.spill import SpillMerger

    per = a.records // a.runs
    task = MapTask(per, 1, dev)
    host_runs, nrecs, fp = [], [], 0
    for r in range(a.runs):
        recs = K.teragen(per, 4242, first_row=r * per, framed=True, device=dev)
        _b, f, _c = K.validate_fixed(recs, per, TERA_STRIDE, 2)
        fp = (fp + f) & 0xFFFFFFFFFFFFFFFF
        mof = task.run(recs, per, uniform_splitters_w0(1))
        h = torch.empty(per * TERA_STRIDE, dtype=torch.uint8, pin_memory=True)
        h.copy_(mof.data[: per * TERA_STRIDE])
        host_runs.append(h)
        nrecs.append(per)
        del recs, mof
    del task
    torch.cuda.empty_cache()
    total = sum(nrecs)
    ring = [torch.empty(256 << 20, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    s_d2h = torch.cuda.Stream(dev)
    check = torch.zeros(3, dtype=torch.int64, device=dev)
    state = {"i": 0, "bad": 0, "fp": 0, "cnt": 0, "verify": False}

    def sink(block, nbytes):
        # consumer: the merged stream leaves for host memory in 256 MiB blocks (what a JVM reducer would read)
        cur = torch.cuda.current_stream(dev)
        if state["verify"]:
            n = (nbytes - (2 if (nbytes % TERA_STRIDE) else 0)) // TERA_STRIDE
            b, f, c = K.validate_fixed(block, n, TERA_STRIDE, 2)
            state["bad"] += b; state["fp"] = (state["fp"] + f) & 0xFFFFFFFFFFFFFFFF; state["cnt"] += c
        ev = torch.cuda.Event()
        ev.record(cur)
        with torch.cuda.stream(s_d2h):
            s_d2h.wait_event(ev)
            for off in range(0, nbytes, 256 << 20):
                ln = min(256 << 20, nbytes - off)
                ring[state["i"] % 2][:ln].copy_(block[off:off + ln], non_blocking=True)
                state["i"] += 1
        cur.wait_stream(s_d2h)  # the block may be overwritten by the next range only after it has left

    sm = SpillMerger(TERA_STRIDE, 2, 10, int(a.budget_gb * (1 << 30)), dev)
    state["verify"] = True
    sm.merge(host_runs, nrecs, sink)
    torch.cuda.synchronize()
    ok = state["bad"] == 0 and state["cnt"] == total and state["fp"] == fp
    state["verify"] = False
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        sm.merge(host_runs, nrecs, sink)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    print(json.dumps({"metric": "TeraSort merge with pinned host-DRAM spill GB/s (1 GPU, device-timed)",
                      "value": total * 100 / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms, "records": total, "runs": a.runs,
                      "hbm_budget_gb": a.budget_gb, "data_gb": total * TERA_STRIDE / 1e9, "stats": sm.stats,
                      "pcie_bytes_per_step": int(total * TERA_STRIDE * 4), "validated": bool(ok)}), flush=True)


if __name__ == "__main__":
    main()
