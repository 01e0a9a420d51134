#!/usr/bin/env python
"""One table row per profiled kernel launch of an ncu report: time, DRAM traffic and achieved rate against the measured
copy roofline (MEASURED_PEAKS.json), issue utilisation, occupancy limiters.

    python tools/ncu_summary.py gpurun_out/ncu/*.ncu-rep > profiles/r2/ncu_kernels.md
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0   # the profiling recipe's fallback


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    if len(r) < 3:
        return []
    hdr, units = r[0], r[1]
    res = []
    for vals in r[2:]:
        d = dict(zip(hdr, vals))
        u = dict(zip(hdr, units))
        res.append((d, u))
    return res


def num(d, k, default=0.0):
    try:
        return float(str(d.get(k, "")).replace(",", ""))
    except ValueError:
        return default


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)


def to_us(v, unit):
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)


def main():
    print(f"| report | kernel | time (us) | DRAM read + write (MB) | DRAM GB/s | of measured copy roofline ({PEAK:.0f} GB/s) | issue active % | regs | CTAs/SM limit (regs / smem) | warps active % |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for rep in sys.argv[1:]:
        for d, u in rows_of(rep):
            name = d.get("Kernel Name", "?").split("(")[0].replace("void ", "").replace("uda::", "")
            t = to_us(num(d, "gpu__time_duration.sum"), u.get("gpu__time_duration.sum", "us"))
            rd = to_bytes(num(d, "dram__bytes_read.sum"), u.get("dram__bytes_read.sum", "byte"))
            wr = to_bytes(num(d, "dram__bytes_write.sum"), u.get("dram__bytes_write.sum", "byte"))
            gbs = (rd + wr) / max(t, 1e-9) / 1e3
            print(f"| {os.path.basename(rep).replace('.ncu-rep', '')} | `{name[:48]}` | {t:.1f} | {(rd + wr) / 1e6:.1f} | {gbs:.0f} | {100 * gbs / PEAK:.0f} % | "
                  f"{num(d, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.0f} | {num(d, 'launch__registers_per_thread'):.0f} | "
                  f"{num(d, 'launch__occupancy_limit_registers'):.0f} / {num(d, 'launch__occupancy_limit_shared_mem'):.0f} | "
                  f"{num(d, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.0f} |")


if __name__ == "__main__":
    main()
