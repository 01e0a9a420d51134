#!/usr/bin/env python
"""Per-source-line view of an ncu capture without the GUI.

`ncu --page source --csv` lists SASS instructions with their sample / instruction counters but no source line;
`nvdisasm -g` lists the same instructions with `//## File ..., line N` markers. This joins the two by instruction
offset inside the kernel and prints the hottest source lines (share of executed warp instructions and of stall
samples), which is what the profiling recipe's "read the stall reasons on the source page" needs on a box
without a display.

    python tools/ncu_lines.py gpurun_out/ncu/prof.ncu-rep uda_b200/lib/libuda_kernels.so fr_merge_kernelILb1ELi102 [top]
"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, lib, pat = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    seq = {}
    for cub in sorted(os.listdir(tmp)):
        if not cub.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
        infn, cur = False, None
        for l in txt.splitlines():
            if l.startswith("//---------------------"):
                infn = pat in l
                cur = None
                continue
            if not infn:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
            if m:
                seq[int(m.group(1), 16)] = cur
        if seq:
            break
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hdr]
    ia, isamp = h.index("Instructions Executed"), h.index("# Samples")
    data = rows[hdr + 1:]
    base = int(data[0][0], 16)
    agg = defaultdict(lambda: [0, 0])
    tot = ts = 0
    for r in data:
        ln = seq.get(int(r[0], 16) - base)
        agg[ln][0] += int(r[ia]); agg[ln][1] += int(r[isamp])
        tot += int(r[ia]); ts += int(r[isamp])
    cache = {}

    def text(ln):
        if ln is None:
            return "?"
        f, n = ln
        if f not in cache:
            for d in ("csrc/kernels", "csrc/runtime", "include"):
                p = os.path.join(ROOT, d, f)
                if os.path.exists(p):
                    cache[f] = open(p).read().splitlines()
                    break
            else:
                cache[f] = []
        a = cache[f]
        return a[n - 1].strip()[:110] if 0 < n <= len(a) else f

    print(f"warp instructions {tot}, stall samples {ts}")
    for ln, (i, s) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        where = f"{ln[0]}:{ln[1]}" if ln else "?"
        print(f"{where:26s} inst {100 * i / max(tot, 1):5.1f}%  samples {100 * s / max(ts, 1):5.1f}%  {text(ln)}")


if __name__ == "__main__":
    main()
