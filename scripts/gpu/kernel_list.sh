#!/bin/bash
# Per-kernel device time of one timed step (ncu launch list; cold-cache, serialised: compare shares, not absolutes).
#   scripts/gpu/kernel_list.sh <maps-per-gpu> [extra bench args]   -> gpurun_out/ncu/launches_k<maps>.csv + table on stdout
set -u
cd "$(dirname "$0")/../.."
k=$1; shift
mkdir -p gpurun_out/ncu
B="--maps-per-gpu $k --steps 2 --warmup 1 --no-e2e --no-check $*"
python bench.py $B > gpurun_out/ncu/plain_l$k.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fr_|push_kernel|sums_kernel|apply_kernel" -s 20 -c 14 --csv \
    --log-file gpurun_out/ncu/launches_k$k.csv python bench.py $B > gpurun_out/ncu/ncu_l$k.log 2>&1
python - "$k" <<'PY'
import csv, sys
k = sys.argv[1]
rows = list(csv.reader(open(f'gpurun_out/ncu/launches_k{k}.csv')))
h = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
hd = rows[h]; kn = hd.index('Kernel Name'); mv = hd.index('Metric Value')
print(f"k={k}")
for r in rows[h + 1:]:
    if len(r) > mv:
        print(f"   {r[kn][:70]:72s} {float(r[mv].replace(',', '')) / 1e3:9.1f} us")
PY
