#!/usr/bin/env bash
# Merge approach 3 (device merge through the plug-in surface) against approach 1 (CPU heap) without Python in the loop:
# provider role over TCP loop-back, consumer roles, byte-counting sink (tests/native/cpu_merge_bench.cc).
#   gpurun --timeout 600 -- 'scripts/gpu/approach3_native.sh > gpurun_out/approach3_native.txt 2>&1'
# Few maps favour the streaming CPU merge (approach 3 is store-and-forward); many maps are where the device should win.
set -uo pipefail
cd "$(dirname "$0")/../.."
python tools/build.py > /dev/null
for shape in "4 2 10000000" "64 1 20000000" "256 1 20000000"; do
    set -- $shape
    for a in 1 3; do
        echo "# maps=$1 reducers=$2 records=$3 approach=$a"
        timeout 300 build/bin/cpu_merge_bench x --tcp --maps "$1" --reducers "$2" --records "$3" --steps 3 --warmup 1 --approach "$a" --check
    done
done
