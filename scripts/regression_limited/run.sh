#!/usr/bin/env bash
# Regression for accounts with limited permissions (counterpart of the reference's
# scripts/regression_for_limited_permissions_cluster/: a copy of its cluster regression that does without sudo, without
# writing outside the user's directories and without restarting services). Here the full driver already needs no
# privileges; what this variant removes is everything a locked-down account may not have: GPUs (CPU rows only), a writable
# repository (the build goes to a private copy under $UDA_LIMITED_HOME, default ~/.uda_b200_regression) and the shared
# report directory (no --publish).
#   scripts/regression_limited/run.sh [--full]        (--full: use the whole matrix when GPUs are visible)
set -euo pipefail
root="$(cd "$(dirname "$0")/../.." && pwd)"
home="${UDA_LIMITED_HOME:-$HOME/.uda_b200_regression}"
mkdir -p "$home"
matrix="$root/scripts/regression_limited/matrix_cpu.csv"
[ "${1:-}" = "--full" ] && matrix="$root/scripts/regression/matrix.csv"
work="$root"
if [ ! -w "$root/uda_b200" ]; then
    # read-only checkout: build in a private copy
    work="$home/tree"
    rm -rf "$work"; mkdir -p "$work"
    (cd "$root" && tar --exclude=.git --exclude=gpurun_out --exclude=build --exclude='*.so' -cf - .) | (cd "$work" && tar -xf -)
fi
cd "$work"
python tools/build.py > "$home/build.log" 2>&1 || { tail -20 "$home/build.log"; exit 1; }
python scripts/regression/auto_tester.py --matrix "$matrix" --out "$home/report"
echo "report: $home/report"
