#!/usr/bin/env bash
# compute-sanitizer over the single-GPU kernel tests (needs a GPU: run it through gpurun, ONE tool per call).
# usage: scripts/ci/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
set -euo pipefail
tool="${1:-memcheck}"
expr="${2:-radix or crc32 or ifile or merge or gather or push or scan or partition or var_merger or tiebreak}"
root="$(cd "$(dirname "$0")/../.." && pwd)"
cd "$root"
mkdir -p gpurun_out
# small inputs only: the sanitizer slows kernels down 10-100x
compute-sanitizer --tool "$tool" --error-exitcode 86 --log-file "gpurun_out/sanitizer_${tool}.log" \
    python -m pytest tests/test_kernels_gpu.py -x -q -k "($expr) and not 6_000_000 and not 3_000_000 and not large" -p no:cacheprovider
echo "compute-sanitizer $tool: exit $?"
tail -5 "gpurun_out/sanitizer_${tool}.log"
