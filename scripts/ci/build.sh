#!/usr/bin/env bash
# One CI pass (counterpart of the reference's daily build, scripts/build/build.sh:1-277: fetch, build, package,
# static analysis). Here: native build for sm_100a with -Werror, CPU test suite, host sanitizers, SASS evidence
# check, wheel. GPU stages (pytest -m gpu, bench, compute-sanitizer) run when a GPU is visible.
set -euo pipefail
root="$(cd "$(dirname "$0")/../.." && pwd)"
cd "$root"
echo "== build =="
python tools/build.py -v | tail -20
echo "== version =="
python -c "import ctypes; from uda_b200 import _lib; L=_lib.runtime(); L.uda_rt_version.restype=ctypes.c_char_p; print(L.uda_rt_version().decode())"
echo "== CPU tests =="
python -m pytest tests/ -x -q -m "not gpu"
echo "== host sanitizers =="
scripts/ci/host_sanitizers.sh address
scripts/ci/host_sanitizers.sh thread
scripts/ci/host_sanitizers.sh undefined
echo "== coverage of the host runtime =="
scripts/ci/coverage.sh | tail -3
echo "== reference arm (the reference's own reducer sources, unmodified) =="
python baseline/build_ref.py > /dev/null && python bench.py --impl reference --steps 1 --warmup 0 --records-per-gpu 500000 | cut -c1-160
echo "== CPU regression =="
UDA_LIMITED_HOME="${TMPDIR:-/tmp}/uda_ci_regression" scripts/regression_limited/run.sh | tail -8
echo "== static analysis =="
scripts/ci/static_analysis.sh
echo "== SASS evidence =="
python tools/sass_listing.py > /dev/null
grep -q "UBLKCP" profiles/SASS_SUMMARY.md || { echo "no TMA bulk copies in the merge/push kernels"; exit 1; }
grep -q "REDG.E.ADD.64.STRONG.SYS" profiles/SASS_SUMMARY.md || { echo "no system-scope release reduction in the push kernel"; exit 1; }
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
    echo "== GPU tests =="
    python -m pytest tests/ -x -q -m gpu
    echo "== bench =="
    python bench.py --steps 5 --warmup 3
else
    echo "no GPU visible: GPU stages skipped (run them with: gpurun -- 'make test-gpu bench')"
fi
echo "CI pass complete"
