#!/usr/bin/env bash
# Line coverage of the host runtime (csrc/runtime, csrc/jni) by the CPU test-suite (the reference's daily build ran a
# Bullseye coverage job: scripts/build/coverage*.sh, BULLSEYE_EXCLUDE markers in src/). gcc --coverage + gcov; no lcov needed.
#   scripts/ci/coverage.sh            -> build/coverage/summary.txt (one line per source file + total)
set -euo pipefail
root="$(cd "$(dirname "$0")/../.." && pwd)"
cd "$root"
out="build/coverage"
rm -rf "$out"; mkdir -p "$out/obj"
flags="-O0 -g --coverage -std=c++17 -fPIC -pthread -Iinclude -Icsrc/runtime -Icsrc/jni -DVERSION_UDA=\"coverage\""
objs=()
for src in csrc/runtime/*.cc csrc/jni/*.cc; do
    o="$out/obj/$(basename "$(dirname "$src")")_$(basename "${src%.cc}").o"
    g++ $flags -c "$src" -o "$o"
    objs+=("$o")
done
g++ -shared --coverage -pthread -o "$out/libuda.so" "${objs[@]}" -ldl
g++ $flags tests/native/test_jni_mock.cc -o "$out/test_jni_mock" -L"$out" -luda -Wl,-rpath,"$root/$out" -ldl
tmp="$(mktemp -d)"; trap 'rm -rf "$tmp"' EXIT
"$out/test_jni_mock" "$out/libuda.so" "$tmp" > "$out/test_jni_mock.log" 2>&1
UDA_KERNELS_LIB=/nonexistent.so UDA_B200_SKIP_BUILD=1 UDA_B200_RUNTIME_LIB="$root/$out/libuda.so" \
    python -m pytest tests/test_runtime_cpu.py tests/test_fuzz_cpu.py -x -q -p no:cacheprovider -k "not reader_bench and not regression_driver and not jni" \
    > "$out/pytest.log" 2>&1 || { tail -20 "$out/pytest.log"; exit 1; }
tail -1 "$out/pytest.log"
(cd "$out/obj" && gcov -r -s "$root" *.gcda > ../gcov.log 2>&1 || true)
python - "$out" <<'PY'
import re, sys
out = sys.argv[1]
txt = open(f"{out}/gcov.log").read()
rows, tot_l, tot_c = [], 0, 0.0
for m in re.finditer(r"File '([^']+)'\nLines executed:([0-9.]+)% of (\d+)", txt):
    f, pct, n = m.group(1), float(m.group(2)), int(m.group(3))
    if not f.startswith("csrc/"):
        continue
    rows.append((f, pct, n)); tot_l += n; tot_c += pct * n / 100
rows.sort()
with open(f"{out}/summary.txt", "w") as fh:
    for f, pct, n in rows:
        fh.write(f"{f:40s} {pct:6.1f} % of {n} lines\n")
    fh.write(f"{'TOTAL':40s} {100 * tot_c / max(tot_l, 1):6.1f} % of {tot_l} lines\n")
print(open(f"{out}/summary.txt").read())
PY
