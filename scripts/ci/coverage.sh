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
    python -m pytest tests/test_runtime_cpu.py tests/test_fuzz_cpu.py tests/test_device_merge_mock_cpu.py -x -q -p no:cacheprovider -k "not reader_bench and not regression_driver and not jni" \
    > "$out/pytest.log" 2>&1 || { tail -20 "$out/pytest.log"; exit 1; }
tail -1 "$out/pytest.log"
# one JSON report per translation unit; a header's inline functions appear in every unit that includes it and count as
# covered when any unit executed them (a per-unit summary reports them as missed in the units that never call them)
python - "$out" <<'PY'
import glob, json, subprocess, sys
out = sys.argv[1]
lines = {}   # file -> {line: max count}
for gcda in sorted(glob.glob(f"{out}/obj/*.gcda")):
    p = subprocess.run(["gcov", "-j", "-t", "-o", f"{out}/obj", gcda], capture_output=True, text=True)
    if p.returncode != 0 or not p.stdout.strip():
        continue
    for f in json.loads(p.stdout)["files"]:
        name = f["file"]
        if not name.startswith("csrc/"):
            continue
        d = lines.setdefault(name, {})
        for ln in f["lines"]:
            d[ln["line_number"]] = max(d.get(ln["line_number"], 0), ln["count"])
tot = cov = 0
with open(f"{out}/summary.txt", "w") as fh:
    for name in sorted(lines):
        n = len(lines[name]); c = sum(1 for v in lines[name].values() if v > 0)
        tot += n; cov += c
        fh.write(f"{name:40s} {100 * c / max(n, 1):6.1f} % of {n} lines\n")
    fh.write(f"{'TOTAL':40s} {100 * cov / max(tot, 1):6.1f} % of {tot} lines\n")
    with open(f"{out}/missed.txt", "w") as mh:
        for name in sorted(lines):
            miss = sorted(k for k, v in lines[name].items() if v == 0)
            if miss:
                mh.write(f"{name}: {' '.join(map(str, miss))}\n")
print(open(f"{out}/summary.txt").read())
PY
