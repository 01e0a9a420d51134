#!/usr/bin/env bash
# Static analysis pass over the host runtime (counterpart of the reference's Coverity target, src/Makefile.am:62-63,
# scripts/build/codeAnalysis.sh:39-52): a stricter warning set than the -Wall -Werror build, cppcheck when it is
# installed, and nvcc's own warnings for the kernels. Exit code 1 when the strict compile reports anything.
set -uo pipefail
root="$(cd "$(dirname "$0")/../.." && pwd)"
cd "$root"
out="${1:-build/static_analysis.log}"
mkdir -p "$(dirname "$out")"
: > "$out"
cxx="${CXX:-g++}"
flags="-std=c++17 -fsyntax-only -Wall -Wextra -Wshadow -Wnon-virtual-dtor -Woverloaded-virtual -Wcast-align -Wformat=2 -Wnull-dereference -Wduplicated-cond -Wlogical-op -Wno-unused-parameter -Iinclude -Icsrc/runtime -Icsrc/jni"
n=0
for f in csrc/runtime/*.cc csrc/jni/*.cc tests/native/*.cc; do
    if ! $cxx $flags "$f" >> "$out" 2>&1; then n=$((n + 1)); fi
done
warnings=$(grep -c "warning:" "$out" || true)
echo "strict compile: $warnings warning(s), $n file(s) with errors (log: $out)"
if command -v cppcheck >/dev/null; then
    cppcheck --enable=warning,performance,portability --std=c++17 --quiet -Iinclude -Icsrc/runtime csrc/runtime csrc/jni 2>> "$out"
    echo "cppcheck: done"
else
    echo "cppcheck not installed: skipped"
fi
[ "$warnings" = "0" ] && [ "$n" = "0" ]
