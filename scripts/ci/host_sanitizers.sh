#!/usr/bin/env bash
# Host runtime under ASAN / TSAN / UBSAN: builds build/<kind>/libuda.so + the native tests and runs
#   * the mock-JVM test of the JNI bridge (threads attach/detach, up-calls, failure latch)
#   * the plumbing self-tests through the C API (queues, thread wrapper)
# usage: scripts/ci/host_sanitizers.sh address|thread|undefined
set -euo pipefail
kind="${1:-address}"
root="$(cd "$(dirname "$0")/../.." && pwd)"
cd "$root"
python tools/build.py --sanitize "$kind"
# leak detection is off: the mock JVM of the test never frees its fake jobjects (a real JVM collects local references)
export ASAN_OPTIONS="detect_leaks=0:abort_on_error=1" TSAN_OPTIONS="halt_on_error=1 second_deadlock_stack=1" UBSAN_OPTIONS="halt_on_error=1"
tmp="$(mktemp -d)"
trap 'rm -rf "$tmp"' EXIT
# the test program dlopens the library it is given: the sanitizer runtime must come first in the process
"build/$kind/test_jni_mock" "build/$kind/libuda.so" "$tmp"
# the same flow with the JNI consumer on merge approach 3, against the host-only mock of the kernel library
UDA_KERNELS_LIB="$root/build/$kind/libuda_kernels_mock.so" "build/$kind/test_jni_mock" "build/$kind/libuda.so" "$tmp" 3
if [ "$kind" = "address" ] && [ "${UDA_SAN_PYTESTS:-1}" = "1" ]; then
    # the whole CPU suite of the runtime (plumbing over TCP, hybrid merge, Snappy / LZO segments, approach 3 fallback)
    # against the instrumented library: the sanitizer runtime and libstdc++ are preloaded into the Python process
    asan_lib="$(/usr/bin/g++ -print-file-name=libasan.so)"
    stdcpp="$(/usr/bin/g++ -print-file-name=libstdc++.so.6)"
    UDA_B200_SKIP_BUILD=1 UDA_B200_RUNTIME_LIB="$root/build/address/libuda.so" LD_PRELOAD="$asan_lib $stdcpp" \
        ASAN_OPTIONS="detect_leaks=0:abort_on_error=0:protect_shadow_gap=0" \
        python -m pytest tests/test_runtime_cpu.py tests/test_fuzz_cpu.py tests/test_device_merge_mock_cpu.py -x -q -p no:cacheprovider -k "not reader_bench and not regression_driver"
fi
if [ "$kind" = "thread" ] && [ "${UDA_SAN_PYTESTS:-1}" = "1" ]; then
    # the multi-threaded part of the suite (provider, consumers, transport, hybrid merge, decompression) under TSAN
    tsan_lib="$(/usr/bin/g++ -print-file-name=libtsan.so)"
    stdcpp="$(/usr/bin/g++ -print-file-name=libstdc++.so.6)"
    log="$(mktemp)"
    # reports go to files of their own (log_path): pytest captures the stderr of passing tests, which hid a report once
    UDA_KERNELS_LIB=/nonexistent.so UDA_B200_SKIP_BUILD=1 UDA_B200_RUNTIME_LIB="$root/build/thread/libuda.so" LD_PRELOAD="$tsan_lib $stdcpp" \
        TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 log_path=$tmp/tsan" \
        python -m pytest tests/test_runtime_cpu.py tests/test_device_merge_mock_cpu.py -x -q -p no:cacheprovider -k "plumbing or queues or thread_wrapper or mock" > "$log" 2>&1 || true
    tail -1 "$log"
    # the native benchmark as a stress: several reducers in one process, compressed map outputs (decompress threads), hybrid merge
    for args in "--tcp --maps 8 --reducers 3 --records 600000 --codec lzo" "--tcp --maps 8 --reducers 3 --records 600000 --codec snappy --approach 2" \
                "--maps 16 --records 600000 --approach 2"; do
        TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 log_path=$tmp/tsan" "build/thread/cpu_merge_bench" x $args --steps 2 --warmup 0 --check | tail -1
    done
    if ls "$tmp"/tsan.* > /dev/null 2>&1; then
        grep -h -A12 "WARNING: ThreadSanitizer" "$tmp"/tsan.* | head -80
        grep -h "SUMMARY: ThreadSanitizer" "$tmp"/tsan.* | sort | uniq -c
        exit 1
    fi
    grep -q " passed" "$log" && ! grep -q " failed" "$log" || { tail -30 "$log"; exit 1; }
fi
if [ "$kind" = "undefined" ] && [ "${UDA_SAN_PYTESTS:-1}" = "1" ]; then
    # parsers, decoders and the plumbing under UBSAN (signed overflow, misaligned access, bad shifts in the hand-written codecs)
    ub_lib="$(/usr/bin/g++ -print-file-name=libubsan.so)"
    stdcpp="$(/usr/bin/g++ -print-file-name=libstdc++.so.6)"
    UDA_B200_SKIP_BUILD=1 UDA_B200_RUNTIME_LIB="$root/build/undefined/libuda.so" LD_PRELOAD="$ub_lib $stdcpp" \
        UBSAN_OPTIONS="halt_on_error=0:print_stacktrace=1:log_path=$tmp/ubsan" \
        python -m pytest tests/test_fuzz_cpu.py tests/test_runtime_cpu.py -x -q -p no:cacheprovider \
        -k "not reader_bench and not regression_driver and not cpu_merge_bench" | tail -1
    if ls "$tmp"/ubsan.* > /dev/null 2>&1; then
        grep -h "runtime error" "$tmp"/ubsan.* | sort | uniq -c | head -20
        exit 1
    fi
fi
echo "host sanitizer run ($kind): ok"
