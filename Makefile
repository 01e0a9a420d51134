# Top-level build (counterpart of the reference's Makefile:4-20 -> src/Makefile.am + plugins/Makefile).
# The work is done by tools/build.py (nvcc for sm_100a, g++ -Wall -Werror for the runtime + JNI bridge);
# this file only gives the usual verbs.
PY ?= python

.PHONY: all native java test test-gpu bench cpu-bench ci sass sanitize asan tsan analyze coverage wheel rpm deb install reference clean

all: native

native:
	$(PY) tools/build.py

# javac is not part of every image: the plug-in sources build wherever a JDK and the Hadoop jars are present
java:
	$(MAKE) -C java

test: native
	$(PY) -m pytest tests/ -x -q -m "not gpu"

test-gpu: native
	$(PY) -m pytest tests/ -x -q -m gpu

bench: native
	$(PY) bench.py

sass: native
	$(PY) tools/sass_listing.py

# compute-sanitizer over the single-GPU kernel tests (memcheck by default: TOOL=racecheck|synccheck|initcheck)
sanitize: native
	scripts/ci/sanitize.sh $(TOOL)

# host runtime under AddressSanitizer / ThreadSanitizer (CPU tests of the runtime and the mock-JVM test)
asan:
	scripts/ci/host_sanitizers.sh address
tsan:
	scripts/ci/host_sanitizers.sh thread

# stricter warning set + cppcheck (when installed) over the host runtime
analyze:
	scripts/ci/static_analysis.sh

# line coverage of the host runtime by the CPU suite (gcc --coverage + gcov) -> build/coverage/summary.txt
coverage:
	scripts/ci/coverage.sh

# the reference arm of bench.py: the reference's own reducer sources compiled unmodified (needs /root/reference or a copy)
reference:
	$(PY) baseline/build_ref.py

# the CPU mergers side by side, both fed from memory (this runtime's reducer and the reference's own), then BASELINE config 1
# natively (provider role over TCP loop-back, two reducers)
cpu-bench: native reference
	build/bin/cpu_merge_bench x --maps 8 --records 20000000 --steps 3 --warmup 1 --check
	baseline/_ref/bin/ref_merger --maps 8 --records 20000000 --steps 3 --warmup 1 --check
	build/bin/cpu_merge_bench x --tcp --maps 4 --reducers 2 --records 10000000 --steps 3 --warmup 1 --check

# one full CI pass (scripts/ci/build.sh): build, CPU suite, sanitizers, coverage, static analysis, reference arm, regression
ci:
	scripts/ci/build.sh

wheel: native
	$(PY) -m pip wheel --no-deps --no-build-isolation -w dist .

rpm:
	packaging/buildrpm.sh

deb: native
	$(PY) packaging/build_deb.py

# unpacked-tree install + Hadoop hook: make install UDA_DIR=/opt/uda [HADOOP_CONF=/etc/hadoop/conf]
install: native
	packaging/install-uda.sh --uda-dir=$(UDA_DIR) $(if $(HADOOP_CONF),--uda-hadoop-conf-dir=$(HADOOP_CONF))

clean:
	rm -rf build dist uda_b200/lib/*.so *.egg-info
