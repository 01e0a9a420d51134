# Top-level build (counterpart of the reference's Makefile:4-20 -> src/Makefile.am + plugins/Makefile).
# The work is done by tools/build.py (nvcc for sm_100a, g++ -Wall -Werror for the runtime + JNI bridge);
# this file only gives the usual verbs.
PY ?= python

.PHONY: all native java test test-gpu bench sass sanitize asan tsan analyze wheel rpm clean

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

wheel: native
	$(PY) -m pip wheel --no-deps --no-build-isolation -w dist .

rpm:
	packaging/buildrpm.sh

clean:
	rm -rf build dist uda_b200/lib/*.so *.egg-info
