#!/usr/bin/env bash
# Source tarball + rpm (reference: build/buildrpm.sh — tarball without jars, premake, makerpm). Falls back to the
# tarball + wheel when rpmbuild is not installed (it is not part of the GPU image).
set -euo pipefail
root="$(cd "$(dirname "$0")/.." && pwd)"
cd "$root"
ver="1.0.$(git rev-list --count HEAD 2>/dev/null || echo 0)"
mkdir -p dist
tar --exclude-vcs --exclude='./build' --exclude='./dist' --exclude='./gpurun_out' --exclude='*.so' --exclude='./baseline/_ref' \
    --transform "s,^\.,uda_b200-$ver," -czf "dist/uda_b200-$ver.tar.gz" .
echo "source tarball: dist/uda_b200-$ver.tar.gz"
if command -v rpmbuild >/dev/null; then
    top="$(mktemp -d)"
    mkdir -p "$top"/{SOURCES,SPECS,BUILD,RPMS,SRPMS}
    cp "dist/uda_b200-$ver.tar.gz" "$top/SOURCES/"
    rpmbuild --define "_topdir $top" --define "_uda_version $ver" -ba packaging/uda_b200.spec
    find "$top/RPMS" -name '*.rpm' -exec cp {} dist/ \;
else
    echo "rpmbuild not found: building the wheel instead"
    python tools/build.py
    python -m pip wheel --no-deps --no-build-isolation -w dist . >/dev/null && ls dist/*.whl
fi
