#!/usr/bin/env python
"""In-tree native build (replaces the reference's autotools flow: src/premake.sh:22-36,
src/configure.ac:31-114, src/Makefile.am:25-63 — ``-std=gnu++0x -Wall -Werror -O3 -g`` and a
``VERSION_UDA`` define derived from release + git revision count).

Outputs (git-ignored, shipped to the GPU box by gpurun):
  uda_b200/lib/libuda_kernels.so   sm_100a CUDA kernels behind the C ABI in include/uda_b200.h
  uda_b200/lib/libuda.so           C++ runtime + JNI bridge (same file name as the reference's library)

Only sm_100a code is generated (``-gencode arch=compute_100a,code=sm_100a``); ``-lineinfo`` keeps the
ncu source page usable.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shlex
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB_DIR = ROOT / "uda_b200" / "lib"
# UDA_BUILD_TAG=<tag>: an A/B variant of the kernel library (UDA_FR_* / UDA_TP_* defines from the environment) is
# written to libuda_kernels_<tag>.so with its own object directory; select it with UDA_B200_KERNELS_LIB
TAG = os.environ.get("UDA_BUILD_TAG", "")
OBJ_DIR = ROOT / "build" / ("obj_" + TAG if TAG else "obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
    "-Xptxas", "-v",
    "-I", str(ROOT / "include"),
]
CXX_FLAGS = ["-O2", "-g", "-std=c++17", "-fPIC", "-Wall", "-Werror", "-pthread",
             "-I", str(ROOT / "include"), "-I", str(ROOT / "csrc" / "runtime"), "-I", str(ROOT / "csrc" / "jni")]


def uda_version() -> str:
    """``<release>-<git rev count>`` like the reference's VERSION_UDA (src/configure.ac:31-39)."""
    try:
        n = subprocess.run(["git", "rev-list", "--count", "HEAD"], cwd=ROOT, capture_output=True, text=True, timeout=10).stdout.strip()
    except Exception:
        n = ""
    return f"b200-1.0.{n or '0'}"


def _run(cmd, log_path: Path | None = None) -> str:
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True)
    out = p.stdout + p.stderr
    if log_path is not None:
        log_path.write_text(" ".join(shlex.quote(c) for c in cmd) + "\n" + out)
    if p.returncode != 0:
        sys.stderr.write(out)
        raise RuntimeError(f"build step failed: {' '.join(cmd[:3])} ... (see output above)")
    return out


def _stamp(src: Path, flags) -> str:
    h = hashlib.sha256()
    h.update(src.read_bytes())
    h.update(" ".join(flags).encode())
    for hdr in sorted(list((ROOT / "include").glob("*.h")) + list(src.parent.glob("*.cuh")) + list(src.parent.glob("*.h"))
                      + list((ROOT / "csrc" / "runtime").glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()


def _compile(src: Path, compiler: str, flags) -> Path:
    obj = OBJ_DIR / (src.parent.name + "_" + src.stem + ".o")
    stamp_file = obj.with_suffix(".stamp")
    stamp = _stamp(src, flags)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return obj
    _run([compiler, *flags, "-c", str(src), "-o", str(obj)], obj.with_suffix(".log"))
    stamp_file.write_text(stamp)
    return obj


def build(verbose: bool = False) -> dict:
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    version = uda_version()
    cu = sorted((ROOT / "csrc" / "kernels").glob("*.cu"))
    cc = sorted((ROOT / "csrc" / "runtime").glob("*.cc")) + sorted((ROOT / "csrc" / "jni").glob("*.cc"))
    vflag = [f'-DVERSION_UDA="{version}"'] + [f"-D{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith(("UDA_FR_", "UDA_TP_", "UDA_DC_"))]
    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        cu_objs = list(ex.map(lambda s: _compile(s, NVCC, NVCC_FLAGS + vflag), cu))
        cc_objs = list(ex.map(lambda s: _compile(s, CXX, CXX_FLAGS + vflag), cc))
    out = {}
    klib = LIB_DIR / (f"libuda_kernels_{TAG}.so" if TAG else "libuda_kernels.so")
    _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(klib), *map(str, cu_objs)])
    out["kernels"] = klib
    if TAG:
        return out
    if cc_objs:
        rlib = LIB_DIR / "libuda.so"
        _run([CXX, "-shared", "-pthread", "-o", str(rlib), *map(str, cc_objs), "-ldl"])
        out["runtime"] = rlib
    # native test programs (mock-JVM test of the JNI bridge)
    bin_dir = ROOT / "build" / "bin"
    bin_dir.mkdir(parents=True, exist_ok=True)
    for src in sorted((ROOT / "tests" / "native").glob("*.cc")):
        exe = bin_dir / src.stem
        if cc_objs:
            _run([CXX, *CXX_FLAGS, str(src), "-o", str(exe), f"-L{LIB_DIR}", "-luda", f"-Wl,-rpath,{LIB_DIR}", "-ldl"])
            out[src.stem] = exe
    # host-only stand-in for the kernel library: lets the CPU tests drive csrc/runtime/device_merge.cc (UDA_KERNELS_LIB)
    mock = ROOT / "tests" / "native" / "mock" / "mock_kernels.cc"
    if mock.exists():
        mlib = bin_dir / "libuda_kernels_mock.so"
        if not mlib.exists() or mlib.stat().st_mtime < max(mock.stat().st_mtime, (ROOT / "include" / "uda_b200.h").stat().st_mtime):
            _run([CXX, *CXX_FLAGS, "-shared", str(mock), "-o", str(mlib)])
        out["kernels_mock"] = mlib
    if verbose:
        for o in cu_objs:
            log = o.with_suffix(".log")
            if log.exists():
                for line in log.read_text().splitlines():
                    if "registers" in line or "spill" in line.lower() and "0 bytes spill" not in line:
                        print(f"[{o.stem}] {line.strip()}")
    return out


def build_host_sanitized(kind: str) -> dict:
    """Host runtime + native tests under ``-fsanitize=<kind>`` (address | thread | undefined) in build/<kind>/.
    The reference has no sanitizer targets; its known latent races are exactly what TSAN is for."""
    assert kind in ("address", "thread", "undefined")
    out_dir = ROOT / "build" / kind
    out_dir.mkdir(parents=True, exist_ok=True)
    cc = sorted((ROOT / "csrc" / "runtime").glob("*.cc")) + sorted((ROOT / "csrc" / "jni").glob("*.cc"))
    flags = [f for f in CXX_FLAGS if f != "-O2"] + ["-O1", "-fno-omit-frame-pointer", f"-fsanitize={kind}", f'-DVERSION_UDA="{uda_version()}"']
    rlib = out_dir / "libuda.so"
    # the sanitizer runtimes ship with the distribution compiler, not necessarily with $CXX
    cxx = os.environ.get("SAN_CXX", "/usr/bin/g++" if Path("/usr/bin/g++").exists() else CXX)
    _run([cxx, *flags, "-shared", "-o", str(rlib), *map(str, cc), "-ldl"])
    out = {"runtime": rlib}
    for src in sorted((ROOT / "tests" / "native").glob("*.cc")):
        exe = out_dir / src.stem
        _run([cxx, *flags, str(src), "-o", str(exe), f"-L{out_dir}", "-luda", f"-Wl,-rpath,{out_dir}", "-ldl"])
        out[src.stem] = exe
    mock = ROOT / "tests" / "native" / "mock" / "mock_kernels.cc"
    if mock.exists():
        mlib = out_dir / "libuda_kernels_mock.so"
        _run([cxx, *flags, "-shared", str(mock), "-o", str(mlib)])
        out["kernels_mock"] = mlib
    return out


if __name__ == "__main__":
    if "--sanitize" in sys.argv:
        for k, v in build_host_sanitized(sys.argv[sys.argv.index("--sanitize") + 1]).items():
            print(f"built {k}: {v}")
        sys.exit(0)
    res = build(verbose="-v" in sys.argv)
    for k, v in res.items():
        print(f"built {k}: {v}")
