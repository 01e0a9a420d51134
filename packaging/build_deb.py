#!/usr/bin/env python
"""Debian package of the built tree (counterpart of the reference's build/build-deb-from-rpm.sh:36-107, which unpacks the
rpm and re-wraps its payload with a `debian/` template directory; metadata comes from the rpm: version, release, packager,
summary, architecture x86_64 -> amd64).

Here the .deb is written directly — `ar` archive of `debian-binary`, `control.tar.gz`, `data.tar.gz` — so neither rpm nor
debhelper is needed on the build machine (the image has neither). With `--from-rpm file.rpm` the version, release and
summary are taken from the rpm instead of the tree (needs `rpm` on PATH), which is what the reference's script does.

    python packaging/build_deb.py [--out dist] [--from-rpm dist/uda_b200-1.0.N-1.x86_64.rpm]

Payload layout = the rpm's (packaging/uda_b200.spec): everything under /usr/lib64/uda, plus /usr/share/doc/uda-b200.
"""
from __future__ import annotations

import argparse
import gzip
import io
import os
import subprocess
import sys
import tarfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PREFIX = "usr/lib64/uda"


def tree_version() -> str:
    try:
        n = subprocess.run(["git", "rev-list", "--count", "HEAD"], cwd=ROOT, capture_output=True, text=True, timeout=10).stdout.strip()
    except Exception:
        n = ""
    return f"1.0.{n or '0'}"


def rpm_field(rpm: str, field: str) -> str:
    return subprocess.run(["rpm", "-qp", "--queryformat", "%{" + field + "}", rpm], capture_output=True, text=True, check=True).stdout.strip()


def payload_files():
    """(source path, path inside the package) pairs."""
    out = []
    for lib in ("libuda.so", "libuda_kernels.so"):
        p = ROOT / "uda_b200" / "lib" / lib
        if p.exists():
            out.append((p, f"{PREFIX}/{lib}"))
    for base in ("uda_b200", "java", "utils"):
        for p in sorted((ROOT / base).rglob("*")):
            if p.is_file() and "__pycache__" not in p.parts and p.suffix not in (".so", ".pyc", ".o"):
                out.append((p, f"{PREFIX}/{p.relative_to(ROOT)}"))
    for doc in ("README.md", "INTEGRATION.md", "DESIGN.md", "docs/journal.txt"):
        p = ROOT / doc
        if p.exists():
            out.append((p, f"{PREFIX}/{Path(doc).name}"))
            out.append((p, f"usr/share/doc/uda-b200/{Path(doc).name}"))
    out.append((ROOT / "packaging" / "install-uda.sh", f"{PREFIX}/install-uda.sh"))
    return out


def tar_gz(members) -> bytes:
    """members: (arcname, bytes, mode). Directories are created on the way."""
    buf = io.BytesIO()
    with gzip.GzipFile(fileobj=buf, mode="wb", mtime=0) as gz, tarfile.open(fileobj=gz, mode="w", format=tarfile.GNU_FORMAT) as tf:
        seen = set()
        for name, data, mode in members:
            parts = Path(name).parts
            for i in range(1, len(parts)):
                d = "./" + "/".join(parts[:i])
                if d not in seen:
                    seen.add(d)
                    ti = tarfile.TarInfo(d)
                    ti.type, ti.mode, ti.mtime = tarfile.DIRTYPE, 0o755, int(time.time())
                    tf.addfile(ti)
            ti = tarfile.TarInfo("./" + name)
            ti.size, ti.mode, ti.mtime = len(data), mode, int(time.time())
            tf.addfile(ti, io.BytesIO(data))
    return buf.getvalue()


def ar_archive(members) -> bytes:
    """The common `ar` format: global header, then 60-byte member headers, data padded to even length."""
    out = io.BytesIO()
    out.write(b"!<arch>\n")
    for name, data in members:
        hdr = f"{name:<16}{int(time.time()):<12}{0:<6}{0:<6}{'100644':<8}{len(data):<10}`\n".encode()
        assert len(hdr) == 60
        out.write(hdr)
        out.write(data)
        if len(data) % 2:
            out.write(b"\n")
    return out.getvalue()


def build(out_dir: Path, from_rpm: str | None = None) -> Path:
    version, release = tree_version(), "1"
    summary = "B200-native MapReduce shuffle-and-merge engine (UDA plug-in surface)"
    packager = os.environ.get("DEBFULLNAME", "uda_b200 maintainers")
    if from_rpm:
        version, release = rpm_field(from_rpm, "VERSION"), rpm_field(from_rpm, "RELEASE")
        summary = rpm_field(from_rpm, "SUMMARY") or summary
        packager = rpm_field(from_rpm, "PACKAGER") or packager
    files = payload_files()
    data_members = [(dst, src.read_bytes(), 0o755 if (src.suffix in (".so", ".sh") or os.access(src, os.X_OK)) else 0o644) for src, dst in files]
    installed_kb = sum(len(d) for _, d, _ in data_members) // 1024 + 1
    control = (
        f"Package: uda-b200\nVersion: {version}-{release}\nSection: misc\nPriority: optional\nArchitecture: amd64\n"
        f"Installed-Size: {installed_kb}\nMaintainer: {packager}\nDepends: python3\n"
        f"Description: {summary}\n"
        " Hadoop shuffle provider/consumer plug-in (UdaBridge JNI surface, libuda.so) whose data path runs on NVIDIA B200\n"
        " GPUs: map outputs resident in HBM, push shuffle over NVLink peer memory, k-way merge kernels, pinned-DRAM spill.\n"
    ).encode()
    postinst = (
        "#!/bin/sh\nset -e\n# register the libraries and point Hadoop at the plug-in when its configuration directory is known\n"
        f"if [ -n \"${{UDA_HADOOP_CONF_DIR:-}}\" ]; then /{PREFIX}/install-uda.sh --uda-dir=/{PREFIX} --prefix=/ "
        "--uda-hadoop-conf-dir=\"$UDA_HADOOP_CONF_DIR\" || true; fi\nexit 0\n"
    ).encode()
    deb = ar_archive([
        ("debian-binary", b"2.0\n"),
        ("control.tar.gz", tar_gz([("control", control, 0o644), ("postinst", postinst, 0o755)])),
        ("data.tar.gz", tar_gz(data_members)),
    ])
    out_dir.mkdir(parents=True, exist_ok=True)
    path = out_dir / f"uda-b200_{version}-{release}_amd64.deb"
    path.write_bytes(deb)
    return path


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "dist"))
    ap.add_argument("--from-rpm", default=None)
    a = ap.parse_args()
    print(build(Path(a.out), a.from_rpm))
