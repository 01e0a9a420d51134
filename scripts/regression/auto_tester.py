#!/usr/bin/env python
"""Single-node regression driver: the counterpart of the reference's 12-phase cluster regression
(scripts/regression/autoTester.sh:186-350 — prepare, input, control, configure, setup-cluster, setup-tests, execute,
collect, analyze, view, distribute, exit) for a box with 1..8 B200s instead of a Hadoop cluster.

A test matrix (CSV, the reference parses one with scripts/regression/parseTests.awk) names the workload, the number
of GPUs and the knobs; every row is run in its own process group under torchrun, its JSON result and logs are
collected into a per-run directory, the analyzer applies the same kinds of verdicts as
scripts/regression/testStatusAnalyzer.sh (validation oracle, "did the accelerated path really run", fallback
expected / forbidden, no new core files) and an HTML + JSON report is written.

    python scripts/regression/auto_tester.py --matrix scripts/regression/matrix.csv --out gpurun_out/regression

Phases can be selected with --phases (comma separated) the way the reference's -b/-s/-e/-c/-a/-v flags do.
"""
from __future__ import annotations

import argparse
import csv
import html
import json
import os
import shutil
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent.parent
PHASES = ["prepare", "input", "control", "configure", "setup", "setup-tests", "execute", "collect", "analyze", "view", "distribute", "exit"]

WORKLOADS = {
    # name -> (script relative to the repo, argv builder)
    "terasort": ("bench.py", lambda r, n: ["--gpus", str(n), "--steps", r.get("steps", "3"), "--warmup", r.get("warmup", "3"),
                                          "--records-per-gpu", r.get("records", "2000000"), "--maps-per-gpu", r.get("maps", "4"),
                                          "--no-e2e", *(r.get("extra", "").split())]),
    "wordcount": ("benchmarks/wordcount_bench.py", lambda r, n: ["--records-per-gpu", r.get("records", "2000000"), *(r.get("extra", "").split())]),
    "spill": ("benchmarks/spill_bench.py", lambda r, n: [*(r.get("extra", "").split())]),
    "sort": ("benchmarks/sort_bench.py", lambda r, n: ["--records", r.get("records", "200000"), *(r.get("extra", "").split())]),
    "plumbing": ("benchmarks/plumbing_bench.py", lambda r, n: [*(r.get("extra", "").split())]),
    # the reference arm of bench.py: the reference's own reducer code, CPU only (baseline/build_ref.py)
    "reference": ("bench.py", lambda r, n: ["--impl", "reference", "--gpus", str(n), "--steps", r.get("steps") or "2", "--warmup", r.get("warmup") or "1",
                                           "--records-per-gpu", r.get("records") or "2000000", "--maps-per-gpu", r.get("maps") or "4",
                                           *(r.get("extra", "").split())]),
}
CPU_WORKLOADS = ("plumbing", "reference")   # rows that run without a GPU


def row_env(t):
    """Optional `env` column: `NAME=value;NAME2=value2`, `$ROOT` expands to the repository (e.g. the host-only mock of the
    kernel library for merge approach 3: UDA_KERNELS_LIB=$ROOT/build/bin/libuda_kernels_mock.so)."""
    out = {}
    for item in (t.get("env") or "").split(";"):
        if "=" in item:
            k, v = item.split("=", 1)
            out[k.strip()] = v.strip().replace("$ROOT", str(ROOT))
    return out


def sh(cmd, **kw):
    return subprocess.run(cmd, capture_output=True, text=True, **kw)


def phase_prepare(ctx):
    """Pre-flight checks (scripts/regression/preReq.sh:4-14: RAM, CPU idle, swap, disk space, permissions)."""
    p = sh([sys.executable, str(ROOT / "scripts" / "regression" / "pre_req.py"), "--json"])
    ctx["prereq"] = json.loads(p.stdout) if p.returncode in (0, 1) and p.stdout.strip().startswith("{") else {"error": p.stderr[-500:]}
    ctx["out"].mkdir(parents=True, exist_ok=True)


def phase_input(ctx):
    rows = []
    with open(ctx["matrix"]) as f:
        for row in csv.DictReader(l for l in f if not l.lstrip().startswith("#")):
            if row.get("enabled", "1").strip() not in ("0", "false", "no"):
                rows.append({k.strip(): (v or "").strip() for k, v in row.items() if isinstance(k, str) and isinstance(v, (str, type(None)))})
    ctx["tests"] = rows


def phase_control(ctx):
    import torch

    ctx["ngpu"] = torch.cuda.device_count() if torch.cuda.is_available() else 0
    for t in ctx["tests"]:
        need = int(t.get("gpus", "1") or 1)
        t["skip"] = "" if need <= ctx["ngpu"] or t["workload"] in CPU_WORKLOADS else f"needs {need} GPUs, {ctx['ngpu']} visible"


def phase_configure(ctx):
    """Per-test environment (the reference writes exports.sh + Hadoop XML per test, parseTests.awk)."""
    for i, t in enumerate(ctx["tests"]):
        d = ctx["out"] / f"test_{i:03d}_{t['name']}"
        d.mkdir(parents=True, exist_ok=True)
        t["dir"] = str(d)
        conf = {k[5:]: v for k, v in t.items() if k.startswith("conf.") and v}
        (d / "conf.json").write_text(json.dumps(conf, indent=1))
        (d / "exports.sh").write_text("".join(f"export {k}={v!r}\n" for k, v in sorted(t.items()) if isinstance(v, str)))


def phase_setup(ctx):
    """Build the native libraries once (the reference installs the rpm on every node here)."""
    p = sh([sys.executable, str(ROOT / "tools" / "build.py")], cwd=ROOT)
    ctx["build_ok"] = p.returncode == 0
    (ctx["out"] / "build.log").write_text(p.stdout + p.stderr)


def phase_setup_tests(ctx):
    ctx["cores_before"] = sorted(str(p) for p in Path("/tmp").glob("core*"))


def phase_execute(ctx):
    for t in ctx["tests"]:
        if t.get("skip") or not ctx.get("build_ok", True):
            continue
        script, argv = WORKLOADS[t["workload"]]
        n = int(t.get("gpus", "1") or 1)
        args = argv(t, n)
        if n > 1:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
                   "--master-port", str(29600 + n), str(ROOT / script), *args]
        else:
            cmd = [sys.executable, str(ROOT / script), *args]
        t0 = time.time()
        try:
            p = sh(cmd, cwd=ROOT, timeout=int(t.get("timeout", "900") or 900), env={**os.environ, "UDA_REGRESSION_DIR": t["dir"], **row_env(t)})
            rc, out, err = p.returncode, p.stdout, p.stderr
        except subprocess.TimeoutExpired as e:
            rc, out, err = 124, (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "timeout"
        t["rc"], t["seconds"] = rc, round(time.time() - t0, 2)
        d = Path(t["dir"])
        (d / "stdout.log").write_text(out)
        (d / "stderr.log").write_text(err)
        (d / "cmd.txt").write_text(" ".join(cmd))


def phase_collect(ctx):
    """Log collection (utils/master/job-log-collector.sh in the reference): the job's result line + log snippets."""
    for t in ctx["tests"]:
        if "rc" not in t:
            continue
        d = Path(t["dir"])
        res = None
        for line in reversed((d / "stdout.log").read_text().splitlines()):
            line = line.strip()
            if line.startswith("{") and line.endswith("}"):
                try:
                    res = json.loads(line)
                    break
                except ValueError:
                    pass
        t["result"] = res
        sh([sys.executable, str(ROOT / "utils" / "job_log_collector.py"), "--dir", str(d), "--out", str(d / "collected.log")])


def phase_analyze(ctx):
    """Verdicts (scripts/regression/testStatusAnalyzer.sh:168-351)."""
    cores_after = sorted(str(p) for p in Path("/tmp").glob("core*"))
    new_cores = sorted(set(cores_after) - set(ctx.get("cores_before", [])))
    for t in ctx["tests"]:
        if t.get("skip"):
            t["verdict"], t["why"] = "SKIPPED", t["skip"]
            continue
        if "rc" not in t:
            t["verdict"], t["why"] = "NOT RUN", "build failed" if not ctx.get("build_ok", True) else "phase execute not selected"
            continue
        why = []
        r = t.get("result") or {}
        expect_fallback = t.get("expect", "") == "fallback"
        log = (Path(t["dir"]) / "collected.log").read_text() if (Path(t["dir"]) / "collected.log").exists() else ""
        fell_back = "fallback" in log.lower() and "will try to fallback" in log
        if t["rc"] != 0:
            why.append(f"exit code {t['rc']}")
        if not r:
            why.append("no JSON result line")
        if r and r.get("validated") is False:
            why.append("validation oracle failed")
        if r and r.get("ok") is False:
            why.append("job reported ok=false")
        if expect_fallback != fell_back:
            why.append("fallback expected but not seen" if expect_fallback else "unexpected fallback")
        if r and t["workload"] == "terasort" and not r.get("gpu_launches"):
            why.append("no kernels of the native library were launched")
        if new_cores:
            why.append(f"new core files: {new_cores[:3]}")
        t["verdict"], t["why"] = ("PASSED", "") if not why else ("FAILED", "; ".join(why))


def phase_view(ctx):
    rows = []
    for t in ctx["tests"]:
        r = t.get("result") or {}
        rows.append((t["name"], t["workload"], t.get("gpus", "1"), t.get("verdict", ""), t.get("seconds", ""), r.get("value", ""), r.get("unit", ""),
                     t.get("why", "")))
    w = max(len(r[0]) for r in rows) if rows else 10
    print(f"{'test':<{w}}  workload   gpus  verdict   sec      value")
    for r in rows:
        print(f"{r[0]:<{w}}  {r[1]:<9}  {r[2]:<4}  {r[3]:<8}  {str(r[4]):<7}  {r[5]} {r[6]}  {r[7]}")
    body = "".join("<tr>" + "".join(f"<td>{html.escape(str(c))}</td>" for c in r) + "</tr>" for r in rows)
    (ctx["out"] / "report.html").write_text(
        "<html><body><h1>uda_b200 regression</h1><table border=1><tr><th>test</th><th>workload</th><th>gpus</th><th>verdict</th>"
        f"<th>seconds</th><th>value</th><th>unit</th><th>why</th></tr>{body}</table><pre>{html.escape(json.dumps(ctx.get('prereq', {}), indent=1))}</pre></body></html>")
    (ctx["out"] / "report.json").write_text(json.dumps({"tests": ctx["tests"], "prereq": ctx.get("prereq"), "ngpu": ctx.get("ngpu")}, indent=1, default=str))


def phase_distribute(ctx):
    """The reference mails the report; here it is copied next to the other evidence when --publish is given."""
    if ctx.get("publish"):
        dst = ROOT / "profiles" / "regression"
        dst.mkdir(parents=True, exist_ok=True)
        for f in ("report.html", "report.json"):
            if (ctx["out"] / f).exists():
                shutil.copy(ctx["out"] / f, dst / f)


def phase_exit(ctx):
    bad = [t["name"] for t in ctx["tests"] if t.get("verdict") == "FAILED"]
    ctx["exit_code"] = 1 if bad else 0
    if bad:
        print("FAILED:", ", ".join(bad))


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--matrix", default=str(ROOT / "scripts" / "regression" / "matrix.csv"))
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "regression"))
    ap.add_argument("--phases", default=",".join(PHASES))
    ap.add_argument("--publish", action="store_true")
    a = ap.parse_args()
    ctx = {"matrix": a.matrix, "out": Path(a.out), "publish": a.publish, "tests": []}
    ctx["out"].mkdir(parents=True, exist_ok=True)
    fns = {"prepare": phase_prepare, "input": phase_input, "control": phase_control, "configure": phase_configure, "setup": phase_setup,
           "setup-tests": phase_setup_tests, "execute": phase_execute, "collect": phase_collect, "analyze": phase_analyze,
           "view": phase_view, "distribute": phase_distribute, "exit": phase_exit}
    for ph in PHASES:
        if ph in a.phases.split(","):
            fns[ph](ctx)
    return ctx.get("exit_code", 0)


if __name__ == "__main__":
    sys.exit(main())
