#!/usr/bin/env python
"""Scheduler-facing wrapper around the regression driver (counterpart of the reference's cases/uda_wrapper.py, which
subclasses a MARS `StandaloneWrapper` and returns the shell command of the run stage). The MARS framework is internal to
the reference's vendor and not available here, so this wrapper is self-contained: it maps the case arguments to a
command, runs it, prints SUCCESS / FAILURE like cases/runRegression_1.sh does, and exits with the command's status.

    cases/uda_b200_wrapper.py --rows cpu|all [--out DIR]
"""
import argparse
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def get_command(rows: str, out: str):
    cmd = [sys.executable, os.path.join(ROOT, "scripts", "regression", "auto_tester.py"), "--out", out]
    if rows == "cpu":
        cmd += ["--matrix", os.path.join(ROOT, "scripts", "regression_limited", "matrix_cpu.csv")]
    return cmd


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", default="cpu", choices=["cpu", "all"])
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "regression_case"))
    a = ap.parse_args()
    rc = subprocess.call(get_command(a.rows, a.out), cwd=ROOT)
    print("SUCCESS!" if rc == 0 else "FAILURE!")
    return rc


if __name__ == "__main__":
    sys.exit(main())
