"""Per-phase device timing, NVTX ranges and structured metrics.

The reference has no tracer: it has a levelled log macro, a backtrace in every exception and a handful of wait-time
counters on the reduce task (src/include/IOUtility.h:165, src/CommUtils/IOUtility.cc:479-498,
src/Merger/reducer.h:80-88), plus per-slave ``dstat`` CSVs collected by the regression.
Here every phase is bracketed by CUDA events on the launching stream and an NVTX range (visible in Nsight when a
profiler is attached, free otherwise); the result is one JSON document per run with bytes, GB/s and the fraction of
the measured roofline (MEASURED_PEAKS.json), reduced as the maximum over ranks.
"""
from __future__ import annotations

import contextlib
import json
import os
import time
from pathlib import Path
from typing import Dict, List, Optional

import torch

_ROOT = Path(__file__).resolve().parent.parent.parent


def measured_peaks() -> Dict[str, float]:
    """``{"hbm_gbs": …, "nvlink_gbs": …}`` from MEASURED_PEAKS.json (driver-written), with the profiling recipe's
    fallbacks when the file is absent."""
    out = {"hbm_gbs": 6541.5, "nvlink_gbs": 770.0}
    p = _ROOT / "MEASURED_PEAKS.json"
    try:
        d = json.loads(p.read_text())
        for k in ("hbm_copy_gbs", "hbm_gbs", "copy_gbs", "hbm_copy_GBps"):
            if k in d:
                out["hbm_gbs"] = float(d[k])
        for k in ("nvlink_p2p_gbs", "nvlink_gbs", "p2p_gbs"):
            if k in d:
                out["nvlink_gbs"] = float(d[k])
    except (OSError, ValueError):
        pass
    return out


class PhaseTimer:
    """``with timer.phase("push", bytes_moved=…):`` records a pair of CUDA events and an NVTX range per phase.
    :meth:`summary` synchronises once and turns the events into milliseconds."""

    def __init__(self, device: Optional[torch.device] = None, nvtx: bool = True):
        self.device = device
        self.nvtx = nvtx and torch.cuda.is_available()
        self._open: List[tuple] = []
        self._done: List[tuple] = []

    @contextlib.contextmanager
    def phase(self, name: str, bytes_moved: int = 0, hbm_bytes: int = 0):
        if self.nvtx:
            torch.cuda.nvtx.range_push(name)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        try:
            yield
        finally:
            b.record()
            if self.nvtx:
                torch.cuda.nvtx.range_pop()
            self._done.append((name, a, b, bytes_moved, hbm_bytes))

    def summary(self) -> Dict[str, dict]:
        """``{phase: {"ms": total, "calls": n, "bytes": …, "gbs": …, "hbm_roofline_frac": …}}``."""
        torch.cuda.synchronize(self.device)
        peaks = measured_peaks()
        out: Dict[str, dict] = {}
        for name, a, b, nbytes, hbm in self._done:
            e = out.setdefault(name, {"ms": 0.0, "calls": 0, "bytes": 0, "hbm_bytes": 0})
            e["ms"] += a.elapsed_time(b)
            e["calls"] += 1
            e["bytes"] += nbytes
            e["hbm_bytes"] += hbm
        for e in out.values():
            if e["ms"] > 0 and e["bytes"]:
                e["gbs"] = e["bytes"] / e["ms"] / 1e6
            if e["ms"] > 0 and e["hbm_bytes"]:
                e["hbm_roofline_frac"] = e["hbm_bytes"] / e["ms"] / 1e6 / peaks["hbm_gbs"]
        return out

    def reset(self) -> None:
        self._done.clear()


def reduce_max_over_ranks(summary: Dict[str, dict], device, group=None) -> Dict[str, dict]:
    """Phase times as the maximum over ranks (multi-GPU numbers are never wall-clock, never one rank's view)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return summary
    names = sorted(summary)
    t = torch.tensor([summary[n]["ms"] for n in names], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    out = {n: dict(summary[n]) for n in names}
    for n, v in zip(names, t.cpu().tolist()):
        out[n]["ms_max_over_ranks"] = v
    return out


def write_metrics(name: str, doc: dict, directory: Optional[str] = None) -> Path:
    """Writes ``<directory or gpurun_out>/<name>.json`` (the scratch directory that travels back from the GPU box)."""
    d = Path(directory) if directory else Path(os.environ.get("GRAFT_REPO_ROOT", str(_ROOT))) / "gpurun_out"
    d.mkdir(parents=True, exist_ok=True)
    doc = dict(doc)
    doc.setdefault("written_at", time.strftime("%Y-%m-%dT%H:%M:%S"))
    p = d / f"{name}.json"
    p.write_text(json.dumps(doc, indent=1, sort_keys=True))
    return p
