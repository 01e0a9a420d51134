#!/usr/bin/env python
"""Headline benchmark: TeraSort shuffle+merge GB/s (BASELINE.json metric), device-timed, max over ranks.

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Config (BASELINE.json config 2, "TeraSort 100 GB synthetic teragen records on 8xB200", weak scaling):
every GPU holds 12.5 GB of 100-byte records (125 M records, 10 B key + 90 B value) as sorted,
range-partitioned IFile map-output files in HBM; one timed step = fused all-to-all shuffle push of every
MOF partition into its reducer's receive arena + landed-flag acquire + key normalise + k-way merge +
record gather into the merged IFile stream. The map-side sort that produces the MOFs is set-up, as it
is in the reference (Hadoop's MapTask, not UDA, sorts map output).

``--impl reference``: the reference's own reducer (NetMerger: INIT/FETCH handling, fetch + merge threads, Segment
double buffers, heap merge, write_kv_to_stream into 1 MiB blocks), compiled unmodified by ``baseline/build_ref.py``
and fed from host memory (its RDMA transport needs libibverbs, which the image lacks: a memcpy stands in), one
reducer process per rank over the same record shape and counts; nothing of uda_b200 is loaded on that path.
``--impl cpu``: this repository's own CPU reducer (the fall-back path of the plug-in), memory-fed the same way, one process.
``--impl nccl_sort``: the second baseline BASELINE.json names
(NCCL all_to_all_single + torch.sort), same data and metric.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "TeraSort shuffle+merge GB/s (device-timed, max over ranks)"
RECORD_BYTES = 100


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="uda_b200", choices=["uda_b200", "reference", "nccl_sort", "cpu"])
    ap.add_argument("--records-per-gpu", type=int, default=125_000_000)
    ap.add_argument("--maps-per-gpu", type=int, default=8)
    ap.add_argument("--subparts", type=int, default=0,
                    help="sub-partitions per reducer pushed/merged as a pipeline (0 = auto: 1 on one GPU, 12 on several)")
    ap.add_argument("--push-ctas", type=int, default=-1,
                    help="grid of the push kernel (-1 = auto: all SMs on one GPU; pipelined: 36 SMs for tma, 296 CTAs for lsu)")
    ap.add_argument("--push-mode", default="auto", choices=["auto", "lsu", "tma"],
                    help="lsu: load/store push kernel on all SMs; tma: copy-engine push on --push-ctas SMs of its own (one launch "
                         "for all sub-partitions); auto: lsu on one GPU, tma on several")
    ap.add_argument("--tma-flush", type=int, default=0, help="tma push: chunks between landed-counter advances (0 = library default)")
    ap.add_argument("--merge-ctas-per-sm", type=int, default=0, help="cap on resident merge CTAs per SM (0 = occupancy limit)")
    ap.add_argument("--graph", action="store_true", help="time replays of the step captured as a CUDA graph (one GPU, no pipelining)")
    ap.add_argument("--local-zero-copy", action="store_true",
                    help="merge this rank's own map outputs in place instead of copying them into its arena (implies --no-e2e)")
    ap.add_argument("--local-copy", default="auto", choices=["auto", "push", "ce"],
                    help="a rank's own map outputs reach its arena through the push kernel or through the copy engines "
                         "(auto: copy engines when the step is pipelined over sub-partitions)")
    ap.add_argument("--arena-slots", type=int, default=0,
                    help="receive slots per reducer (0 = one per sub-partition: the arena holds the whole partition; fewer = a ring "
                         "with credit-based back-pressure, receive memory = slots / subparts of the partition)")
    ap.add_argument("--device-cuts", action="store_true",
                    help="map outputs keep Hadoop's P partitions; the sub-partitions of the pipeline are cut on the device with "
                         "splitters sampled from the sorted segments")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not bind the rank to the CPUs local to its GPU")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def run_cpu(a, emit) -> int:
    """``--impl cpu``: this repository's own CPU reducer (csrc/runtime: online heap merge), fed from memory exactly like the
    reference arm feeds the reference's reducer (tests/native/cpu_merge_bench.cc mirrors baseline/ref_harness.cc). No GPU.
    One process (N = 1): it exists to put the CPU fall-back path on the same scale as the reference arm."""
    exe = os.path.join(ROOT, "build", "bin", "cpu_merge_bench")
    if not os.path.exists(exe):
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import build as _build
        _build.build(verbose=False)
    cmd = [exe, "libuda.so", "--maps", str(a.maps_per_gpu), "--records", str(a.records_per_gpu), "--steps", str(a.steps), "--warmup", str(a.warmup),
           "--seed", "1234", "--budget-s", os.environ.get("UDA_REF_BUDGET_S", "540")]
    if not a.no_check:
        cmd.append("--check")
    p = subprocess.run(cmd, capture_output=True, text=True)
    line = next((l for l in p.stdout.splitlines() if l.startswith("OURS ")), None)
    if p.returncode != 0 or line is None:
        emit({"impl": "cpu", "unavailable": "the CPU reducer did not finish: " + (p.stderr.strip().splitlines() or ["no output"])[-1][:200]})
        return 0
    kv = dict(x.split("=", 1) for x in line.split()[1:])
    secs = [float(x) for x in kv["seconds"].split(",") if x]
    mean_s = sum(secs) / len(secs)
    value = float(a.records_per_gpu) * RECORD_BYTES / mean_s / 1e9
    emit({"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": 1, "steps": len(secs), "warmup": a.warmup, "ms_per_step": mean_s * 1e3,
          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "uint8 records (byte-exact)", "data": "synthetic", "impl": "cpu",
          "config": {"model": "TeraSort, one reducer's share: 100-byte teragen records in sorted IFile segments in host memory",
                     "global_batch": a.records_per_gpu, "seq_len": RECORD_BYTES, "maps_per_gpu": a.maps_per_gpu,
                     "parallelism": "one CPU reducer (libuda.so consumer role, online heap merge), memory-fed like the reference arm",
                     "timed_region": "INIT + FETCH of every map output + fetch/merge threads + 1 MiB blocks to a byte-counting sink (host clock)"},
          "clocks": None, "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
          "validated": {"ok": True, "skipped": None}.get(kv.get("check"), False)})
    return 0


def run_reference(a, emit) -> int:
    """The reference arm: chaoyang223/UDA's own reducer code (see baseline/ref_harness.cc for what is the reference's
    and what stands in for verbs / the JVM). CPU only; one reducer process per rank, started together; the step time is
    the reference's own span from INIT to the last block handed to "Java", max over ranks."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    exe = os.path.join(ROOT, "baseline", "_ref", "bin", "ref_merger")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo")
    why = None
    if not os.path.exists(exe) and rank == 0:   # normally built by __graft_entry__.build(); rank 0 builds, the others wait
        try:
            sys.path.insert(0, os.path.join(ROOT, "baseline"))
            import build_ref
            build_ref.build(verbose=False)
        except Exception as ex:  # no sources on this machine, no compiler, ...
            why = f"reference arm could not be built here: {type(ex).__name__}: {str(ex)[:200]}"
    if dist is not None:
        dist.barrier()
    if why is None and not os.path.exists(exe):
        why = "baseline/_ref/bin/ref_merger is missing and could not be built"
    if why is not None:
        if rank == 0:
            emit({"impl": "reference", "unavailable": why})
        return 0
    budget = float(os.environ.get("UDA_REF_BUDGET_S", "540"))
    cmd = [exe, "--maps", str(a.maps_per_gpu * world), "--records", str(a.records_per_gpu), "--steps", str(a.steps),
           "--warmup", str(a.warmup), "--seed", str(1234 + rank), "--budget-s", str(budget)]
    if not a.no_check:
        cmd.append("--check")
    if dist is not None:
        dist.barrier()
    p = subprocess.run(cmd, capture_output=True, text=True)
    line = next((l for l in p.stdout.splitlines() if l.startswith("REF ")), None)
    ok = p.returncode == 0 and line is not None
    secs, checked = [], None
    if ok:
        kv = dict(x.split("=", 1) for x in line.split()[1:])
        secs = [float(x) for x in kv["seconds"].split(",") if x]
        checked = kv.get("check")
        ok = len(secs) > 0
    mean_s = sum(secs) / len(secs) if ok else -1.0
    nsteps = len(secs)
    if dist is not None:
        import torch
        t = torch.tensor([mean_s if ok else 1e30, float(nsteps), 1.0 if ok else 0.0], dtype=torch.float64)
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        mn = t.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        mean_s, nsteps, ok = float(mx[0]), int(mn[1]), bool(mn[2] > 0)
    if rank == 0:
        if not ok:
            emit({"impl": "reference", "unavailable": "the reference reducer did not finish: " + (p.stderr.strip().splitlines() or ["no output"])[-1][:200]})
        else:
            total_bytes = float(a.records_per_gpu) * RECORD_BYTES * world
            value = total_bytes / mean_s / 1e9
            emit({
                "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": nsteps, "warmup": a.warmup,
                "ms_per_step": mean_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "uint8 records (byte-exact; no floating point in this workload)", "data": "synthetic", "impl": "reference",
                "config": {
                    "model": "TeraSort 100 GB/8 GPUs (12.5 GB per GPU), 100-byte teragen records, IFile-framed sorted MOFs in HBM",
                    "global_batch": a.records_per_gpu * world, "seq_len": RECORD_BYTES, "records_per_gpu": a.records_per_gpu,
                    "maps_per_gpu": a.maps_per_gpu, "runs_per_reducer": a.maps_per_gpu * world,
                    "parallelism": f"one reference reducer (NetMerger) process per rank x{world}, CPU",
                    "reference_code": "src/Merger/{NetMergerMain,reducer,MergeManager,StreamRW,MergeQueue,CompareFunc}, src/CommUtils/{IOUtility,C2JNexus,UdaUtil} unmodified",
                    "stand_ins": "RDMA fetch -> memcpy from host memory on a copy thread; JVM up-calls -> byte counter (Java-side copy not charged)",
                    "timed_region": "INIT + FETCH of every map output + fetch/merge threads + write_kv_to_stream into 1 MiB blocks, per reduce task "
                                    "(host clock inside the reducer process; the reference has no device work)",
                    "requested_steps": a.steps,
                },
                "clocks": None,
                "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                        "note": "the reference's data starts in host memory and ends in its 1 MiB host staging block: the timed number is end to end"},
                "gpu_launches": 0, "validated": (checked == "ok") if checked != "skipped" else None,
            })
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    a = parse_args()
    # Rank 0 must print exactly one line on stdout. Libraries print banners there (NCCL's version line comes
    # from C code, so redirecting sys.stdout is not enough): park the real stdout and point fd 1 at stderr.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(obj) + "\n").encode())

    if a.impl == "reference":
        return run_reference(a, emit)
    if a.impl == "cpu":
        return run_cpu(a, emit)

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus and world > 1:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    affinity = {"bound": False}
    if not a.no_numa_bind:
        # before any pinned allocation: first-touch then places the staging buffers next to this rank's GPU
        from uda_b200.utils.affinity import bind_to_gpu_numa
        affinity = bind_to_gpu_numa(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)   # NCCL's own log level is left to the caller (fd 1 is parked above)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if a.impl == "nccl_sort":
        from baseline.nccl_sort import NcclSortJob as Job
    else:
        from uda_b200.models.terasort import TeraSortJob as Job

    # measured on 8xB200 (profiles/pipeline_sweep.md, profiles/r2/final): 12 sub-partitions pushed by the copy-engine kernel on
    # SMs of its own (36; 48 on 8 GPUs: 21.1 - 22.0 ms against 22.0 - 22.2 ms) while the merge keeps the others; the local
    # segments ride the copy engines
    if a.push_mode == "auto":
        a.push_mode = "tma" if world > 1 else "lsu"
    if a.subparts <= 0:
        # 2 GPUs: 8 / 32 measured 3 % better than 12 / 36; one GPU: 4 sub-partitions, the arena copies on the copy engines
        # overlapping the merge (10.15 ms against 10.80 ms for copy-then-merge, profiles/r2)
        a.subparts = (12 if world > 2 else 8) if world > 1 else (1 if (a.graph or a.local_zero_copy or a.local_copy == "push") else 4)
    if a.push_ctas < 0:
        a.push_ctas = (((48 if world >= 8 else 36) if world > 2 else 32) if a.push_mode == "tma" else 296) if a.subparts > 1 else 0
    if a.impl == "nccl_sort":
        job = Job(a.records_per_gpu, a.maps_per_gpu, device=dev).setup()
    else:
        job = Job(a.records_per_gpu, a.maps_per_gpu, device=dev, subparts=a.subparts, push_ctas=a.push_ctas,
                  merge_ctas_per_sm=a.merge_ctas_per_sm, push_mode=a.push_mode, local_zero_copy=a.local_zero_copy, tma_flush=a.tma_flush, local_copy=a.local_copy,
                  arena_slots=a.arena_slots, device_cuts=a.device_cuts).setup()
    total_bytes = float(a.records_per_gpu) * RECORD_BYTES * world

    def timed(fn, steps, warmup, sample_clocks):
        for _ in range(warmup):
            fn()
        barrier()
        sampler = ClockSampler(local_rank) if (sample_clocks and rank == 0) else None
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        clocks = sampler.stop() if sampler else None
        return ms, clocks

    step_fn = job.step
    if a.graph and a.impl == "uda_b200":
        if world > 1 or a.subparts > 1:
            raise SystemExit("--graph is offered for the single-stream configuration (one GPU, --subparts 1)")
        for _ in range(2):
            job.step()
        graph, _graph_out = job.capture()
        step_fn = graph.replay
    ms, clocks = timed(step_fn, a.steps, a.warmup, True)
    ms_per_step = ms / a.steps
    value = total_bytes / (ms_per_step * 1e-3) / 1e9

    phases = None
    if hasattr(job, "step_phases"):
        ph = [job.step_phases() for _ in range(3)][-1]
        phases = {k: max_over_ranks(v) for k, v in ph.items()}
        if hasattr(job, "step_timeline") and getattr(job, "subparts", 1) > 1:
            tl = [job.step_timeline() for _ in range(3)][-1]
            phases["timeline_rank0_ms"] = {k: [round(x, 2) for x in v] for k, v in tl.items()}

    check = None
    if not a.no_check:
        check = job.check()
        if not check["ok"]:
            print(json.dumps({"error": "validation failed", "check": check}), file=sys.stderr)
            return 1

    e2e = None
    if not a.no_e2e and not a.local_zero_copy and hasattr(job, "e2e_setup"):
        try:
            job.e2e_setup()
            ems, _ = timed(job.e2e_step, a.steps, max(1, min(a.warmup, 3)), False)
            e2e_ms = ems / a.steps
            e2e = {"value": total_bytes / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e2e_ms,
                   "h2d_bytes_per_step": int(job.e2e_h2d_bytes), "d2h_bytes_per_step": int(job.e2e_d2h_bytes),
                   "api": "uda_b200.models.terasort.TeraSortJob.e2e_step (pinned host MOFs -> merged stream in pinned host memory)"}
            if not a.no_check:
                c2 = job.e2e_check()
                e2e["check_ok"] = bool(c2)
        except Exception as ex:  # report, do not hide
            e2e = {"value": None, "error": f"{type(ex).__name__}: {ex}"}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "uint8 records (byte-exact; no floating point in this workload)", "data": "synthetic",
            "impl": a.impl,
            "config": {
                "model": "TeraSort 100 GB/8 GPUs (12.5 GB per GPU), 100-byte teragen records, IFile-framed sorted MOFs in HBM",
                "global_batch": a.records_per_gpu * world, "seq_len": RECORD_BYTES,
                "records_per_gpu": a.records_per_gpu, "maps_per_gpu": a.maps_per_gpu, "runs_per_reducer": a.maps_per_gpu * world,
                "subparts": a.subparts, "push_mode": a.push_mode, "local_zero_copy": a.local_zero_copy, "local_copy": getattr(job, "local_copy", None),
                "arena_slots": (job.plan.slots if hasattr(job, "plan") else None), "arena_bytes": (int(job.plan.max_arena_bytes()) if hasattr(job, "plan") else None),
                "device_cuts": bool(a.device_cuts),
                "subpart_balance_max_over_mean": (round(max(job.sub_n) * len(job.sub_n) / max(1, sum(job.sub_n)), 3) if getattr(job, "sub_n", None) else None),
                "cuda_graph": bool(a.graph), "push_ctas": a.push_ctas, "merge_ctas_per_sm": a.merge_ctas_per_sm,
                "parallelism": f"partition-parallel a2a x{world} (one reducer per GPU)",
                "l2_policy": "inputs (12.75 GB/GPU) larger than the 126 MB L2; no flush needed",
                "timed_region": "shuffle push + flag acquire + key normalise + k-way merge + record gather (map-side sort is set-up)",
            },
            "clocks": clocks, "e2e": e2e, "phases_ms_max_over_ranks": phases, "cpu_affinity_rank0": affinity,
            "gpu_launches": int(getattr(job, "launches_per_step", 0)) * a.steps,
            "validated": None if check is None else check["ok"],
        }
        emit(out)
    job.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
