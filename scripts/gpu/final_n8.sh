#!/bin/bash
# Final 8-GPU session of a round (run under `gpurun --gpus 8`): headline bench with the end-to-end number, one push-SM
# variant, the host copy ceiling, the all-to-all sweep (config 5) and the 4-GPU point. Every command has its own time limit.
set -u
cd "$(dirname "$0")/../.."
O=gpurun_out/final; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    ph={k:(round(v,2) if isinstance(v,float) else "..") for k,v in (d.get("phases_ms_max_over_ranks") or {}).items()}
    print(sys.argv[2], "ms/step", round(d["ms_per_step"],3), "GB/s", round(d["value"],1), ph, "e2e", (d.get("e2e") or {}).get("value"), "ok", d.get("validated"))
except Exception as e:
    print(sys.argv[2], "no result", e)
PY
}
NCCL_DEBUG=INFO timeout 300 $TR --nproc-per-node 8 --master-port 29581 bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_n8.json 2> $O/bench_n8.err; show $O/bench_n8.json n8_default
grep -c "NCCL INFO.*nranks 8" $O/bench_n8.err | sed "s/^/   nccl 'nranks 8' lines: /"
timeout 200 $TR --nproc-per-node 8 --master-port 29582 bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e --push-ctas 48 > $O/bench_n8_p48.json 2> $O/bench_n8_p48.err; show $O/bench_n8_p48.json n8_p48
timeout 120 $TR --nproc-per-node 8 --master-port 29583 benchmarks/host_copy_ceiling.py --reps 2 > $O/host_copy_ceiling_n8.json 2> $O/hcc8.err; cat $O/host_copy_ceiling_n8.json | cut -c1-400
timeout 150 $TR --nproc-per-node 8 --master-port 29584 benchmarks/a2a_sweep.py --iters 5 > $O/a2a_n8.jsonl 2> $O/a2a_n8.err; cut -c1-260 $O/a2a_n8.jsonl
timeout 200 $TR --nproc-per-node 4 --master-port 29585 bench.py --gpus 4 --steps 10 --warmup 3 --no-e2e > $O/bench_n4.json 2> $O/bench_n4.err; show $O/bench_n4.json n4_default
