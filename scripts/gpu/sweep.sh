#!/bin/bash
# Sub-partition count x push-SM sweep of the headline step on N GPUs:  scripts/gpu/sweep.sh N "subparts:ctas ..." [extra bench args]
set -u
cd "$(dirname "$0")/../.."
N=$1; shift; CASES=$1; shift
O=gpurun_out/sweep_n$N; mkdir -p $O
for c in $CASES; do
  s=${c%%:*}; p=${c##*:}
  if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555"; fi
  $L bench.py --gpus $N --steps 8 --warmup 3 --no-e2e --subparts $s --push-ctas $p "$@" > $O/s${s}_p$p.json 2> $O/s${s}_p$p.err || echo "FAILED $c"
  python - "$O/s${s}_p$p.json" "N=$N subparts=$s push_ctas=$p" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    ph={k:round(v,2) for k,v in d["phases_ms_max_over_ranks"].items() if k!="timeline_rank0_ms"}
    print(f"{sys.argv[2]:36s} ms/step {d['ms_per_step']:.2f}  GB/s {d['value']:.0f}  {ph}  ok={d['validated']}")
    tl=d["phases_ms_max_over_ranks"].get("timeline_rank0_ms")
    if tl:
        print("      merge start", tl["merge_start_ms"]); print("      merge end  ", tl["merge_end_ms"]); print("      pushed     ", tl["pushed_ms"][-1:])
except Exception as e:
    print(sys.argv[2], "no result", e)
PY
done
