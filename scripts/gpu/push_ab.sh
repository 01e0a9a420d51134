#!/bin/bash
# A/B of the copy-engine push pipeline shape (stages x chunk x loads-ahead x warps) on N GPUs (default 2).
set -u
cd "$(dirname "$0")/../.."
N=${1:-2}; O=gpurun_out/push_ab; mkdir -p $O
for lib in "" uda_b200/lib/libuda_kernels_tp_*.so; do
  name=$(basename "${lib:-default}" .so)
  UDA_B200_KERNELS_LIB=$lib python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
      bench.py --gpus $N --steps 6 --warmup 3 --no-e2e ${PUSH_ARGS:-} > $O/$name.json 2> $O/$name.err || echo "FAILED $name"
  python - "$O/$name.json" "$name" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    ph={k:round(v,2) for k,v in d["phases_ms_max_over_ranks"].items() if k!="timeline_rank0_ms"}
    print(f"{sys.argv[2]:36s} ms/step {d['ms_per_step']:.2f}  GB/s {d['value']:.0f}  {ph}  ok={d['validated']}")
except Exception as e:
    print(sys.argv[2], "no result", e)
PY
done
