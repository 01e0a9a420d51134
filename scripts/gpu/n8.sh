#!/bin/bash
# 8-GPU session: the multi-GPU test file, then the headline bench at N=8 (and a push-SM sweep). Run under `gpurun --gpus 8`.
set -u
cd "$(dirname "$0")/../.."
O=gpurun_out/n8; mkdir -p $O
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q -rs > $O/pytest_multi_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_multi_gpu.log
run() { # name N args...
  local name=$1 n=$2; shift 2
  NCCL_DEBUG=INFO timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29577 \
      bench.py --gpus $n --steps 10 --warmup 3 "$@" > $O/$name.json 2> $O/$name.err || echo "FAILED $name"
  grep -c "NCCL INFO.*nranks" $O/$name.err | sed "s/^/   nccl nranks lines: /"
  python - "$O/$name.json" "$name" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    ph=d.get("phases_ms_max_over_ranks") or {}
    print(sys.argv[2], "ms/step", round(d["ms_per_step"],3), "GB/s", round(d["value"],1), {k:(round(v,2) if isinstance(v,float) else "..") for k,v in ph.items()}, "e2e", (d.get("e2e") or {}).get("value"), "ok", d.get("validated"))
except Exception as e:
    print(sys.argv[2], "no result", e)
PY
}
run n8_default 8 --no-e2e
run n8_p48 8 --no-e2e --push-ctas 48
run n8_p24 8 --no-e2e --push-ctas 24
run n8_s8 8 --no-e2e --subparts 8
run n4_default 4 --no-e2e
run n2_default 2 --no-e2e
