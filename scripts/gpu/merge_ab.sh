#!/bin/bash
# A/B of the fused record merge on one B200: tile ordering (key bins vs merge tree), tile capacity / CTAs per SM,
# run count. Writes one JSON line per variant to gpurun_out/merge_ab/.
set -u
cd "$(dirname "$0")/../.."
O=gpurun_out/merge_ab; mkdir -p $O
run() { # name, env..., -- args
  local name=$1; shift
  local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e "$@" > $O/$name.json 2> $O/$name.err || echo "FAILED $name" >> $O/failed.txt
  python - "$O/$name.json" "$name" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "ms/step", round(d["ms_per_step"],3), "GB/s", round(d["value"],1), d.get("phases_ms_max_over_ranks"), "ok", d.get("validated"))
except Exception as e:
    print(sys.argv[2], "no result", e)
PY
}
C3=uda_b200/lib/libuda_kernels_c3.so
for k in 8 64; do
  run bins_c2_k$k UDA_X=1 -- --maps-per-gpu $k
  [ -f $C3 ] && run bins_c3_k$k UDA_B200_KERNELS_LIB=$C3 -- --maps-per-gpu $k
done
run tree_c2_k8 UDA_FR_FORCE_TREE=1 -- --maps-per-gpu 8
run bins_c2_k16 UDA_X=1 -- --maps-per-gpu 16
run bins_c2_k32 UDA_X=1 -- --maps-per-gpu 32
run bins_c2_cap384_k8 UDA_FR_CAP=384 -- --maps-per-gpu 8
run bins_c2_cap512_k8 UDA_FR_CAP=512 -- --maps-per-gpu 8
