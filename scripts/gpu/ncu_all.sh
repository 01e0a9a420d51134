#!/bin/bash
# ncu --set full captures of the hot kernels on ONE GPU (each after the same command exited 0 without ncu), plus the per-kernel
# launch list of a timed step. Reports land in gpurun_out/ncu/*.ncu-rep; summarise them with tools/ncu_lines.py and
# `ncu -i ... --page raw --csv`, and copy what matters to profiles/.
set -u
cd "$(dirname "$0")/../.."
O=gpurun_out/ncu; mkdir -p $O
cap() { # name kernel-regex skip count -- command...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 5
  timeout 200 "$@" > $O/$name.plain.log 2>&1 || { echo "plain run of $name failed"; return; }
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c $cnt -f -o $O/$name "$@" > $O/$name.ncu.log 2>&1
  echo "$name: ncu rc=$?"
}
A="--records-per-gpu 20000000 --steps 1 --warmup 1 --no-e2e --no-check"
cap fr_merge_k8 "fr_merge_kernel" 2 1 -- python bench.py $A
cap fr_merge_k64 "fr_merge_kernel" 2 1 -- python bench.py $A --maps-per-gpu 64
cap prepass_k64 "fr_partition_kernel|fr_sample_hist" 2 2 -- python bench.py $A --maps-per-gpu 64
cap prepass_k8 "fr_partition_search" 2 1 -- python bench.py $A
cap mapside "rs_hist|rs_scatter|rs_scan|crc32_kernel|gather_fixed|extract_keys_fixed|teragen" 0 9 -- python bench.py $A --maps-per-gpu 1 --subparts 1 --local-copy push
cap push_lsu "push_kernel" 1 1 -- python bench.py $A --subparts 1 --local-copy push
cap wordcount "partition_|combine_" 0 6 -- python benchmarks/wordcount_bench.py --records-per-gpu 20000000 --vocab 200000 --steps 1 --warmup 0
cap sort_var "ifile_|tiebreak|gather_var|extract_keys_var|merge_tiles|mg_" 0 12 -- python benchmarks/sort_bench.py --records 2000000 --steps 1 --warmup 0 --no-side-index
bash scripts/gpu/kernel_list.sh 8 > $O/kernel_list_k8.txt 2>&1; cat $O/kernel_list_k8.txt | tail -16
ls -la $O/*.ncu-rep | awk '{print $5, $9}'
