#!/bin/bash
# round 2 evidence: launch list of the default bench command (short timed region), then one
# `--set full` capture of a timed fused launch of each step kernel. Every ncu run follows the
# same plain command (&&). Reports are exported to csv here (gpurun returns <= 64 MiB).
mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --no-extra-legs --only-fused --target-ms 20 --steps 20 --warmup 5"
$B > gpurun_out/p2_plain_c2.json 2> gpurun_out/p2_plain_c2.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p2_launches_c2.csv \
    $B > gpurun_out/p2_ncu_launches.log 2>&1
cap() {   # workload, kernel regex, tag
  local C="python bench.py --workload $1 --no-cpu-baseline --only-fused --target-ms 5 --steps 20 --warmup 5"
  $C > gpurun_out/p2_plain_$1.json 2> gpurun_out/p2_plain_$1.err &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -s 2 -c 1 -f -o gpurun_out/p2_$1 \
      $C > gpurun_out/p2_ncu_full_$1.log 2>&1
  tail -n 1 gpurun_out/p2_ncu_full_$1.log
  if [ -f gpurun_out/p2_$1.ncu-rep ]; then
    ncu -i gpurun_out/p2_$1.ncu-rep --page raw --csv > gpurun_out/p2_$1_raw.csv 2>/dev/null
    ncu -i gpurun_out/p2_$1.ncu-rep --page source --csv > gpurun_out/p2_$1_source.csv 2>/dev/null
    ls -la gpurun_out/p2_$1.ncu-rep
    rm -f gpurun_out/p2_$1.ncu-rep
  fi
}
cap c2 ring1_hot
cap c5 ring1_hot
cap c3 ring_ml
cap c4 highway_kernel
cap mg merge_kernel
cap f8 fig8_kernel
ls -la gpurun_out | grep p2_ | head -30
