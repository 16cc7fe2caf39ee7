#!/bin/bash
# single-step (K = 1 launches) kernel time against threads per env: bench `single_step` leg
mkdir -p gpurun_out
for wl in c5 c2; do
  for sh in 1 2 4 8 16 32; do
    FS_RING_SHAPE=$sh python bench.py --workload $wl --steps 300 --warmup 10 --no-cpu-baseline \
      > gpurun_out/ss_${wl}_${sh}.json 2> gpurun_out/ss_${wl}_${sh}.err
  done
done
