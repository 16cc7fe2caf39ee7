#!/bin/bash
# round profile: plain bench (must exit 0) -> ncu launch list of the same command -> one
# `--set full` capture of the dominant kernel for c2 (default workload) and c5.
mkdir -p gpurun_out
python bench.py --steps 300 --warmup 100 --no-cpu-baseline > gpurun_out/prof_plain_c2.json 2> gpurun_out/prof_plain_c2.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
    python bench.py --steps 300 --warmup 100 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ring1_hot -s 3 -c 1 -o gpurun_out/prof_c2_final \
    python bench.py --steps 1000 --warmup 500 --no-cpu-baseline --only-fused > gpurun_out/ncu_full_c2.log 2>&1
python bench.py --workload c5 --steps 200 --warmup 100 --no-cpu-baseline --only-fused > gpurun_out/prof_plain_c5.json 2> gpurun_out/prof_plain_c5.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:ring1_hot -s 3 -c 1 -o gpurun_out/prof_c5_final \
    python bench.py --workload c5 --steps 200 --warmup 100 --no-cpu-baseline --only-fused > gpurun_out/ncu_full_c5.log 2>&1
# one SINGLE-STEP launch (k = 1) of c5: the HBM-bound variant
FS_NO_GRAPH=1 FS_NO_PDL=1 ncu --set full --clock-control none --import-source on -k regex:ring1_hot -s 12 -c 1 -o gpurun_out/prof_c5_single \
    python bench.py --workload c5 --steps 40 --warmup 10 --fuse 40 --no-cpu-baseline > gpurun_out/ncu_full_c5_single.log 2>&1
for f in gpurun_out/ncu_full_c2.log gpurun_out/ncu_full_c5.log; do tail -n 2 $f; done
