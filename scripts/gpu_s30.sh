#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 6
python bench.py --no-cpu-baseline > gpurun_out/s30b_c2.json 2> gpurun_out/s30b_c2.err
FS_NO_PDL=1 python bench.py --no-cpu-baseline > gpurun_out/s30b_c2_nopdl.json 2> gpurun_out/s30b_c2_nopdl.err
FS_RING_SHAPE=4 python bench.py --no-cpu-baseline > gpurun_out/s30b_c2_sh4.json 2> gpurun_out/s30b_c2_sh4.err
python bench.py --workload c5 --steps 300 --warmup 10 --no-cpu-baseline > gpurun_out/s30b_c5.json 2> gpurun_out/s30b_c5.err
FS_NO_PDL=1 python bench.py --workload c5 --steps 300 --warmup 10 --no-cpu-baseline > gpurun_out/s30b_c5_nopdl.json 2> gpurun_out/s30b_c5_nopdl.err
python bench.py --workload mah --steps 1000 --warmup 3 --no-cpu-baseline > gpurun_out/s30b_mah.json 2> gpurun_out/s30b_mah.err
tail -c 400 gpurun_out/s30b_mah.err
