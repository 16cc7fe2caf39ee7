#!/bin/bash
# parity tests, then sweep the hot kernel's threads-per-env shapes on c2 and c5
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
show() { python -c "
import json,sys; d=json.load(open(sys.argv[1])); print(sys.argv[2],'fused %.3e single %.3e e2e %.3e e2e1 %.3e us/step %.3f frac %.3f crash %.3f'%(d['value'],d['single_step']['value'],d['e2e']['value'],d['e2e_single_step']['value'],d['ms_per_step']*1e3,d['roofline']['frac'],d['crashed_env_frac']))" $1 "$2"; }
for sh in ${C2_SHAPES:-8 16 32 4}; do FS_RING_SHAPE=$sh python bench.py --steps 1000 --warmup 100 --no-cpu-baseline > gpurun_out/c2_shape$sh.json 2>gpurun_out/c2_shape$sh.err && show gpurun_out/c2_shape$sh.json "c2 shape $sh"; done
for sh in ${C5_SHAPES:-1 2 4 8 16 32}; do FS_RING_SHAPE=$sh python bench.py --workload c5 --steps 300 --warmup 100 --no-cpu-baseline > gpurun_out/c5_shape$sh.json 2>gpurun_out/c5_shape$sh.err && show gpurun_out/c5_shape$sh.json "c5 shape $sh"; done
for sh in ${C2_NONOISE:-32}; do FS_RING_SHAPE=$sh python bench.py --steps 1000 --warmup 100 --no-cpu-baseline --noise 0 > gpurun_out/c2_shape${sh}_nonoise.json 2>/dev/null && show gpurun_out/c2_shape${sh}_nonoise.json "c2 shape $sh noise0"; done
for sh in ${C5_NONOISE:-1}; do FS_RING_SHAPE=$sh python bench.py --workload c5 --steps 300 --warmup 100 --no-cpu-baseline --noise 0 > gpurun_out/c5_shape${sh}_nonoise.json 2>/dev/null && show gpurun_out/c5_shape${sh}_nonoise.json "c5 shape $sh noise0"; done
if [ -n "$NCU" ]; then
S2=${NCU_C2_SHAPE:-32}; S5=${NCU_C5_SHAPE:-1}
FS_RING_SHAPE=$S2 python bench.py --steps 200 --warmup 100 --no-cpu-baseline > gpurun_out/plain_c2.log 2>&1 &&
FS_RING_SHAPE=$S2 ncu --set full --clock-control none --import-source on -k regex:ring1_hot -s 4 -c 2 -o gpurun_out/prof_c2_hot_g$S2 python bench.py --steps 200 --warmup 100 --no-cpu-baseline > gpurun_out/ncu_c2.log 2>&1
FS_RING_SHAPE=$S5 python bench.py --workload c5 --steps 200 --warmup 100 --no-cpu-baseline > gpurun_out/plain_c5.log 2>&1 &&
FS_RING_SHAPE=$S5 ncu --set full --clock-control none --import-source on -k regex:ring1_hot -s 4 -c 2 -o gpurun_out/prof_c5_hot_g$S5 python bench.py --workload c5 --steps 200 --warmup 100 --no-cpu-baseline > gpurun_out/ncu_c5.log 2>&1
fi
