#!/bin/bash
# round 2: multi-lane kernel with persistent leaders - parity tests, then the c3 line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_multilane.py tests/test_gpu_sorted.py tests/test_gpu_fuzz.py tests/test_flow_api.py -m gpu -x -q 2>&1 | tail -15
python scripts/lc_count.py
python bench.py --workload c3 --only-fused --steps 3000 --warmup 100 > gpurun_out/r2_c3.json 2> gpurun_out/r2_c3.err; echo "c3 rc=$?"; tail -c 300 gpurun_out/r2_c3.err
FS_ML_GENERIC=1 python bench.py --workload c3 --only-fused --steps 3000 --warmup 100 > gpurun_out/r2_c3_generic.json 2> gpurun_out/r2_c3_generic.err; echo "c3g rc=$?"
python -c "
import json
for f in ('r2_c3','r2_c3_generic'):
    d=json.load(open('gpurun_out/%s.json'%f)); print(f, d['value'], d['ms_per_step'], d['crashed_env_frac'])
"
timeout 900 python -m pytest tests/test_gpu_hot_kernel.py -m gpu -x -q 2>&1 | tail -8
