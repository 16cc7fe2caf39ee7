#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_highway.py tests/test_highway_ma.py tests/test_gpu_fuzz.py tests/test_flow_api.py -m gpu -x -q -k "highway or hw or Highway" 2>&1 | tail -4
for w in c4 mah; do
  python bench.py --workload $w --only-fused --steps 3000 > gpurun_out/r2_$w.json 2> gpurun_out/r2_$w.err
  python -c "
import json; d=json.load(open('gpurun_out/r2_$w.json')); print('$w', '%.4g' % d['value'], '%.4g us/step' % (1e3*d['ms_per_step']), 'active', d['active_vehicles_per_env'], 'crash', d['crashed_env_frac'])"
done
