#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hot_kernel.py -k closed_loop -m gpu -x -q 2>&1 | tail -12
python bench.py --gpus 1 --steps 20 --warmup 5 --no-extra-legs --no-cpu-baseline > gpurun_out/r2_cl.json 2> gpurun_out/r2_cl.err; echo "rc=$?"; tail -c 400 gpurun_out/r2_cl.err
python -c "
import json
d=json.load(open('gpurun_out/r2_cl.json')); print('value', d['value']); e=d['e2e']; print('e2e', e['value'], 'blocking', e['blocking']['value'], e['blocking']['ms_per_step'], 'closed', e['closed_loop']['value'], e['closed_loop']['ms_per_step'])
"
