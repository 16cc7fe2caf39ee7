#!/bin/bash
# c2 / c5 thread-shape sweep of the hot kernel (mode 25 = kSafe instantiation), 3000 steps each
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_speed_mode.py tests/test_gpu_hot_kernel.py -m gpu -x -q 2>&1 | tail -4
for sh in 0 2 4 8 16; do
  FS_RING_SHAPE=$sh python bench.py --only-fused --steps 3000 > gpurun_out/sw_c2_$sh.json 2> gpurun_out/sw_c2_$sh.err
done
for sh in 0 1 2; do
  FS_RING_SHAPE=$sh python bench.py --workload c5 --only-fused --steps 3000 > gpurun_out/sw_c5_$sh.json 2> gpurun_out/sw_c5_$sh.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/sw_c*.json')):
    try:
        d = json.load(open(f)); print(f, '%.4g' % d['value'], '%.4g us/step' % (1e3 * d['ms_per_step']))
    except Exception as ex: print(f, 'ERR', open(f.replace('.json', '.err')).read()[-300:])
PY
