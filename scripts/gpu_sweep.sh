python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for sh in 1 2 4 8; do FS_RING_SHAPE=$sh python bench.py --steps 600 --warmup 100 --no-cpu-baseline $NOISE > gpurun_out/c2_shape$sh.json 2>gpurun_out/c2_shape$sh.err; python -c "
import json; d=json.load(open('gpurun_out/c2_shape$sh.json')); print('c2 shape $sh fused %.3e single %.3e e2e %.3e e2e1 %.3e us/step %.2f'%(d['value'],d['single_step']['value'],d['e2e']['value'],d['e2e_single_step']['value'],d['ms_per_step']*1e3))"; done
for sh in 1 2 4; do FS_RING_SHAPE=$sh python bench.py --workload c5 --steps 300 --warmup 100 --no-cpu-baseline $NOISE > gpurun_out/c5_shape$sh.json 2>gpurun_out/c5_shape$sh.err; python -c "
import json; d=json.load(open('gpurun_out/c5_shape$sh.json')); print('c5 shape $sh fused %.3e single %.3e e2e %.3e e2e1 %.3e us/step %.2f'%(d['value'],d['single_step']['value'],d['e2e']['value'],d['e2e_single_step']['value'],d['ms_per_step']*1e3))"; done
