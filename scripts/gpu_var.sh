#!/bin/bash
# differential timing of hot-kernel variants (library built with FS_HOT_EXPERIMENTS=1)
mkdir -p gpurun_out
show() { python -c "
import json,sys; d=json.load(open(sys.argv[1])); print(sys.argv[2],'fused %.3e us/step %.3f'%(d['value'],d['ms_per_step']*1e3))" $1 "$2"; }
for sh in ${SHAPES:-8 16 32}; do for var in ${VARS:-0 1 2 4 6}; do
FS_HOT_VAR=$var FS_RING_SHAPE=$sh python bench.py --steps 1000 --warmup 100 --no-cpu-baseline --only-fused > gpurun_out/var_${sh}_$var.json 2>gpurun_out/var.err && show gpurun_out/var_${sh}_$var.json "c2 shape $sh var $var"
done; done
