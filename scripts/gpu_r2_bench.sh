#!/bin/bash
# round 2: the driver's own bench call, the long-K cross-check and the reference arm
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_k20.json 2> gpurun_out/r2_bench_k20.err
echo "k20 rc=$?"
python bench.py --gpus 1 --steps 3000 --warmup 300 --no-cpu-baseline --no-extra-legs > gpurun_out/r2_bench_k3000.json 2> gpurun_out/r2_bench_k3000.err
echo "k3000 rc=$?"
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
echo "ref rc=$?"
tail -c 600 gpurun_out/r2_bench_k20.err
