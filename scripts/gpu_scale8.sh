#!/bin/bash
# 8-GPU lines of one box: weak c2 (the driver's contract), weak c5, strong c5 (65 536 envs split)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 8 --steps 1000 --warmup 100 > gpurun_out/scale_c2_n8.json 2> gpurun_out/scale_c2_n8.err
$TR --master-port 29522 bench.py --gpus 8 --workload c5 --steps 300 --warmup 10 > gpurun_out/scale_c5_n8.json 2> gpurun_out/scale_c5_n8.err
$TR --master-port 29523 bench.py --gpus 8 --workload c5 --scaling strong --steps 300 --warmup 10 > gpurun_out/scale_c5_n8_strong.json 2> gpurun_out/scale_c5_n8_strong.err
tail -c 300 gpurun_out/scale_c5_n8_strong.err
