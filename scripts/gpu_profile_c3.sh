#!/bin/bash
# c3 / c4: plain bench (must exit 0) -> one `--set full` capture of the step kernel.
mkdir -p gpurun_out
for w in c3 c4 mah; do
  python bench.py --workload $w --steps 1000 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || { tail -n 5 gpurun_out/bench_$w.err; exit 1; }
done
ncu --set full --clock-control none --import-source on -k regex:ring_ml -s 3 -c 1 -o gpurun_out/prof_c3 \
    python bench.py --workload c3 --steps 200 --warmup 3 --no-cpu-baseline --only-fused > gpurun_out/ncu_full_c3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:highway_kernel -s 3 -c 1 -o gpurun_out/prof_c4 \
    python bench.py --workload c4 --steps 200 --warmup 3 --no-cpu-baseline --only-fused > gpurun_out/ncu_full_c4.log 2>&1
tail -n 2 gpurun_out/ncu_full_c3.log gpurun_out/ncu_full_c4.log
python - <<PY
import json
for w in ("c3","c4"):
    d=json.load(open("gpurun_out/bench_%s.json"%w))
    print(w, "fused %.3e single %.3e e2e %.3e blk %.3e" % (d["value"], d["single_step"]["value"], d["e2e"]["value"], d["e2e_blocking"]["value"]), "ms/step", d["ms_per_step"], "frac", d["roofline"]["frac"], "crash", d["crashed_env_frac"], "speed", d["mean_speed"])
PY
