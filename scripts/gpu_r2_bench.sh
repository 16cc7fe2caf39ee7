#!/bin/bash
# round 2: the driver's own bench call, A/B of the speed mode, the other workloads, the reference arm
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_k20.json 2> gpurun_out/r2_bench_k20.err; echo "k20 rc=$?"
python bench.py --gpus 1 --steps 3000 --warmup 300 --only-fused > gpurun_out/r2_bench_k3000.json 2> gpurun_out/r2_bench_k3000.err; echo "k3000 rc=$?"
python bench.py --speed-mode 0 --only-fused --steps 3000 > gpurun_out/r2_c2_sm0.json 2> gpurun_out/r2_c2_sm0.err; echo "sm0 rc=$?"
python bench.py --workload c5 --speed-mode 0 --only-fused --steps 3000 > gpurun_out/r2_c5_sm0.json 2> gpurun_out/r2_c5_sm0.err
for w in c3 c4 ma mah mg f8; do
  python bench.py --workload $w --only-fused --steps 3000 > gpurun_out/r2_$w.json 2> gpurun_out/r2_$w.err; echo "$w rc=$?"
done
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
tail -c 400 gpurun_out/r2_bench_k20.err
python - <<'PY'
import json
def L(f):
    try: return json.load(open('gpurun_out/%s.json' % f))
    except Exception as ex: return None
for f in ('r2_bench_k20','r2_bench_k3000','r2_c2_sm0','r2_c5_sm0','r2_c3','r2_c4','r2_ma','r2_mah','r2_mg','r2_f8','r2_bench_ref'):
    d = L(f)
    if not d: print(f, 'MISSING'); continue
    print(f, '%.4g' % d['value'], 'ms/step %.3g' % d['ms_per_step'], 'crash', d.get('crashed_env_frac'), 'frac', (d.get('roofline') or {}).get('frac'))
d = L('r2_bench_k20')
if d:
    e = d['e2e']; print('e2e %.4g blocking %.4g closed %.4g (%.2f us)' % (e['value'], e['blocking']['value'], e['closed_loop']['value'], 1e3*e['closed_loop']['ms_per_step']))
    print('single %.4g frac %.3f' % (d['single_step']['value'], d['roofline_single_step']['frac']))
    for k, v in d.get('legs', {}).items(): print('leg', k, '%.4g' % v['value'], v['roofline']['frac'], v.get('roofline_single_step', {}).get('frac'))
    print('cpu', d['cpu_baseline']['kind'], '%.4g' % d['cpu_baseline']['value'], 'port %.4g' % d['cpu_baseline'].get('port', {}).get('value', 0))
PY
