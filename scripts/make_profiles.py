#!/usr/bin/env python
"""Condense gpurun_out/ ncu exports (scripts/gpu_r2_profile.sh) into the tracked profiles/
summaries of a round.   usage: make_profiles.py r02"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, 'gpurun_out'), os.path.join(ROOT, 'profiles')
tag = sys.argv[1] if len(sys.argv) > 1 else 'r02'
CMD = 'python bench.py --no-cpu-baseline --only-fused --target-ms 5 --steps 20 --warmup 5'


def short(name):
    for k in ('ring1_hot_kernel', 'ring1_fast_kernel', 'ring1_kernel', 'ring_reset_kernel',
              'ring_ml_kernel', 'highway_kernel', 'merge_kernel', 'fig8_kernel'):
        if k in name:
            i = name.index(k)
            return name[i:name.index('(', i)] if '(' in name[i:] else name[i:]
    return 'torch/other: ' + name.split('(')[0][-60:]


# 1. launch list -> per-kernel totals and shares
ll = os.path.join(G, 'p2_launches_c2.csv')
if os.path.exists(ll):
    rows = [r for r in csv.reader(open(ll)) if len(r) > 10 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        a = agg.setdefault(short(r[4]), [0, 0.0, r[7], r[8]])
        a[0] += 1
        a[1] += float(r[-1])
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(P, '%s_launches_c2_summary.csv' % tag), 'w') as f:
        f.write('# ncu --metrics gpu__time_duration.sum --clock-control none -c 400, python bench.py '
                '--no-cpu-baseline --no-extra-legs --only-fused --target-ms 20 --steps 20 --warmup 5 (c2); '
                'per-launch times are serialised and cold-cache: compare SHARES, not absolutes. '
                'ring1_hot_kernel<float, 3, 8, 1, 1, 1> = the timed fused launches (500 steps each), '
                '<..., 1, 0, 1> = the 750 warm-up steps inside fs_reset (outside the timed brackets)\n')
        f.write('kernel,launches,total_us,share,block,grid\n')
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write('"%s",%d,%.1f,%.4f,"%s","%s"\n' % (k, a[0], a[1] / 1e3, a[1] / tot, a[2], a[3]))
    with open(os.path.join(P, '%s_launches_c2.csv' % tag), 'w') as f:
        f.write('id,kernel,block,grid,gpu_time_ns\n')
        for r in rows:
            f.write('%s,"%s","%s","%s",%s\n' % (r[0], short(r[4]), r[7], r[8], r[-1]))

# 2. full captures -> summary csv + per-env-step traffic / issue figures for bench.py
KERNEL = {'c2': 'ring1_hot', 'c5': 'ring1_hot', 'c3': 'ring_ml', 'c4': 'highway_kernel',
          'mg': 'merge_kernel', 'f8': 'fig8_kernel'}
traffic = {}
for wl in ('c2', 'c5', 'c3', 'c4', 'mg', 'f8'):
    raw = os.path.join(G, 'p2_%s_raw.csv' % wl)
    plain = os.path.join(G, 'p2_plain_%s.json' % wl)
    if not (os.path.exists(raw) and os.path.exists(plain)):
        continue
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'scripts', 'ncu_summary.py'), raw],
                         capture_output=True, text=True).stdout
    name = '%s_ncu_full_%s_%s_f32.csv' % (tag, wl, KERNEL[wl])
    b = json.load(open(plain))
    fuse, E = b['config']['fuse'], b['config']['envs_per_gpu']
    open(os.path.join(P, name), 'w').write(
        '# ncu --set full --clock-control none --import-source on -k regex:%s -s 2 -c 1: one timed '
        'fused launch (%d steps x %d envs) of %s --workload %s (speed mode 25, idm noise 0.2)\n'
        % (KERNEL[wl], fuse, E, CMD, wl) + out)
    rows = list(csv.reader(open(raw)))
    h, u, r = rows[0], rows[1], rows[2]

    def val(k):
        x = float(r[h.index(k)].replace(',', ''))
        return x * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u[h.index(k)], 1)
    env_steps = float(fuse) * E
    # instruction counts are per SIMULATION step of an env (5 per Env.step on the merge, else 1)
    sim_steps = env_steps * int(b['config'].get('sims_per_step', 1))
    veh = b['active_vehicles_per_env']
    winst = val('smsp__inst_executed.sum')
    tinst = val('smsp__thread_inst_executed.sum') if 'smsp__thread_inst_executed.sum' in h else \
        winst * val('smsp__thread_inst_executed_per_inst_executed.ratio')
    traffic['%s_f32' % wl] = {
        'dram_bytes_per_env_step': (val('dram__bytes_read.sum') + val('dram__bytes_write.sum')) / env_steps,
        'issue_active_pct': val('smsp__issue_active.avg.pct_of_peak_sustained_active'),
        'warps_active_pct': val('sm__warps_active.avg.pct_of_peak_sustained_active'),
        'warp_inst_per_env_step': winst / sim_steps,
        'thread_inst_per_vehicle_step': tinst / sim_steps / veh,
        'fma_pipe_pct': val('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active'),
        'alu_pipe_pct': val('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'),
        'xu_pipe_pct': val('sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active'),
        'launch_us_under_ncu': val('gpu__time_duration.sum') * {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(
            u[h.index('gpu__time_duration.sum')], 1e-3),
        'capture': 'profiles/' + name}
if traffic:
    json.dump(traffic, open(os.path.join(P, 'ncu_traffic.json'), 'w'), indent=1)
print(json.dumps(traffic, indent=1))
