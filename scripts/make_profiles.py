#!/usr/bin/env python
"""Condense gpurun_out/ ncu exports into the tracked profiles/ summaries of a round.
usage: make_profiles.py r01"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, 'gpurun_out'), os.path.join(ROOT, 'profiles')
tag = sys.argv[1] if len(sys.argv) > 1 else 'r01'


def short(name):
    for k in ('ring1_hot_kernel', 'ring1_fast_kernel', 'ring1_kernel', 'ring_reset_kernel',
              'ring_ml_kernel', 'highway_kernel'):
        if k in name:
            i = name.index(k)
            return name[i:name.index('(', i)] if '(' in name[i:] else name[i:]
    return 'torch/other: ' + name.split('(')[0][-60:]


# 1. launch list -> per-kernel totals and shares
ll = os.path.join(G, 'launches_c2.csv')
if os.path.exists(ll):
    rows = [r for r in csv.reader(open(ll)) if len(r) > 10 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        k = short(r[4])
        a = agg.setdefault(k, [0, 0.0, r[7], r[8]])
        a[0] += 1
        a[1] += float(r[-1])
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(P, '%s_launches_c2_summary.csv' % tag), 'w') as f:
        f.write('# ncu --metrics gpu__time_duration.sum --clock-control none, python bench.py '
                '--steps 300 --warmup 100 --no-cpu-baseline (c2); per-launch times are '
                'serialised and cold-cache: compare SHARES, not absolutes\n')
        f.write('kernel,launches,total_us,share,block,grid\n')
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write('"%s",%d,%.1f,%.4f,"%s","%s"\n' % (k, a[0], a[1] / 1e3, a[1] / tot, a[2], a[3]))
    # the raw list, ring kernels only, one line per launch
    with open(os.path.join(P, '%s_launches_c2.csv' % tag), 'w') as f:
        f.write('id,kernel,block,grid,gpu_time_ns\n')
        for r in rows:
            f.write('%s,"%s","%s","%s",%s\n' % (r[0], short(r[4]), r[7], r[8], r[-1]))

# 2. full captures -> summary csv + traffic json
traffic = {}
KERNEL = {'c2': 'ring1_hot', 'c5': 'ring1_hot', 'c3': 'ring_ml', 'c4': 'highway_kernel'}
for wl in ('c2', 'c5', 'c3', 'c4'):
    raw = os.path.join(G, 'prof_%s_final_raw.csv' % wl)
    if not os.path.exists(raw) and os.path.exists(os.path.join(G, 'prof_%s_raw.csv' % wl)):
        raw = os.path.join(G, 'prof_%s_raw.csv' % wl)     # exported on the GPU box (gpu_final.sh)
    rep = os.path.join(G, ('prof_%s_final.ncu-rep' if wl in ('c2', 'c5') else 'prof_%s.ncu-rep') % wl)
    if os.path.exists(rep) and (not os.path.exists(raw) or os.path.getmtime(raw) < os.path.getmtime(rep)):
        with open(raw, 'w') as f:
            subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], stdout=f, stderr=subprocess.DEVNULL)
    if not os.path.exists(raw):
        continue
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'scripts', 'ncu_summary.py'), raw],
                         capture_output=True, text=True).stdout
    open(os.path.join(P, '%s_ncu_full_%s_%s_f32.csv' % (tag, wl, KERNEL[wl] if wl in ('c3', 'c4') else 'ring1_hot')), 'w').write(
        '# ncu --set full --clock-control none --import-source on -k regex:%s, one timed '
        'fused launch (%d steps) of python bench.py --workload %s --only-fused\n'
        % (KERNEL[wl], 500 if wl == 'c2' else 100, wl) + out)
    rows = list(csv.reader(open(raw)))
    h, u, r = rows[0], rows[1], rows[2]
    def val(k):
        i = h.index(k)
        x = float(r[i])
        return x * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u[i]]
    traffic['%s_f32_fuse%d' % (wl, 500 if wl == 'c2' else 100)] = \
        val('dram__bytes_read.sum') + val('dram__bytes_write.sum')
rep = os.path.join(G, 'prof_c5_single.ncu-rep')
raw = os.path.join(G, 'prof_c5_single_raw.csv')
if os.path.exists(rep):
    with open(raw, 'w') as f:
        subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], stdout=f, stderr=subprocess.DEVNULL)
if os.path.exists(raw):
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'scripts', 'ncu_summary.py'), raw],
                         capture_output=True, text=True).stdout
    open(os.path.join(P, '%s_ncu_full_c5_single_step.csv' % tag), 'w').write(
        '# ncu --set full --clock-control none --import-source on -k regex:ring1_hot -s 12 -c 1, ONE '
        'single-step launch (k = 1, 65 536 envs) of FS_NO_GRAPH=1 FS_NO_PDL=1 python bench.py '
        '--workload c5 --steps 40 --warmup 10 --fuse 40\n' + out)
if traffic:
    json.dump(traffic, open(os.path.join(P, 'ncu_traffic.json'), 'w'), indent=1)

# 3. bench lines
for wl, src in (('c2', 'final_default.json'), ('c5', 'final_c5.json'), ('c3', 'bench_c3.json'),
                ('c4', 'bench_c4.json'), ('mah', 'bench_mah.json'),
                ('reference_arm', 'final_reference.json')):
    b = os.path.join(G, src)
    if os.path.exists(b):
        open(os.path.join(P, '%s_bench_%s.json' % (tag, wl)), 'w').write(open(b).read())
print(sorted(os.listdir(P)))
