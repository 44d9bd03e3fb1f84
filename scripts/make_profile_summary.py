"""Builds the machine-made part of profiles/rNN_summary.md from one evidence run in gpurun_out/:

    rNN_bench_config2.json      python bench.py                       (the bench line, CUDA-event kernel timings)
    rNN_bench_reference.json    python bench.py --impl reference
    rNN_launches_config2.csv    ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv \
                                    python bench.py --steps 2 --warmup 3 --no-cpu-baseline
    rNN_top4.ncu-rep            ncu --set full --clock-control none --import-source on \
                                    -k "regex:k_lltable2|k_chain_mma|k_round|k_hist2" -c 4 \
                                    python bench.py --no-cpu-baseline --fits 8 --steps 1 --warmup 1

usage: python scripts/make_profile_summary.py r01   -> prints markdown; small inputs are copied to profiles/."""
import collections
import csv
import json
import os
import re
import shutil
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else 'r01'
G = 'gpurun_out'
P = 'profiles'
os.makedirs(P, exist_ok=True)
for f in ('%s_bench_config2.json', '%s_bench_reference.json', '%s_launches_config2.csv', '%s_pytest_gpu.log', '%s_smoke.log',
          '%s_fullfit_parity.json', '%s_fullfit_time.log', '%s_noalloc_alone.log'):
    src = os.path.join(G, f % tag)
    if os.path.exists(src):
        shutil.copy(src, os.path.join(P, f % tag))

print('## Launch list (ncu, cold-cache and serialised: compare shares, not absolutes)\n')
rows = [r for r in csv.reader(open(os.path.join(G, '%s_launches_config2.csv' % tag))) if len(r) > 10]
hdr = rows[0]
iK, iV = hdr.index('Kernel Name'), hdr.index('Metric Value')
agg = collections.OrderedDict()
for r in rows[1:]:
    k = re.sub(r'\(.*', '', r[iK]).replace('void ', '')
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += float(r[iV].replace(',', '')) / 1e3
tot = sum(v[1] for v in agg.values())
print('| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|')
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if v[1] / tot >= 0.0005:
        print('| %s | %d | %.0f | %.1f | %.3f |' % (k, v[0], v[1], v[1] / v[0], v[1] / tot))

b = json.load(open(os.path.join(G, '%s_bench_config2.json' % tag)))
print('\n## Bench line (CUDA events on the batch stream, warm, 48 fits, config2)\n')
print('step %.2f ms, value %.4g %s, e2e %.4g (%.1f ms per step, H2D %.1f MB, D2H %.1f MB), launches %d, clocks %s\n' % (
    b['ms_per_step'], b['value'], b['unit'], b['e2e']['value'], b['e2e']['ms_per_step'],
    b['e2e']['h2d_bytes_per_step'] / 1e6, b['e2e']['d2h_bytes_per_step'] / 1e6, b['gpu_launches'], json.dumps(b['clocks'])))
kt = sum(v['ms_per_launch'] * v['launches'] for v in b['kernels'].values())
print('| kernel class | launches | ms per launch | share of kernel time | bound | GB/s (frac of %.0f) | TFLOP/s FP64 (frac of %.1f) |' % (
    b['roofline'].get('hbm_peak_gbs') or 0., b['roofline']['peak'] if b['roofline']['bound'] == 'fp64' else 0.))
print('|---|---|---|---|---|---|---|')
for k, v in sorted(b['kernels'].items(), key=lambda kv: -kv[1]['ms_per_launch'] * kv[1]['launches']):
    print('| %s | %d | %.3f | %.3f | %s | %s | %s |' % (
        k, v['launches'], v['ms_per_launch'], v['ms_per_launch'] * v['launches'] / kt, v.get('bound', ''),
        '%.0f (%.2f)' % (v['gbs'], v['hbm_frac']) if v.get('gbs') else '',
        '%.2f (%.3f)' % (v['tflops'], v['fp64_frac']) if v.get('tflops') else ''))
if 'cpu_baseline' in b:
    print('\ncpu_baseline: %s' % json.dumps(b['cpu_baseline']))
ref = os.path.join(G, '%s_bench_reference.json' % tag)
if os.path.exists(ref):
    r = json.load(open(ref))
    print('\nreference arm: %.4g %s on %s cores (%s); e2e ratio %.0fx' % (
        r['value'], r['unit'], r['cpu_baseline']['cores'], r['cpu_baseline']['sample'], b['e2e']['value'] / r['value']))

rep = os.path.join(G, '%s_top4.ncu-rep' % tag)
if not os.path.exists(rep):
    rep = os.path.join(G, '%s_top.ncu-rep' % tag)
if os.path.exists(rep):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    want = [('gpu__time_duration.sum', 'duration'), ('launch__grid_size', 'grid'), ('launch__block_size', 'block'),
            ('launch__registers_per_thread', 'registers'), ('dram__bytes_read.sum', 'dram read'),
            ('dram__bytes_write.sum', 'dram write'),
            ('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'FP64 pipe active %'),
            ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue slots busy %'),
            ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps active %'),
            ('smsp__inst_executed.sum', 'warp instructions'),
            ('smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed', 'DFMA thread inst / clk'),
            ('smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed', 'DADD thread inst / clk'),
            ('smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed', 'DMUL thread inst / clk'),
            ('sm__cycles_elapsed.max', 'cycles'),
            ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'shared bank conflicts')]
    print('\n## ncu --set full, one launch each (8 fits; values are per launch)\n')
    for r in rows[2:]:
        name = re.sub(r'\(.*', '', r[hdr.index('Kernel Name')]).replace('void ', '')
        print('### %s\n' % name)
        for n, lab in want:
            if n in hdr:
                print('* %s: %s %s' % (lab, r[hdr.index(n)], units[hdr.index(n)]))
        st = []
        for i, h in enumerate(hdr):
            if 'issue_stalled' in h and 'per_issue_active' in h:
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                if v > 0.3:
                    st.append((v, h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
        print('* stall reasons (warps per issue): ' + ', '.join('%s %.2f' % (b2, a) for a, b2 in sorted(st, reverse=True)))
        print()
