#!/usr/bin/env python3
"""Summarise gpurun_out/<tag>_launches.csv (+ optional <tag>_<kernel>.ncu-rep captures)
into profiles/<tag>_summary.md.  Usage: summarize_ncu.py TAG [capture.ncu-rep ...]"""
import collections
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RAW_METRICS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
    'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers',
    'launch__occupancy_limit_shared_mem', 'smsp__inst_executed.sum',
]


def launches(tag, out):
    path = os.path.join(ROOT, 'gpurun_out', tag + '_launches.csv')
    if not os.path.exists(path):
        return
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    h = rows[0]
    ki, vi, gi = h.index('Kernel Name'), h.index('Metric Value'), h.index('Grid Size')
    agg = collections.OrderedDict()
    for r in rows[1:]:     # the same kernel at another grid size is another workload (e2e chunks)
        agg.setdefault(r[ki].split('(')[0] + ' grid ' + r[gi].replace(' ', ''), []).append(
            float(r[vi].replace(',', '')))
    tot = sum(sum(v) for v in agg.values())
    out.write('## Launch list (ncu --metrics gpu__time_duration.sum --clock-control none; '
              'cold-cache, serialised: compare shares)\n\n')
    out.write('| kernel | launches | mean us | share |\n|---|---|---|---|\n')
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.write('| `%s` | %d | %.1f | %.3f |\n' % (k, len(v), sum(v) / len(v) / 1e3,
                                                    sum(v) / tot))
    out.write('\n')


def capture(path, out):
    raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], stdout=subprocess.PIPE,
                         stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        return
    h, units = rows[0], rows[1]
    out.write('## `ncu --set full` capture: %s\n\n' % os.path.basename(path))
    name_i = h.index('Kernel Name')
    out.write('kernel: `%s`, %d launch(es) captured\n\n| metric | unit | per launch |\n|---|---|---|\n'
              % (rows[2][name_i].split('(')[0], len(rows) - 2))
    for m in RAW_METRICS:
        if m in h:
            i = h.index(m)
            out.write('| %s | %s | %s |\n' % (m, units[i], ', '.join(r[i] for r in rows[2:])))
    src = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--launch-count', '1'],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    if len(rows) > 3:
        h = rows[1]
        si = h.index('# Samples')
        data = [r for r in rows[2:] if len(r) > si]
        stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
        agg = {c: sum(int(r[h.index(c)] or 0) for r in data) for c in stalls}
        tot = sum(int(r[si] or 0) for r in data) or 1
        out.write('\nwarp-stall samples (first launch, %d samples): ' % tot)
        out.write(', '.join('%s %.1f%%' % (k[6:], 100.0 * v / tot)
                            for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:7]))
        out.write('\n\n')


def main():
    tag = sys.argv[1]
    os.makedirs(os.path.join(ROOT, 'profiles'), exist_ok=True)
    dest = os.path.join(ROOT, 'profiles', tag + '_summary.md')
    with open(dest, 'w') as out:
        out.write('# ncu summary %s\n\n' % tag)
        launches(tag, out)
        for p in sys.argv[2:]:
            capture(p, out)
    print(open(dest).read())


if __name__ == '__main__':
    main()
