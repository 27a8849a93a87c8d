#!/usr/bin/env python3
"""Summarise multi-kernel `ncu --set full` captures (one decode per workload) into one markdown file:
per kernel duration, DRAM bytes, achieved GB/s on the bytes that actually moved, DRAM / SM throughput,
warps active, registers, occupancy limiters and the top warp-stall reasons.
Usage: summarize_ncu_multi.py OUT.md TAG:capture.ncu-rep [TAG:capture.ncu-rep ...]"""
import csv
import io
import subprocess
import sys

RAW = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
       'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
       'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
       'launch__block_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
       'launch__occupancy_limit_warps', 'smsp__inst_executed.sum']
UNIT = {'nsecond': 1e-3, 'usecond': 1.0, 'msecond': 1e3, 'second': 1e6, 'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6,
        'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}


def num(v):
    try:
        return float(v.replace(',', ''))
    except ValueError:
        return float('nan')


def stalls(path, launch_index):
    src = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--launch-skip', str(launch_index),
                          '--launch-count', '1'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = next((i for i, r in enumerate(rows) if '# Samples' in r), None)
    if hdr is None:
        return ''
    h = rows[hdr]
    si = h.index('# Samples')
    data = [r for r in rows[hdr + 1:] if len(r) > si]
    cols = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
    agg = {c: sum(int(r[h.index(c)] or 0) for r in data if r[h.index(c)].strip().isdigit()) for c in cols}
    tot = sum(int(r[si] or 0) for r in data if r[si].strip().isdigit()) or 1
    top = sorted(agg.items(), key=lambda kv: -kv[1])[:4]
    return ', '.join('%s %.0f%%' % (k[6:], 100.0 * v / tot) for k, v in top if v)


def main():
    out = ['# ncu `--set full` captures, one decode per workload (cold-cache, serialised: compare shares)', '']
    for spec in sys.argv[2:]:
        tag, path = spec.split(':', 1)
        raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            out += ['## %s: no data in %s' % (tag, path), '']
            continue
        h, units = rows[0], rows[1]
        ix = {m: h.index(m) for m in RAW if m in h}
        ki = h.index('Kernel Name')
        out += ['## %s' % tag, '',
                '| kernel | grid x block | us | DRAM read MB | DRAM write MB | GB/s on DRAM bytes | DRAM % | SM % | '
                'warps active % | regs | limit (regs / smem / warps) | top stalls |', '|' + '---|' * 12]
        total = 0.0
        for li, r in enumerate(rows[2:]):
            def val(m):
                return num(r[ix[m]]) * UNIT.get(units[ix[m]], 1.0) if m in ix else float('nan')
            us = val('gpu__time_duration.sum')
            rd, wr = val('dram__bytes_read.sum'), val('dram__bytes_write.sum')
            total += us
            out.append('| `%s` | %s x %s | %.1f | %.2f | %.2f | %.0f | %.1f | %.1f | %.1f | %d | %s / %s / %s | %s |' % (
                r[ki].split('(')[0].replace('pp::', '').replace('void ', ''), r[ix['launch__grid_size']],
                r[ix['launch__block_size']], us, rd / 1e6, wr / 1e6, (rd + wr) / us / 1e3,
                num(r[ix['gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']]),
                num(r[ix['sm__throughput.avg.pct_of_peak_sustained_elapsed']]),
                num(r[ix['sm__warps_active.avg.pct_of_peak_sustained_active']]),
                int(num(r[ix['launch__registers_per_thread']])),
                r[ix['launch__occupancy_limit_registers']].split('.')[0],
                r[ix['launch__occupancy_limit_shared_mem']].split('.')[0],
                r[ix['launch__occupancy_limit_warps']].split('.')[0], stalls(path, li)))
        out += ['', 'sum of the kernel durations: %.1f us' % total, '']
    open(sys.argv[1], 'w').write('\n'.join(out) + '\n')
    print('\n'.join(out))


if __name__ == '__main__':
    main()
