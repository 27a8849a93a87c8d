#!/usr/bin/env python3
"""Top CUDA source lines of a captured kernel by warp-stall samples.
usage: ncu_lines.py REPORT.ncu-rep [N] [--no-barrier]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 30
nobar = '--no-barrier' in sys.argv
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass',
                      '--launch-count', '1'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = None
fname = ''
out = []
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    elif r and r[0] == 'Line No':
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0] not in ('', 'Line No'):
        si = hdr.index('# Samples'); bi = hdr.index('stall_barrier'); ie = hdr.index('Instructions Executed')
        s = int(r[si] or 0) - (int(r[bi] or 0) if nobar else 0)
        stalls = sorted(((int(r[i] or 0), hdr[i][6:]) for i in range(len(hdr))
                         if hdr[i].startswith('stall_') and 'Not Issued' not in hdr[i]
                         and not (nobar and hdr[i] == 'stall_barrier')), reverse=True)[:2]
        out.append((s, fname, r[0], r[1].strip()[:80], r[ie], stalls))
tot = sum(o[0] for o in out) or 1
print('total samples', tot)
for o in sorted(out, reverse=True)[:n]:
    print('%5d %4.1f%% %s:%s  %s  [inst %s] %s' % (o[0], 100.0 * o[0] / tot, o[1], o[2], o[3], o[4], o[5]))
