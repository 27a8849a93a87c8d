#!/usr/bin/env python3
"""Per-kernel SASS digest of libpifpaf_b200.so: counts of the mnemonics that show how a kernel moves
data (cuobjdump -sass; runs without a GPU).  Usage: sass_digest.py [out.md]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'openpifpaf_b200', 'libpifpaf_b200.so')
WATCH = ['UBLKCP', 'UTMALDG', 'UTMASTG', 'SYNCS', 'LDGSTS', 'REDUX', 'MATCH', 'VOTE', 'SHFL', 'ATOMS', 'ATOMG', 'RED',
         'LDG', 'STG', 'LDS', 'STS', 'BAR', 'HMMA', 'UTCHMMA', 'MUFU', 'FFMA', 'DADD', 'DMUL']
NOTES = {'UBLKCP': 'cp.async.bulk (TMA, no tensor map): global->shared loads with mbarrier / shared->global stores',
         'SYNCS': 'mbarrier operations (arrive.expect_tx / try_wait)', 'LDGSTS': 'cp.async (16-byte-less staging)',
         'REDUX': 'warp-wide integer reductions (redux.sync)', 'HMMA': 'legacy tensor-core path (none expected)',
         'UTCHMMA': 'tcgen05.mma (none expected: nothing on this path is a contraction)',
         'FFMA': 'fused multiply-add: pp_expf spells out fmaf, and the IEEE division / square-root sequences use it '
                 'internally; parity arithmetic itself is never contracted (the library is built -fmad=false)'}


def main():
    sass = subprocess.run(['cuobjdump', '-sass', LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            name = subprocess.run(['c++filt', m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
            cur = kernels.setdefault(re.sub(r'\(.*', '', name), collections.Counter())
            continue
        m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)', line)
        if m and cur is not None:
            cur['_all'] += 1
            op = m.group(1)
            for w in WATCH:
                if op == w or op.startswith(w + '.') or op.startswith(w + '_'):
                    cur[w] += 1
                    break
            else:
                cur[op.split('.')[0]] += 0
    cols = [w for w in WATCH if any(c[w] for c in kernels.values()) or w in ('HMMA', 'UTCHMMA')]
    out = ['# SASS digest of libpifpaf_b200.so (cuobjdump -sass, sm_100a)', '',
           '| kernel | instructions | ' + ' | '.join(cols) + ' |', '|---|---|' + '---|' * len(cols)]
    for name, c in kernels.items():
        out.append('| `%s` | %d | %s |' % (name.replace('pp::', ''), c['_all'], ' | '.join(str(c[w]) for w in cols)))
    out += ['', 'Legend:'] + ['- `%s`: %s' % (k, v) for k, v in NOTES.items() if k in cols]
    text = '\n'.join(out) + '\n'
    if len(sys.argv) > 1:
        open(sys.argv[1], 'w').write(text)
    print(text)


if __name__ == '__main__':
    main()
