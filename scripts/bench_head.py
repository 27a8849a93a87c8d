#!/usr/bin/env python3
"""Time pp_head_epilogue on the bench shape (batch 64, 641 px input: h = w = 41 -> 81 x 81)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from openpifpaf_b200.network import FusedHeadEpilogue

B, h = int(os.environ.get('B', 64)), int(os.environ.get('H', 41))
out = {}
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for name, nf, nv in (('cif', 17, 1), ('caf', 19, 2)):
    ep = FusedHeadEpilogue(nf, nv)
    x = torch.randn(B, nf * (1 + 4 * nv) * 4, h, h, device='cuda')
    o = ep(x)
    ts = []
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); ep(x, out=o); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = sorted(ts)[len(ts) // 2]
    byts = o.numel() * 4 * 2
    out[name] = {'ms': ms, 'GBs': byts / ms / 1e6, 'bytes': byts}
print(json.dumps(out))
