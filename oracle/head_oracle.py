"""CPU restatement (numpy) of the head epilogue.  TEST INFRASTRUCTURE ONLY.

Follows CompositeFieldFused.forward in eval mode (reference
network/heads.py:406-448) and CifCafCollector.forward (network/heads.py:65-88),
written as one gather instead of views.  Pinned against the running reference by
tests/test_oracle_vs_reference.py::test_head_epilogue and against
tests/golden/head_*.npz (made by oracle/gen_golden_heads.py).
"""
import numpy as np


def head_epilogue(x, n_fields, n_vectors, quad=1):
    x = np.asarray(x, np.float32)
    B, C, h, w = x.shape
    nf, nv = n_fields, n_vectors
    nch = 1 + 4 * nv
    if quad:
        # PixelShuffle(2): out[c, 2i+di, 2j+dj] = in[4c + 2di + dj, i, j]; then drop the last row / column
        x = x.reshape(B, C // 4, 2, 2, h, w).transpose(0, 1, 4, 2, 5, 3).reshape(B, C // 4, 2 * h, 2 * w)
        x = x[:, :, :-1, :-1]
    H, W = x.shape[2:]
    conf = x[:, 0:nf]
    regs = x[:, nf:nf + 2 * nv * nf].reshape(B, nf, nv, 2, H, W)
    logb = x[:, nf + 2 * nv * nf:nf + 3 * nv * nf].reshape(B, nf, nv, H, W)
    scale = x[:, nf + 3 * nv * nf:nf + 4 * nv * nf].reshape(B, nf, nv, H, W)
    out = np.empty((B, nf, nch, H, W), np.float32)
    with np.errstate(over='ignore'):
        out[:, :, 0] = np.float32(1) / (np.float32(1) + np.exp(-conf, dtype=np.float32))
        cols = np.arange(W, dtype=np.float32)[None, :]
        rows = np.arange(H, dtype=np.float32)[:, None]
        for v in range(nv):
            out[:, :, 1 + 4 * v] = regs[:, :, v, 0] + cols
            out[:, :, 2 + 4 * v] = regs[:, :, v, 1] + rows
            out[:, :, 3 + 4 * v] = logb[:, :, v]
            out[:, :, 4 + 4 * v] = np.exp(scale[:, :, v], dtype=np.float32)
    return out
