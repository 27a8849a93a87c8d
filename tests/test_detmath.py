"""include/pifpaf_detmath.h: the deterministic float32 exp shared by the CUDA
assembly kernel and the oracle's exp_mode=1 is within 1 ulp of exp on the
decoder's range (x = -0.5 d^2 / sigma^2 <= 0)."""
import numpy as np


def ulp_diff(a, b):
    ai = np.frombuffer(np.float32(a).tobytes(), np.int32)[0]
    bi = np.frombuffer(np.float32(b).tobytes(), np.int32)[0]
    return abs(int(ai) - int(bi))


def test_detexp_accuracy(oracle):
    rng = np.random.RandomState(0)
    xs = np.concatenate([-rng.uniform(0, 104, 4000), -rng.exponential(1.0, 4000),
                         [0.0, -0.0, -1e-8, -0.5, -1.0, -2.0, -87.0, -88.0, -103.9]]).astype(np.float32)
    worst = 0
    for x in xs:
        got = oracle.detexp(x)
        want = np.float32(np.exp(np.float64(x)))
        if want < np.finfo(np.float32).tiny:      # subnormal results: absolute check
            assert abs(got - want) <= 2e-45 * 4
            continue
        worst = max(worst, ulp_diff(got, want))
    assert worst <= 1, worst


def test_detexp_edges(oracle):
    assert oracle.detexp(0.0) == 1.0
    assert oracle.detexp(-200.0) == 0.0
    assert np.isnan(oracle.detexp(float('nan')))
    assert oracle.detexp(float('-inf')) == 0.0
