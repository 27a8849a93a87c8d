"""pp_keypoint_nms_noalloc: the keypoint NMS with its scratch in a static device array of the library
(no cudaMalloc by anyone, no counterpart in the reference).

Collected last on purpose (a failure here must not hide other tests).  Checks,
through the C ABI, that two calls in a row give exactly what pp_keypoint_nms gives on the same inputs
(bit-exact, the bar of the other parity tests), that nothing from the first call leaks into the second
(the array keeps its contents), that a request beyond the built-in size is refused with
PP_ERR_STATIC_LIMIT, and that the call itself allocates no device memory.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip('torch')

pytestmark = pytest.mark.gpu

F = 17


def _poses(rng, B, cap, n_per_image, hw=641):
    """B padded lists of overlapping synthetic poses: (data (B,cap,F,3), scales (B,cap,F), n (B,))."""
    data = np.zeros((B, cap, F, 3), np.float32)
    scales = np.zeros((B, cap, F), np.float32)
    n = np.zeros((B,), np.int32)
    for b in range(B):
        n[b] = n_per_image[b]
        centres = rng.uniform(0.2 * hw, 0.8 * hw, size=(max(1, n[b] // 3 + 1), 2))
        for i in range(n[b]):
            c = centres[rng.randint(len(centres))]          # several poses per centre: suppression happens
            xy = c[None, :] + rng.normal(0.0, 12.0, size=(F, 2))
            v = rng.uniform(0.0, 1.0, size=F)
            v[rng.uniform(size=F) < 0.2] = 0.0                # unset joints
            v[rng.uniform(size=F) < 0.05] = 0.0005            # below the keypoint threshold
            data[b, i, :, :2] = xy
            data[b, i, :, 2] = v
            scales[b, i] = rng.uniform(1.0, 30.0, size=F)
    return data, scales, n


def _run(entry, data, scales, n, cap, params, check_free=True):
    """One NMS call through the C ABI; returns host copies of (data, scales, n, score, source)."""
    from openpifpaf_b200 import _lib
    lib = _lib.lib()
    B = data.shape[0]
    shape = _lib.Shape(B, F, 0, 1, 1, 8)
    caps = _lib.Caps(1, 1, cap, 1)
    d_data = torch.from_numpy(data).cuda()
    d_scales = torch.from_numpy(scales).cuda()
    d_n = torch.from_numpy(n).cuda()
    d_score = torch.full((B, cap), -7.0, dtype=torch.float64, device='cuda')
    d_source = torch.full((B, cap), -7, dtype=torch.int32, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    if entry == 'workspace':
        ws_bytes = lib.pp_workspace_bytes(C.byref(_lib.Shape(B, F, 0, 1, 1, 8)), C.byref(caps))
        assert ws_bytes > 0
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device='cuda')
        rc = lib.pp_keypoint_nms(ptr(d_data), ptr(d_scales), ptr(d_n), ptr(d_score), ptr(d_source), ptr(ws),
                                 ws_bytes, C.byref(shape), C.byref(caps), C.byref(params), stream)
    else:
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        rc = lib.pp_keypoint_nms_noalloc(ptr(d_data), ptr(d_scales), ptr(d_n), ptr(d_score), ptr(d_source),
                                         C.byref(shape), C.byref(caps), C.byref(params), stream)
        torch.cuda.synchronize()
        after = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        assert after[0] == before[0], 'torch allocated device memory during the no-alloc call'
        # (the very first use of the library's module in a process loads it, static array included:
        # that is the driver's module load, not an allocation of this call -- hence the warm-up call)
        assert not check_free or after[1] >= before[1], 'free device memory shrank during the no-alloc call'
    _lib.check(rc)
    torch.cuda.synchronize()
    return tuple(t.cpu().numpy() for t in (d_data, d_scales, d_n, d_score, d_source))


def _params():
    from openpifpaf_b200.decoder.engine import DecoderParams
    return DecoderParams(nms_keypoint_threshold=0.001, nms_instance_threshold=0.05,
                         nms_suppression=0.0).to_c()


def _assert_same(got, want, cap):
    g_data, g_scales, g_n, g_score, g_source = got
    w_data, w_scales, w_n, w_score, w_source = want
    assert np.array_equal(g_n, w_n)
    for b, m in enumerate(w_n.tolist()):
        assert 0 <= m <= cap
        assert np.array_equal(g_data[b, :m], w_data[b, :m]), 'annotations differ (image %d)' % b
        assert np.array_equal(g_scales[b, :m], w_scales[b, :m])
        assert np.array_equal(g_score[b, :m], w_score[b, :m])
        assert np.array_equal(g_source[b, :m], w_source[b, :m])


def test_noalloc_nms_equals_workspace_nms_two_calls_in_a_row():
    from openpifpaf_b200 import _lib
    lib = _lib.lib()
    assert lib.pp_keypoint_nms_noalloc_max_bytes() == 4 * 1024 * 1024
    rng = np.random.RandomState(7)
    params = _params()
    warm = _poses(np.random.RandomState(1), 1, 4, [3])
    _run('noalloc', *warm, 4, params, check_free=False)            # loads the module in a fresh process
    # call 1: 4 images, up to 60 poses each; call 2: other shape, other counts (an empty image too),
    # so stale contents of the shared array would show
    cases = [(4, 64, [60, 37, 1, 64]), (3, 40, [0, 40, 11])]
    inputs = [_poses(rng, B, cap, counts) for B, cap, counts in cases]
    results = []
    for (B, cap, _), (data, scales, n) in zip(cases, inputs):       # the two no-alloc calls, back to back
        need = lib.pp_keypoint_nms_noalloc_bytes(C.byref(_lib.Shape(B, F, 0, 1, 1, 8)),
                                                 C.byref(_lib.Caps(1, 1, cap, 1)))
        assert 0 < need <= lib.pp_keypoint_nms_noalloc_max_bytes()
        results.append(_run('noalloc', data.copy(), scales.copy(), n.copy(), cap, params))
    for (B, cap, _), (data, scales, n), got in zip(cases, inputs, results):
        want = _run('workspace', data.copy(), scales.copy(), n.copy(), cap, params)
        assert want[2].sum() > 0 and (want[2] < n).any(), 'the case must suppress something'
        _assert_same(got, want, cap)
    # and once more after the others: same inputs, same answer
    again = _run('noalloc', inputs[0][0].copy(), inputs[0][1].copy(), inputs[0][2].copy(), cases[0][1], params)
    _assert_same(again, results[0], cases[0][1])


def test_noalloc_nms_rejects_requests_beyond_the_static_scratch():
    from openpifpaf_b200 import _lib
    lib = _lib.lib()
    B, cap = 64, 256                                   # 16384 annotation slots: ~19 MB of scratch
    shape, caps = _lib.Shape(B, F, 0, 1, 1, 8), _lib.Caps(1, 1, cap, 1)
    assert lib.pp_keypoint_nms_noalloc_bytes(C.byref(shape), C.byref(caps)) > \
        lib.pp_keypoint_nms_noalloc_max_bytes()
    data = torch.zeros((B, cap, F, 3), device='cuda')
    scales = torch.zeros((B, cap, F), device='cuda')
    n = torch.zeros((B,), dtype=torch.int32, device='cuda')
    params = _params()
    rc = lib.pp_keypoint_nms_noalloc(C.c_void_p(data.data_ptr()), C.c_void_p(scales.data_ptr()),
                                     C.c_void_p(n.data_ptr()), None, None, C.byref(shape), C.byref(caps),
                                     C.byref(params), None)
    assert rc == _lib.PP_ERR_STATIC_LIMIT
    assert b'static scratch' in lib.pp_error_string(rc)
    with pytest.raises(_lib.PifPafError):
        _lib.check(rc)
