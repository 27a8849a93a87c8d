"""Workload tooling (openpifpaf_b200/workloads.py): the torch statement of the head epilogue used to
generate random-init head fields where no CUDA device is present, against the pinned numpy oracle."""
import numpy as np
import torch


def test_collect_torch_matches_head_oracle():
    from openpifpaf_b200 import workloads
    from oracle.head_oracle import head_epilogue
    rng = np.random.RandomState(3)
    for n_fields, n_vectors, hw in ((17, 1, (5, 7)), (19, 2, (4, 3))):
        x = rng.normal(0, 1.5, (2, n_fields * (1 + 4 * n_vectors) * 4) + hw).astype(np.float32)
        want = head_epilogue(x, n_fields, n_vectors)
        got = workloads.collect_torch(torch.from_numpy(x), n_fields, n_vectors).numpy()
        assert got.shape == want.shape
        # regressions and log-b rows are plain copies / adds; sigmoid and exp go through torch's kernels
        np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-7)


def test_random_init_net_shapes():
    from openpifpaf_b200 import workloads
    trunk, conv_cif, conv_caf = workloads.build_random_init_net(seed=0)
    assert conv_cif.out_channels == 17 * 5 * 4 and conv_caf.out_channels == 19 * 9 * 4
    with torch.no_grad():
        feat = trunk(torch.zeros(1, 3, 65, 97))
    assert feat.shape == (1, 2048, 5, 7)          # stride 16: (65 - 1) / 16 + 1, (97 - 1) / 16 + 1
    cif = workloads.collect_torch(conv_cif(feat).detach(), 17, 1)
    caf = workloads.collect_torch(conv_caf(feat).detach(), 19, 2)
    assert cif.shape == (1, 17, 5, 9, 13) and caf.shape == (1, 19, 9, 9, 13)
    stats = workloads.field_statistics(cif)
    assert 0.0 <= stats['active_fraction'] <= 1.0
