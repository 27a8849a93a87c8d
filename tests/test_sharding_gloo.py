"""N>1 host logic on CPU: two gloo ranks shard a job, decode their halves and
rank 0 gets every image's result in order.  The per-batch decode is stood in by
the CPU oracle (there is no GPU here); on the GPU box the same ShardedDecoder
wraps CifCaf.decode_batch (tests/test_gpu_parity.py::test_sharded_single_rank)."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from openpifpaf_b200.decoder.sharding import ShardedDecoder, chunk_ranges, shard_bounds


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 64, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard_bounds(4096, 8, 3) == (1536, 2048)
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)
    assert list(chunk_ranges(10, 75, 64)) == [(10, 74), (74, 75)]
    assert list(chunk_ranges(5, 5, 64)) == []


N_IMAGES = 5


def _fields(i):
    from openpifpaf_b200.encoder import synthetic_fields
    return synthetic_fields((97, 129), 1 + i % 2, pose_seed=40 + i, jitter_seed=50 + i)[:2]


def _oracle_batch(cif, caf):
    from oracle import oracle as orc
    from openpifpaf_b200.datasets.constants import COCO_PERSON_SKELETON
    return [orc.decode(c, a, COCO_PERSON_SKELETON, orc.Config(exp_mode=1)).data
            for c, a in zip(cif, caf)]


def _worker(rank, world, port, out_path):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        seen = []

        def get_fields(lo, hi):
            seen.append((lo, hi))
            items = [_fields(i) for i in range(lo, hi)]
            return np.stack([c for c, _ in items]), np.stack([a for _, a in items])

        sd = ShardedDecoder(_oracle_batch, chunk=2)
        merged = sd.decode_all(N_IMAGES, get_fields)
        lo, hi = shard_bounds(N_IMAGES, world, rank)
        assert seen == list(chunk_ranges(lo, hi, 2))
        if rank == 0:
            np.savez(out_path, **{'img%d' % i: m for i, m in enumerate(merged)})
        else:
            assert merged is None
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_decode(tmp_path):
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    out = str(tmp_path / 'merged.npz')
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    z = np.load(out)
    assert len(z.files) == N_IMAGES
    for i in range(N_IMAGES):
        cif, caf = _fields(i)
        want = _oracle_batch(cif[None], caf[None])[0]
        assert np.array_equal(z['img%d' % i], want), i


def test_single_process_path():
    sd = ShardedDecoder(lambda cif, caf: [int(c) for c in cif], chunk=3)
    assert sd.decode_all(7, lambda lo, hi: (list(range(lo, hi)), None)) == list(range(7))
