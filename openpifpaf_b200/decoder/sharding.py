"""Image sharding of a decode job over the GPUs of one node.

Decoding is independent per image, so a job of N images is split into
contiguous shards, one per rank (= one process per GPU); every rank walks its
shard in chunks that bound the hr-map memory (64 x 641^2 images = 1.79 GB) and
there is NO collective on the data path.  ``torch.distributed`` is only used to
bring the per-image results (small Python objects) to rank 0 in image order,
which replaces the reference's ``multiprocessing.Pool.starmap`` result list
(decoder/generator/generator.py:93-98).
"""
from typing import Callable, List, Optional, Sequence, Tuple


def shard_bounds(n_items: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of rank; the first ``n_items % world_size`` ranks get one more."""
    if world_size <= 0 or not 0 <= rank < world_size:
        raise ValueError('bad rank %d of %d' % (rank, world_size))
    base, extra = divmod(max(n_items, 0), world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def chunk_ranges(lo: int, hi: int, chunk: int):
    """[lo, hi) in pieces of at most ``chunk`` images."""
    if chunk <= 0:
        raise ValueError('chunk must be positive')
    for start in range(lo, hi, chunk):
        yield start, min(start + chunk, hi)


class ShardedDecoder:
    """Runs ``decode_batch(cif, caf) -> List[List[Annotation]]`` over this rank's shard.

    :param decode_batch: the batch decode callable (``CifCaf.decode_batch``).
    :param chunk: images per call (bounds the hr workspace).
    :param group: torch.distributed process group (None = default / single process).
    """

    def __init__(self, decode_batch: Callable, *, chunk: int = 64, group=None, decode_stream: Callable = None):
        self.decode_batch = decode_batch
        # optional ``CifCaf.decode_stream``: the chunks of the shard are kept in flight (copies, kernels,
        # result copies and Annotation construction of successive chunks overlap) instead of one
        # synchronous call per chunk; same results
        self.decode_stream = decode_stream
        self.chunk = chunk
        self.group = group

    def _dist(self):
        try:
            import torch.distributed as dist
        except ImportError:       # pragma: no cover
            return None
        return dist if dist.is_available() and dist.is_initialized() else None

    def rank_world(self):
        dist = self._dist()
        if dist is None:
            return 0, 1
        return dist.get_rank(self.group), dist.get_world_size(self.group)

    def decode_shard(self, n_images: int, get_fields: Callable[[int, int], Sequence]) -> Tuple[int, List]:
        """Decode this rank's shard. ``get_fields(lo, hi)`` returns ``(cif, caf)`` for images
        [lo, hi) already on (or headed for) this rank's GPU.  Returns (first image index, results)."""
        rank, world = self.rank_world()
        lo, hi = shard_bounds(n_images, world, rank)
        results = []
        ranges = list(chunk_ranges(lo, hi, self.chunk))
        if self.decode_stream is not None:
            outs = self.decode_stream(get_fields(a, b) for a, b in ranges)
        else:
            outs = (self.decode_batch(*get_fields(a, b)) for a, b in ranges)
        for (a, b), out in zip(ranges, outs):
            if len(out) != b - a:
                raise RuntimeError('decode returned %d results for %d images' % (len(out), b - a))
            results.extend(out)
        return lo, results

    def decode_all(self, n_images: int, get_fields, *, dst: int = 0) -> Optional[List]:
        """All ranks decode their shards; rank ``dst`` returns the results in image order."""
        lo, results = self.decode_shard(n_images, get_fields)
        dist = self._dist()
        if dist is None:
            return results
        rank, world = self.rank_world()
        gathered = [None] * world if rank == dst else None
        dist.gather_object((lo, results), gathered, dst=dst, group=self.group)
        if rank != dst:
            return None
        merged = []
        for shard_lo, shard in sorted(gathered, key=lambda t: t[0]):
            assert shard_lo == len(merged), 'shards are not contiguous'
            merged.extend(shard)
        return merged
