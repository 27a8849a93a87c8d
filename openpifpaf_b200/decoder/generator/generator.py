"""Batch driver shared by decoders (reference: decoder/generator/generator.py:14-107).

Keeps the reference's public surface -- ``Generator.batch``, the static
``fields_batch``, ``last_nn_time`` / ``last_decoder_time`` and the
``worker_pool`` argument -- but the heads never leave the GPU: where the
reference copies every head to numpy and maps single images over a process
pool, ``batch`` hands the CUDA tensors of the whole batch to one device decode.
``worker_pool`` is accepted and ignored (the GPU batch replaces the pool).
"""
import logging
import time

import torch

LOG = logging.getLogger(__name__)


class DummyPool:
    """In-process stand-in kept for API compatibility (generator.py:14-17)."""

    @staticmethod
    def starmap(f, iterable):
        return [f(*i) for i in iterable]


def _nested(f, items):
    if items is None:
        return None
    if isinstance(items, (list, tuple)):
        return [_nested(f, i) for i in items]
    return f(items)


class Generator:
    def __init__(self, worker_pool=None):
        if worker_pool not in (None, 0) and not isinstance(worker_pool, DummyPool):
            LOG.info('worker_pool=%r ignored: decoding is batched on the GPU', worker_pool)
        self.worker_pool = DummyPool()
        self.last_decoder_time = 0.0
        self.last_nn_time = 0.0

    def __getstate__(self):
        return {k: v for k, v in self.__dict__.items() if k not in ('worker_pool',)}

    @staticmethod
    def heads_batch(model, image_batch, *, device=None):
        """Network forward; returns the head tensors untouched (still on the GPU)."""
        with torch.no_grad():
            if device is not None:
                image_batch = image_batch.to(device, non_blocking=True)
            return model(image_batch)

    @staticmethod
    def fields_batch(model, image_batch, *, device=None):
        """Reference layout ``[image][head]...`` of numpy arrays (generator.py:43-77).

        Provided for callers that inspect fields; ``batch`` does not use it.
        """
        start = time.time()
        heads = Generator.heads_batch(model, image_batch, device=device)
        heads = _nested(lambda x: x.cpu().numpy(), heads)
        iters = _nested(iter, heads)
        per_image = []
        while True:
            try:
                per_image.append(_nested(next, iters))
            except StopIteration:
                break
        LOG.debug('nn processing time: %.3fs', time.time() - start)
        return per_image

    def __call__(self, fields, *, initial_annotations=None):
        """Single image: fields -> annotations."""
        raise NotImplementedError()

    def decode_heads(self, heads):
        """Whole batch: head tensors (leading batch axis) -> annotations per image."""
        raise NotImplementedError()

    def batch(self, model, image_batch, *, device=None):
        """From image batch straight to annotations batch."""
        start_nn = time.perf_counter()
        heads = self.heads_batch(model, image_batch, device=device)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.last_nn_time = time.perf_counter() - start_nn

        start_decoder = time.perf_counter()
        result = self.decode_heads(heads)
        self.last_decoder_time = time.perf_counter() - start_decoder
        LOG.debug('time: nn = %.3fs, dec = %.3fs', self.last_nn_time, self.last_decoder_time)
        return result
