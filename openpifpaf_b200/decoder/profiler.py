"""Wall-clock + CUDA-event profiling wrapper for ``decode.__call__``
(role of the reference's decoder/profiler.py:9-33, cProfile based)."""
import cProfile
import io
import logging
import pstats

import torch

LOG = logging.getLogger(__name__)


class Profiler:
    def __init__(self, function_to_profile, *, profile=None, out_name=None):
        self.function_to_profile = function_to_profile
        self.profile = profile or cProfile.Profile()
        self.out_name = out_name

    def __call__(self, *args, **kwargs):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.profile.enable()
        result = self.function_to_profile(*args, **kwargs)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.profile.disable()
        stream = io.StringIO()
        pstats.Stats(self.profile, stream=stream).sort_stats('tottime').print_stats(0.1)
        LOG.info(stream.getvalue())
        if self.out_name:
            self.profile.dump_stats(self.out_name)
        return result
