"""``--profile-decoder``: where one ``decode(fields)`` call spends its time.

The reference wraps ``decode.__call__`` in cProfile (decoder/profiler.py:9-33) because its decoder
is Python; here the decoder is a chain of CUDA kernels, so the wrapper switches on the library's
per-stage CUDA-event timer (``pp_profile_*``, ``engine.StageProfile``) for the call and reports
milliseconds per stage next to the wall-clock time of the call.  Same class name, constructor and
call behaviour (the wrapped function's result is returned; ``out_name`` receives the numbers).
"""
from collections import defaultdict
import json
import logging
import time

LOG = logging.getLogger(__name__)


class Profiler:
    def __init__(self, function_to_profile, *, profile=None, out_name=None):
        self.function_to_profile = function_to_profile
        self.profile = profile            # an engine.StageProfile, created on first use
        self.out_name = out_name
        self.calls = 0
        self.wall_ms = 0.0
        self.stage_ms = defaultdict(float)

    def __call__(self, *args, **kwargs):
        from .engine import StageProfile
        from .generator.cifcaf import CifCaf
        if self.profile is None:
            self.profile = StageProfile(64)
        self.profile.reset()
        CifCaf.stage_profile = self.profile
        start = time.perf_counter()
        try:
            result = self.function_to_profile(*args, **kwargs)
        finally:
            CifCaf.stage_profile = None
        self.wall_ms += 1e3 * (time.perf_counter() - start)
        self.calls += 1
        for row in self.profile.read():           # one row per decode launched inside the call
            for name, ms in zip(self.profile.names, row):
                self.stage_ms[name] += float(ms)
        lines = ['decode calls %d, wall %.3f ms per call; device time per stage and call:' % (
            self.calls, self.wall_ms / self.calls)]
        lines += ['  %-14s %8.4f ms' % (n, v / self.calls) for n, v in self.stage_ms.items()]
        print('\n'.join(lines))
        if self.out_name:
            LOG.info('writing profile file %s', self.out_name)
            with open(self.out_name, 'w') as f:
                json.dump({'calls': self.calls, 'wall_ms_per_call': self.wall_ms / self.calls,
                           'stage_ms_per_call': {n: v / self.calls for n, v in self.stage_ms.items()}}, f)
        return result
