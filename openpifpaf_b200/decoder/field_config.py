"""Which head outputs feed the decoder and how (reference: decoder/field_config.py:5-25).

Same attribute names and defaults as the reference dataclass, because
``factory_decode`` and user code set these attributes by name.  List-valued
entries with more than one element describe the multi-scale configuration
(decoder/factory.py:153-180).
"""
import dataclasses
from typing import List, Optional


def _list_of(*values):
    return dataclasses.field(default_factory=lambda: list(values))


@dataclasses.dataclass
class FieldConfig:
    # index of each CIF / CAF head in ``fields`` and its stride in image pixels
    cif_indices: List[int] = _list_of(0)
    caf_indices: List[int] = _list_of(1)
    cif_strides: List[int] = _list_of(8)
    caf_strides: List[int] = _list_of(8)
    # multi-scale filters (0 / None = off)
    cif_min_scales: List[float] = _list_of(0.0)
    caf_min_distances: List[float] = _list_of(0.0)
    caf_max_distances: List[Optional[float]] = _list_of(None)
    # optional per-field switches
    seed_mask: Optional[List[int]] = None
    confidence_scales: Optional[List[float]] = None
    # debug hooks (objects with a ``predicted(field)`` method), see decoder/generator/cifcaf.py
    cif_visualizers: Optional[list] = None
    caf_visualizers: Optional[list] = None

    def verify(self):
        n_cif, n_caf = len(self.cif_indices), len(self.caf_indices)
        assert len(self.cif_strides) == n_cif
        assert len(self.cif_min_scales) == n_cif
        assert len(self.caf_strides) == n_caf
        assert len(self.caf_min_distances) == n_caf
        assert len(self.caf_max_distances) == n_caf

    @property
    def single_scale(self):
        return len(self.cif_indices) == 1 and len(self.caf_indices) == 1
