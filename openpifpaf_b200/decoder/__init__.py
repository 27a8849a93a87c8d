"""Collections of decoders: fields to annotations (reference: openpifpaf/decoder/__init__.py)."""

from .caf_scored import CafScored
from .cif_hr import CifHr
from .cif_seeds import CifSeeds
from .engine import DecoderParams, DeviceDecoder
from .factory import cli, configure, factory_decode, factory_from_args
from .field_config import FieldConfig
from .generator.cifcaf import CifCaf
from .generator.cifdet import CifDet, CifDetHr, CifDetSeeds
from .generator.generator import Generator
from . import nms
from .occupancy import Occupancy
