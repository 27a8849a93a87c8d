"""Device-side epilogue of the head networks (the step between the network's last
convolution and the decoder).  Only this epilogue is built here; base networks
and training heads are outside the decode path."""
from .heads import CifCafCollector, FusedHeadEpilogue, index_field_torch  # noqa: F401
