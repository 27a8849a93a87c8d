"""Post-decode side of the reference's ``transforms`` package: mapping decoded
annotations back to the original image (``Preprocess.annotations_inverse``).
Image-side preprocessing (crop, pad, rescale of PIL images) is outside the path."""
from .preprocess import Preprocess, pack_metas, hflip_target  # noqa: F401
