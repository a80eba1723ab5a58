from .base import AutoModality, BaseModality, BaseModalityConfig, BaseModalityProcessor  # noqa: F401
from .image_modality import B200ImageConfig, B200ImageModality, B200ImageProcessor  # noqa: F401
