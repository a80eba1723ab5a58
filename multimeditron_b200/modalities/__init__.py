from .base import AutoModality, BaseModality, BaseModalityConfig, BaseModalityProcessor  # noqa: F401
from .image_modality import B200ImageConfig, B200ImageModality, B200ImageProcessor  # noqa: F401
from .image_modality_moe import (B200MOEImageConfig, B200MOEImageConfigPEP, B200MOEImageModality,  # noqa: F401
                                 B200MOEImageModalityPEP, B200MOEImageProcessor)
