"""`gluefactory.utils.image` subset on the B200 kernels (reference: gluefactory/utils/image.py:12-128)."""
from glue_factory_b200.image import (  # noqa: F401
    ImagePreprocessor,
    load_image,
    numpy_image_to_torch,
    read_image,
)
