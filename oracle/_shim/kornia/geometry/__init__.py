from . import homography, transform  # noqa: F401
