from oracle.preprocess import kornia_resize as resize  # noqa: F401
