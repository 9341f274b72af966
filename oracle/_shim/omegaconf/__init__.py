"""Stand-in so that the UNMODIFIED reference imports in a container without `omegaconf`.

Test infrastructure only: put `oracle/_shim` on PYTHONPATH next to /root/reference when
generating golden vectors (tests/golden/make_golden.py). Never imported by the product.
"""
from glue_factory_b200._conf import (  # noqa: F401
    ConfigAttributeError,
    ConfigKeyError,
    DictConfig,
    MissingMandatoryValue,
    OmegaConf,
    ReadonlyConfigError,
    open_dict,
    read_write,
)

__b200glue_shim__ = True
ListConfig = list
