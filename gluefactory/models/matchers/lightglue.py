"""`get_model('matchers.lightglue')` -> the B200 LightGlue matcher (see glue_factory_b200.models.lightglue)."""
from glue_factory_b200.models.lightglue import LightGlue as _LightGlue

__main_model__ = _LightGlue
