"""`get_model('gluefactory_nonfree.superglue')` -> the B200 SuperGlue matcher
(see glue_factory_b200.models.superglue; reference: gluefactory_nonfree/superglue.py:204-307)."""
from glue_factory_b200.models.superglue import SuperGlue as _SuperGlue

__main_model__ = _SuperGlue
