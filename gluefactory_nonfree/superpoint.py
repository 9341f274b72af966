"""`get_model('gluefactory_nonfree.superpoint')` -> the B200 MagicLeap-SuperPoint extractor
(see glue_factory_b200.models.superpoint_nonfree; reference: gluefactory_nonfree/superpoint.py:152-356)."""
from glue_factory_b200.models.superpoint_nonfree import SuperPoint as _SuperPoint

__main_model__ = _SuperPoint
