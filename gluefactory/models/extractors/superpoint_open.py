"""`get_model('extractors.superpoint_open')` -> the B200 SuperPoint-open extractor.

The class is defined in `glue_factory_b200.models.superpoint_open`; a re-exported class fails the
registry's `__module__` filter, so the registry falls through to `__main_model__`
(reference: gluefactory/models/__init__.py:22-25, utils/tools.py:202-205).
"""
from glue_factory_b200.models.superpoint_open import SuperPoint as _SuperPoint

__main_model__ = _SuperPoint
