"""`gluefactory.utils.export_predictions` on the B200 kernels (reference: utils/export_predictions.py:17-81)."""
from glue_factory_b200.export import RecordFile, export_predictions  # noqa: F401
