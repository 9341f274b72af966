"""`gluefactory.eval.utils` subset on the B200 kernels (reference: gluefactory/eval/utils.py:40-96,176-196)."""
from glue_factory_b200.eval_utils import (  # noqa: F401
    eval_homography_dlt,
    eval_matches_epipolar,
    eval_matches_homography,
)
