from oracle.eval_utils import find_homography_dlt  # noqa: F401
