"""API-compatible subset of Glue Factory covering the extraction + matching forward path.

Only the model registry, the BaseModel contract and the two-view pipeline are provided;
the kernels live in `glue_factory_b200`.  (reference layer map: SURVEY.md section 1)
"""
import logging

logger = logging.getLogger(__name__)
if not logger.handlers:
    _h = logging.StreamHandler()
    _h.setFormatter(
        logging.Formatter("[%(asctime)s %(name)s %(levelname)s] %(message)s", "%m/%d/%Y %H:%M:%S")
    )
    logger.addHandler(_h)
    logger.setLevel(logging.INFO)
    logger.propagate = False

__module_name__ = __name__
