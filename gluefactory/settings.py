"""Filesystem locations (reference: gluefactory/settings.py:1-7)."""
from pathlib import Path

root = Path(__file__).resolve().parent.parent
DATA_PATH = root / "data"
TRAINING_PATH = root / "outputs" / "training"
EVAL_PATH = root / "outputs" / "results"
