"""Model registry (reference: gluefactory/models/__init__.py:7-30).

Lookup order for `get_model(name)`: the raw name as an import path, then
``gluefactory.models.<name>``, ``…extractors.<name>``, ``…matchers.<name>``.  A module
resolves to the single BaseModel subclass it defines, otherwise to ``__main_model__``.
"""
import importlib
import importlib.util

from ..utils.tools import get_class
from .base_model import BaseModel


def _exists(path):
    try:
        return importlib.util.find_spec(path) is not None
    except (ModuleNotFoundError, ValueError):
        return False


def get_model(name):
    candidates = [
        name,
        f"{__name__}.{name}",
        f"{__name__}.extractors.{name}",
        f"{__name__}.matchers.{name}",
    ]
    for path in candidates:
        if not _exists(path):
            continue
        try:
            return get_class(path, BaseModel)
        except AssertionError:
            module = importlib.import_module(path)
            main = getattr(module, "__main_model__", None)
            if main is not None:
                return main
            print(f"module '{path}' has no attribute '__main_model__'")
    raise RuntimeError(f'Model {name} not found in any of [{" ".join(candidates)}]')
