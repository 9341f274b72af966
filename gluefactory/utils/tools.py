"""Registry helper (reference: gluefactory/utils/tools.py:193-206)."""
import importlib
import inspect
import random

import numpy as np
import torch


def get_class(mod_path, BaseClass):
    """Return the single subclass of `BaseClass` that is *defined* in module `mod_path`.

    Classes merely imported into the module do not count (their ``__module__`` differs);
    zero or several candidates raise AssertionError, which `get_model` uses as the signal
    to fall back on ``module.__main_model__``.
    """
    module = importlib.import_module(mod_path)
    found = [
        cls
        for _, cls in inspect.getmembers(module, inspect.isclass)
        if cls.__module__ == mod_path and issubclass(cls, BaseClass)
    ]
    assert len(found) == 1, found
    return found[0]


def set_seed(seed):
    """Seed python, numpy and torch (reference: gluefactory/utils/tools.py set_seed)."""
    random.seed(seed)
    torch.manual_seed(seed)
    np.random.seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
        torch.cuda.manual_seed_all(seed)
