"""CPU checks of the C-ABI library: it loads without a GPU and exports every symbol of include/b200glue.h."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    from glue_factory_b200 import build

    return build.build()


def test_library_exports_every_declared_symbol(lib_path):
    header = open(os.path.join(ROOT, "include", "b200glue.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", header))
    declared -= {"b200_stream_t"}
    assert len(declared) >= 15
    lib = ctypes.CDLL(lib_path)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing


def test_binding_matches_header(lib_path):
    from glue_factory_b200 import _lib

    header = open(os.path.join(ROOT, "include", "b200glue.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", header))
    assert set(_lib.declared_symbols()) == declared
    assert _lib.lib().b200_abi_version() == _lib.ABI_VERSION
    # struct mirrors have the C sizes (LP64): 34 fields / 21 fields with natural alignment
    assert ctypes.sizeof(_lib.AttnSeq) == 24
    assert ctypes.sizeof(_lib.GemmDesc) % 8 == 0 and ctypes.sizeof(_lib.AttnDesc) % 8 == 0


def test_no_cpu_fallback():
    import torch

    from glue_factory_b200 import _lib, ops

    with pytest.raises(_lib.B200Error):
        ops.sp_nms_border(torch.zeros(1, 16, 16), 2, 0)


def test_errors_are_reported_without_a_gpu(lib_path):
    from glue_factory_b200 import _lib

    lib = _lib.lib()
    rc = lib.b200_sp_select_topk(None, 1, 8, 8, 0.0, 0, None, None, None, None, None)  # k = 0 is rejected up front
    assert rc != 0 and b"select_topk" in lib.b200_last_error()
