"""CPU tests of the drop-in boundary: the C-ABI library builds, loads and exports every symbol
include/rf_b200.h declares, and the product path fails loudly without a CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rf_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from randomfield_b200 import build, _lib
    build.build()
    return _lib


def test_header_symbols_are_exported_and_bound(lib):
    names = _declared_symbols()
    assert len(names) >= 25
    handle = ctypes.CDLL(lib.LIB_PATH)
    for name in names:
        assert hasattr(handle, name), name
    # the ctypes signature table covers exactly the header
    assert sorted(lib.SIGNATURES) == names


def test_version_and_error_string(lib):
    assert lib.lib().rf_version() >= 100
    assert isinstance(lib.last_error(), str)


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="checks the no-device failure mode")
def test_no_cpu_fallback(lib):
    """Without a CUDA device the engine refuses to construct (no silent CPU path)."""
    with pytest.raises(RuntimeError):
        lib.Engine(64, 64, 64)
    import randomfield_b200 as rf
    with pytest.raises(RuntimeError):
        rf.Generator(64, 64, 64, 2.0)


def test_argument_errors_map_to_value_error(lib):
    with pytest.raises(ValueError):
        lib.Engine(63, 64, 64)           # odd dimension: transform.py:172-177
    with pytest.raises(ValueError):
        lib.dtype_code(np.int32)


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under randomfield_b200/ may reference it."""
    pkg = os.path.join(ROOT, "randomfield_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "ref_replay" not in text, f
