"""The C-ABI library builds, loads, and exports exactly what include/funasr_b200.h declares.
No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

from funasr_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "funasr_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fb200_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib_path():
    return build.build()


def test_header_and_binding_agree():
    assert _declared() == _lib.EXPORTS


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    for name in _declared():
        assert hasattr(lib, name), name


def test_version_and_error_string(lib_path):
    lib = _lib.load()
    assert lib.fb200_version() == 100
    assert isinstance(lib.fb200_last_error(), bytes)


def test_bad_arguments_fail_loudly_without_a_gpu(lib_path):
    # argument validation happens before any CUDA call
    with pytest.raises(RuntimeError, match="parts must be 1, 2 or 3"):
        _lib.call("fb200_gemm", None, 1, 4, None, 4, 1, 1, 1, 1, None, 0, None, 0, None, 0, 0, None, 0, 0, None)
    with pytest.raises(RuntimeError, match="head_dim"):
        _lib.call("fb200_attention", None, 8, 0, 8, None, 8, 0, None, 8, 0, 8, 1, None, None, None, 1, 1, 1, 1, 4, 64, 1.0, None,
                  0, None, 1, 0, None)


def test_ops_refuse_cpu_tensors(lib_path):
    import torch

    from funasr_b200 import ops

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.split_rows(torch.zeros(4, 64), 1)


def test_no_product_import_of_oracle():
    pkg = os.path.join(ROOT, "funasr_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), encoding="utf-8").read()
                assert "import oracle" not in src and "from oracle" not in src, f
