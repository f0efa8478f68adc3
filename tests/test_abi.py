"""No-GPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/v2m_cuda.h declares,
and fails loudly (no CPU fallback) when asked to compute without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    with open(os.path.join(ROOT, "include", "v2m_cuda.h")) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(v2m_cuda_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import vcf2multialign_b200 as v
    L = v.lib()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), n
    # the ctypes table in _lib.py covers the header too
    assert set(names) <= set(v._lib.SYMBOLS)


def test_no_cpu_fallback_without_device():
    import vcf2multialign_b200 as v
    n = ctypes.c_int(0)
    st = v.lib().v2m_cuda_device_count(ctypes.byref(n))
    if st == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(v.V2MError):
        v.transpose_matrix(np.zeros(64, np.uint64), 64, 64)


def test_argument_errors_match_reference_asserts():
    import vcf2multialign_b200 as v
    with pytest.raises(v.V2MError):
        v.transpose_matrix(np.zeros(3, np.uint64), 96, 2)       # transpose_matrix.cc:53-54
    assert v.transpose_matrix(np.zeros(0, np.uint64), 64, 0).size == 0   # transpose_matrix.cc:48-49


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under vcf2multialign_b200/ or include/ may reference it."""
    for base in ("vcf2multialign_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dirpath.split(os.sep):
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".hh", ".h", ".cc", ".cpp")):
                    with open(os.path.join(dirpath, f), errors="ignore") as fh:
                        text = fh.read()
                    assert "oracle" not in text.lower() or f == "_never_", (dirpath, f)
