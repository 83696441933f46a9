"""CPU: the C-ABI library loads and exports exactly what include/dvf_b200.h declares (no compute calls)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "dvf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dvf_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from dynamo_release_b200 import _lib

    lib = _lib.load()
    names = _header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "libdvf_b200.so does not export %s" % n
    # and the ctypes table covers the header one to one
    assert sorted(_lib.SIGNATURES) == names
    assert lib.dvf_abi_version() == 2


def test_argument_errors_are_reported_without_a_gpu():
    from dynamo_release_b200 import _lib

    lib = _lib.load()
    rc = lib.dvf_eval(None, 10, 30, None, 5, 0.1, None, None, None)
    assert rc == _lib.DVF_ERR_ARG
    assert b"NULL" in lib.dvf_last_error()
    with pytest.raises(_lib.DvfError):
        _lib.check(lib.dvf_jacobian(None, 10, 300, None, 5, 0.1, None, None, None))
    assert lib.dvf_fit_row_width(500, 30) == 576
    assert lib.dvf_fit_row_width(100, 30) == 192


def test_no_cpu_fallback_in_the_product():
    """The product package must not import the oracle and must refuse to run without CUDA."""
    import torch

    import dynamo_release_b200 as pkg

    pkg_dir = os.path.dirname(pkg.__file__)
    for fn in os.listdir(pkg_dir):
        if fn.endswith(".py"):
            text = open(os.path.join(pkg_dir, fn)).read()
            assert "oracle" not in text.replace("vfc_oracle.prelude", ""), fn + " mentions the oracle"
    if not torch.cuda.is_available():
        import numpy as np

        with pytest.raises(RuntimeError):
            pkg.con_K(np.zeros((3, 2)), np.zeros((2, 2)))
