"""CPU: the C-ABI library loads and exports every symbol include/ufe_b200.h declares; compute calls fail loudly without a GPU."""
import ctypes as C
import os
import re

import pytest

import ufe_b200 as ub

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_all_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "ufe_b200.h")).read()
    declared = set(re.findall(r"\b(ufe_b200_[a-zA-Z0-9_]+)\s*\(", hdr))
    assert declared == set(ub.ABI_SYMBOLS), declared ^ set(ub.ABI_SYMBOLS)
    L = ub.lib()
    for s in declared:
        assert hasattr(L, s), s


def test_config_struct_matches_header_defaults():
    L = ub.lib()
    c = ub.Config()
    L.ufe_b200_config_defaults(C.byref(c))
    d = ub.default_config()
    for name, _ in ub.Config._fields_:
        a, b = getattr(c, name), getattr(d, name)
        if hasattr(a, "__len__"):
            assert list(a) == list(b), name
        else:
            assert a == b, name


def test_no_cpu_fallback():
    if os.path.exists("/dev/nvidiactl"):
        pytest.skip("a GPU is present")
    with pytest.raises(ub.UfeError, match="no CUDA device|CPU fallback"):
        ub.Context()


def test_no_alloc_limit_follows_the_header():
    hdr = open(os.path.join(ROOT, "include", "ufe_b200.h")).read()
    assert int(re.search(r"#define UFE_B200_NOALLOC_MAX_VERTICES (\d+)", hdr).group(1)) == ub.NOALLOC_MAX_VERTICES
