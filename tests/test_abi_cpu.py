"""CPU checks of the C ABI: the library loads, exports every symbol that
include/npge_anchor.h declares, and refuses to compute without a device."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, has_gpu
from npge_b200 import _abi, build


def test_library_builds_and_exports_all_declared_symbols():
    path = build.build()
    lib = C.CDLL(path)
    header = open(os.path.join(ROOT, "include", "npge_anchor.h")).read()
    declared = set(re.findall(r"\b(npge_(?:anchor|fragment|block)_[a-z_]+)\s*\(", header))
    assert declared == set(_abi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_version_and_struct_layout():
    lib = _abi.load()
    assert b"sm_100a" in lib.npge_anchor_version()
    assert C.sizeof(_abi.Params) == 16
    assert C.sizeof(_abi.Result) == 96
    assert C.sizeof(_abi.Stats) == 152


@pytest.mark.skipif(has_gpu(), reason="checks the no-device error path")
def test_no_cpu_fallback():
    lib = _abi.load()
    ctx = C.c_void_p()
    rc = lib.npge_anchor_create(C.byref(ctx), 0)
    assert rc == _abi.ERR_CUDA
    assert b"no CPU fallback" in lib.npge_anchor_last_error()
    from npge_b200.engine import AnchorEngine, AnchorError
    with pytest.raises(AnchorError):
        AnchorEngine()
