"""The C-ABI library loads, exports every symbol include/gcn_b200.h declares, and refuses to
compute without a GPU (no CPU fallback).  CPU only: no compute calls."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from gcn_b200 import _ffi

HEADER = os.path.join(ROOT, "include", "gcn_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"^\s*(?:const\s+)?(?:int|int64_t|void|char\s*\*|const char\s*\*)\s+\*?(\w+)\s*\(",
                                 text, flags=re.M)))


def test_header_symbols_all_exported_and_bound():
    syms = declared_symbols()
    assert len(syms) == len(_ffi.SIGNATURES) and "gcnb_spmm" in syms and "flexspmm" in syms
    lib = ctypes.CDLL(_ffi.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/gcn_b200.h but not exported"
        assert s in _ffi.SIGNATURES, f"{s} has no ctypes signature in gcn_b200/_ffi.py"
    assert set(_ffi.SIGNATURES) == set(syms)


def test_abi_version_and_error_string():
    assert _ffi.lib.gcnb_abi_version() == 1
    rc = _ffi.lib.gcnb_order_deg(None, None, 3, 0, 1, None)
    assert rc == -1 and "null" in _ffi.last_error().lower()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    assert _ffi.lib.gcnb_device_count() == 0
    rp = np.array([0, 1], dtype=np.int32)
    ci = np.array([0], dtype=np.int32)
    va = np.ones(1, dtype=np.float32)
    h = ctypes.c_void_p()
    rc = _ffi.lib.gcnb_plan_create(_ffi.np_ptr(rp), _ffi.np_ptr(ci), _ffi.np_ptr(va), 1, 1, 1, 0, 0,
                                   ctypes.byref(h))
    assert rc == _ffi.ERR_NOGPU and "no CPU fallback" in _ffi.last_error()
    from gcn_b200 import layers, ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.spmm(torch.eye(2).to_sparse(), torch.ones(2, 2))
    layer = layers.GraphConvolution(4, 2)
    with pytest.raises(RuntimeError, match="CPU"):
        layer(torch.ones(3, 4), torch.eye(3).to_sparse())


def test_product_does_not_import_oracle():
    """Nothing under gcn_b200/ may reference oracle/ (test infrastructure only)."""
    pkg = os.path.join(ROOT, "gcn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle|oracle/|liboracle", text, flags=re.M), f
