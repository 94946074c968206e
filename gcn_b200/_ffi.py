"""ctypes binding of libgcn_b200.so (include/gcn_b200.h).

The reference binds its five libraries with ``ctypes.cdll.LoadLibrary('./x.so')`` relative to the
CWD and declares no argtypes (pygcn/gcn6.py:21-25).  Here: one library, absolute path inside the
package, argtypes/restype declared, and every call checked — a non-zero status raises
``GcnB200Error`` carrying ``gcnb_last_error()``.  There is no fallback: if the library is not
built, importing this module fails.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgcn_b200.so")

F32, BF16 = 0, 1
EPI_NONE, EPI_BIAS, EPI_RELU, EPI_ACCUM, EPI_DROPOUT, EPI_MASK = 0, 1, 2, 4, 8, 16
PLAN_DEFAULT, PLAN_ASSUME_SYMMETRIC, PLAN_NO_TRANSPOSE, PLAN_PANEL, PLAN_NO_PANEL = 0, 1, 2, 4, 8
ERR_NOGPU = -3


class GcnB200Error(RuntimeError):
    def __init__(self, fn: str, code: int, text: str):
        super().__init__(f"{fn} failed with status {code}: {text}")
        self.code = code


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is not built. gcn_b200 has no Python/CPU fallback: build the CUDA library "
        "with `make -C gcn_b200/csrc` (or `python -c 'import __graft_entry__ as g; g.build()'`).")

lib = ctypes.CDLL(LIB_PATH)

_i32p = POINTER(c_int32)
_i64p = POINTER(c_int64)
_f32p = POINTER(c_float)

# name -> (restype, argtypes); every symbol include/gcn_b200.h declares
SIGNATURES = {
    "gcnb_abi_version": (c_int, []),
    "gcnb_last_error": (c_char_p, []),
    "gcnb_device_count": (c_int, []),
    "gcnb_plan_create": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int,
                                 c_int, POINTER(c_void_p)]),
    "gcnb_plan_destroy": (c_int, [c_void_p]),
    "gcnb_plan_info": (c_int, [c_void_p, _i64p]),
    "gcnb_spmm": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_void_p, c_int64, c_int, c_void_p,
                          c_int, c_void_p]),
    "gcnb_spmm_t": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_void_p, c_int64, c_int, c_void_p]),
    "gcnb_spmm_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "gcnb_permute_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "gcnb_xw": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_int, c_void_p, c_int64,
                        c_int64, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "gcnb_xw_ex": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_int, c_void_p, c_int64,
                           c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int64, ctypes.c_float,
                           ctypes.c_uint32, c_int, c_void_p]),
    "gcnb_xw_tn_colsum": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int,
                                  c_int, c_void_p]),
    "gcnb_dropout_keeps": (c_int, [c_int64, c_int, c_int, ctypes.c_float, ctypes.c_uint32]),
    "gcnb_colsum_noalloc": (c_int, [c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p]),
    "gcnb_colsum_noalloc_max_floats": (c_int64, []),
    "gcnb_logsoftmax_nll_fwd": (c_int, [c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p, c_int64, c_int64,
                                        c_void_p, c_int64, c_void_p, c_void_p]),
    "gcnb_logsoftmax_nll_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int,
                                        c_int, c_void_p, c_int64, c_void_p]),
    "gcnb_peer_alloc": (c_int, [c_int, c_int64, POINTER(c_void_p), c_void_p]),
    "gcnb_peer_free": (c_int, [c_int, c_void_p]),
    "gcnb_peer_open": (c_int, [c_int, c_void_p, POINTER(c_void_p)]),
    "gcnb_peer_close": (c_int, [c_int, c_void_p]),
    "gcnb_peer_barrier": (c_int, [c_void_p, c_int, c_int, ctypes.c_uint32, c_void_p]),
    "gcnb_pull_rows": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_int64,
                               c_int64, c_int64, c_int64, c_void_p]),
    "gcnb_order_deg": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "gcnb_order_rcm": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p]),
    "gcnb_order_gorder": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "gcnb_perm_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64]),
    "gcnb_renumber_gorder": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int]),
    "gcnb_renumber_dfs": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64]),
    "gcnb_renumber_rabbit": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64]),
    "gcnb_renumber_rcm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64]),
    "gcnb_renumber_deg": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int]),
    "gcnb_tile_count": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, _i64p]),
    "gcnb_csr2tile": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, _i64p]),
    # compatibility shim: the reference's names, all void
    "flexspmm": (None, [c_void_p] * 5 + [c_int] * 4 + [c_void_p] * 2),
    "cuspmm": (None, [c_void_p] * 5 + [c_int] * 4),
    "csr2tile": (None, [c_void_p] * 3 + [c_int] * 3 + [c_void_p] * 6 + [c_int, c_void_p]),
    "csr2seg_Cmajor": (None, [c_int] + [c_void_p] * 3 + [c_int] * 3 + [c_void_p] * 4 + [c_int, c_void_p]),
    "dfs": (None, [c_void_p] * 4 + [c_int] * 3),
    "gorder": (None, [c_void_p] * 4 + [c_int] * 3),
    "rabbit": (None, [c_void_p] * 4 + [c_int] * 3),
    "perm_apply": (None, [c_void_p] * 4 + [c_int] * 3),
    "permutate": (None, [c_void_p] * 3 + [c_int] * 3),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header and library out of sync
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return (lib.gcnb_last_error() or b"").decode("utf-8", "replace")


def check(fn: str, code: int) -> None:
    if code != 0:
        raise GcnB200Error(fn, code, last_error())


def call(fn: str, *args) -> None:
    """Call a status-returning entry point and raise on failure."""
    check(fn, getattr(lib, fn)(*args))


def np_ptr(a):
    """void* of a C-contiguous numpy array (caller keeps the array alive)."""
    return a.ctypes.data_as(c_void_p)
