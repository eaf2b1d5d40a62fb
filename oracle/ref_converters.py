"""TEST INFRASTRUCTURE ONLY — ctypes view of oracle/_ref/libhet_ref_converters.so,
i.e. the reference's own C++ converters (MyHyb.h) compiled by oracle/build_ref.py.

Mirrors what the reference's torch ops do around them
(hrt/include/DGLHackKernel/OpExport/DataConverters.inc.h:10-344): apply the
permutation TransposeCSR returns to rel_type / eids, infer num_rels as max+1.
"""
from __future__ import annotations

import ctypes
from pathlib import Path

import numpy as np

LIB_PATH = Path(__file__).resolve().parent / "_ref" / "libhet_ref_converters.so"
_P = ctypes.POINTER(ctypes.c_int64)
_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(str(LIB_PATH))
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.int64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_P)


def _c(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def transpose_csr(row_ptrs, col, rel, eid):
    """DataConverters.inc.h:283-344 -> MyHyb.h:371-397 -> TransposeCSR (MyHyb.h:31-75)."""
    rp, ci = _c(row_ptrs).copy(), _c(col).copy()
    perm = np.empty(ci.size, dtype=np.int64)
    _load().ref_transpose_csr(_p(rp), ctypes.c_int64(rp.size - 1), _p(ci), ctypes.c_int64(ci.size), _p(perm))
    return rp, ci, _c(rel)[perm], _c(eid)[perm]


def _sep(fn, row, col, rel, eid, num_rows, num_rels, ptr_len):
    row, col, rel, eid = map(_c, (row, col, rel, eid))
    e = row.size
    rel_ptrs = np.zeros(num_rels + 1, dtype=np.int64)
    second = np.zeros(ptr_len, dtype=np.int64)
    oc, oe = np.zeros(e, dtype=np.int64), np.zeros(e, dtype=np.int64)
    fn(_p(row), _p(col), _p(rel), _p(eid), ctypes.c_int64(e), ctypes.c_int64(num_rows),
       ctypes.c_int64(num_rels), _p(rel_ptrs), _p(second), _p(oc), _p(oe))
    return rel_ptrs, second, oc, oe


def coo_to_separate_coo(row, col, rel, eid, num_rows, num_rels):
    """ConvertIntegratedCOOToSeparateCOO (MyHyb.h:1047-1097) -> rel_ptrs,row,col,eid"""
    return _sep(_load().ref_coo_to_separate_coo, row, col, rel, eid, num_rows, num_rels, len(row))


def coo_to_separate_csr(row, col, rel, eid, num_rows, num_rels):
    """ConvertIntegratedCOOToSeparateCSR (MyHyb.h:996-1045) -> rel_ptrs,row_ptrs,col,eid"""
    return _sep(_load().ref_coo_to_separate_csr, row, col, rel, eid, num_rows, num_rels,
                num_rows * num_rels + 1)
