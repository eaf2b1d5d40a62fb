"""CPU oracle for relation sort / CSR construction — TEST INFRASTRUCTURE ONLY.

numpy restatement (int64, bit-exact) of the reference's converters. Every
ordering is defined with *stable* sorts so the result is unique:

* ``coo2csr``            python/utils_lite/sparse_matrix_converters.py:7-39
                         (the reference's argsort is unstable, so its in-row
                         order is unspecified; we define it as stable)
* ``transpose_csr``      MyHyb/MyHyb.h:31-75 (new row = old col; inside a row:
                         ascending old row, then original position)
* ``to_separate_coo``    MyHyb/MyHyb.h:1047-1150 (order rel, row, original)
* ``sort_by_etype_eids`` python/utils/coo_sorters.py:156-170 (per relation by eid)
* ``to_separate_csr``    MyHyb/MyHyb.h:996-1045
* ``canonicalize_eids``  python/utils/mydgl_graph.py:736-824

Pinned against the reference's own C++ (MyHyb.h compiled into oracle/_ref by
oracle/build_ref.py) in tests/test_oracle_converters.py, and against the one
property the reference's test asserts (transpose twice preserves row_ptrs,
python/test/test_kernel_correctness.py:44).
"""
from __future__ import annotations

import numpy as np

I64 = np.int64


def _ptrs(keys: np.ndarray, n: int) -> np.ndarray:
    p = np.zeros(n + 1, dtype=I64)
    np.cumsum(np.bincount(keys, minlength=n), out=p[1:])
    return p


def expand_rows(row_ptrs: np.ndarray) -> np.ndarray:
    return np.repeat(np.arange(row_ptrs.size - 1, dtype=I64), np.diff(row_ptrs))


def coo2csr(src, dst, rel, eid, num_rows):
    o = np.argsort(src, kind="stable")
    return _ptrs(src, num_rows), dst[o].astype(I64), rel[o].astype(I64), eid[o].astype(I64)


def transpose_csr(row_ptrs, col, rel, eid):
    n = row_ptrs.size - 1
    rows = expand_rows(row_ptrs)
    o = np.argsort(col, kind="stable")
    return _ptrs(col, n), rows[o], rel[o].astype(I64), eid[o].astype(I64)


def to_separate_coo(row, col, rel, eid, num_rows, num_rels):
    """order (rel, row, original position) -> rel_ptrs, row, col, eid"""
    o = np.argsort(rel * I64(num_rows) + row, kind="stable")
    return _ptrs(rel, num_rels), row[o].astype(I64), col[o].astype(I64), eid[o].astype(I64)


def csr_to_separate_coo(row_ptrs, col, rel, eid):
    num_rels = int(rel.max()) + 1 if rel.size else 0
    return to_separate_coo(expand_rows(row_ptrs), col, rel, eid, row_ptrs.size - 1, num_rels)


def to_separate_csr(row, col, rel, eid, num_rows, num_rels):
    key = rel * I64(num_rows) + row
    o = np.argsort(key, kind="stable")
    row_ptrs = _ptrs(key, num_rows * num_rels)
    return row_ptrs[::num_rows].copy(), row_ptrs, col[o].astype(I64), eid[o].astype(I64)


def separate_csr_rel_ptrs_ref_quirk(row_ptrs, num_rows, num_rels):
    """What the reference actually returns as rel_ptrs from
    ConvertIntegratedCOOToSeparateCSR (MyHyb.h:1039-1041): it adds the GLOBAL running
    row_ptrs value at the end of each relation instead of that relation's edge count, so
    rel_ptrs[r+1] = sum_{q<=r} (edges in relations 0..q). A reference bug (documented,
    not copied: ``to_separate_csr`` returns the true offsets); kept so the golden
    vectors pin the rest of the function bit-exactly."""
    ends = row_ptrs[num_rows::num_rows][:num_rels]
    out = np.zeros(num_rels + 1, dtype=I64)
    np.cumsum(ends, out=out[1:])
    return out


def sort_by_etype_eids(rel_ptrs, row, col, eid):
    row, col, eid = row.copy(), col.copy(), eid.copy()
    for r in range(rel_ptrs.size - 1):
        a, b = int(rel_ptrs[r]), int(rel_ptrs[r + 1])
        o = np.argsort(eid[a:b], kind="stable")
        row[a:b], col[a:b], eid[a:b] = row[a:b][o], col[a:b][o], eid[a:b][o]
    return row, col, eid


def canonicalize_eids(separate_coo_eids, *other_eid_arrays):
    """remap so the separate-COO eids become 0..E-1 in order."""
    remap = np.empty(separate_coo_eids.size, dtype=I64)
    remap[separate_coo_eids] = np.arange(separate_coo_eids.size, dtype=I64)
    return (np.arange(separate_coo_eids.size, dtype=I64),) + tuple(remap[a] for a in other_eid_arrays)


def unique_node_indices_single_sided_ref(rel_ptrs, row_indices, col_indices):
    """generate_separate_unique_node_indices_single_sided_for_each_etype
    (hrt/python/utils_lite/mydgl_graph_methods.py:11-157): for every relation, the sorted unique
    sources (``_row``) / destinations (``_col``) of its edges with the inverse map, concatenated
    with running offsets. Edge ids are separate-COO positions (canonicalised), so the inverse map
    in COO order is the reference's ``edata_idx_to_inverse_idx``."""
    out = {}
    R = len(rel_ptrs) - 1
    for side, nodes in (("row", row_indices), ("col", col_indices)):
        ptrs, lists, inv = [0], [], np.empty(len(nodes), dtype=np.int64)
        for r in range(R):
            a, b = int(rel_ptrs[r]), int(rel_ptrs[r + 1])
            uniq, inverse = np.unique(nodes[a:b], return_inverse=True)
            inv[a:b] = inverse + ptrs[-1]
            lists.append(uniq)
            ptrs.append(ptrs[-1] + len(uniq))
        out["rel_ptrs_" + side] = np.asarray(ptrs, dtype=np.int64)
        out["node_indices_" + side] = (np.concatenate(lists) if lists else np.empty(0)).astype(np.int64)
        out["inverse_idx_" + side] = inv
    return out
