"""Generates tests/golden/converters_*.npz from the REFERENCE's own C++ converters
(oracle/_ref/libhet_ref_converters.so = /root/reference/hrt/include/MyHyb/MyHyb.h compiled by
oracle/build_ref.py). Run in the build container (needs /root/reference):

    python -m oracle.build_ref && python tests/golden/make_converter_golden.py

Inputs are seeded integrated COOs; outputs are what TransposeCSR /
ConvertIntegratedCOOToSeparateCOO / ConvertIntegratedCOOToSeparateCSR return.
The case "gnp100" reproduces the shape of the reference's only unit test
(hrt/python/test/test_kernel_correctness.py:8-44: G(n=100, p=0.1), seed 0, random rel types).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import converters_ref as C  # noqa: E402
from oracle import ref_converters as R  # noqa: E402

CASES = {
    # name: (num_nodes, num_rels, num_edges, seed)
    "gnp100": (100, 4, 1000, 0),
    "tiny": (5, 3, 12, 1),
    "ragged": (64, 17, 300, 2),      # several empty relations and empty rows
    "onerel": (40, 1, 200, 3),
    "multi": (300, 11, 5000, 4),     # duplicate (src,dst,rel) edges
}


def make_case(n, r, e, seed):
    rng = np.random.default_rng(seed)
    src = rng.integers(0, n, e, dtype=np.int64)
    dst = rng.integers(0, n, e, dtype=np.int64)
    # leave some relations empty (but keep max(rel) == r-1 so max+1 inference agrees)
    rel = rng.integers(0, max(1, r // 2), e, dtype=np.int64) * 2 % r if r > 2 else rng.integers(0, r, e, dtype=np.int64)
    rel[0] = r - 1
    eid = rng.permutation(e).astype(np.int64)
    return src, dst, rel, eid


def main():
    assert R.available(), "build oracle/_ref first"
    for name, (n, r, e, seed) in CASES.items():
        src, dst, rel, eid = make_case(n, r, e, seed)
        out = {"src": src, "dst": dst, "rel": rel, "eid": eid, "num_nodes": n, "num_rels": r}
        rp, col, rl, ei = C.coo2csr(src, dst, rel, eid, n)  # input CSR for the transpose
        out.update(csr_row_ptrs=rp, csr_col=col, csr_rel=rl, csr_eid=ei)
        t = R.transpose_csr(rp, col, rl, ei)
        out.update(t_row_ptrs=t[0], t_col=t[1], t_rel=t[2], t_eid=t[3])
        s = R.coo_to_separate_coo(src, dst, rel, eid, n, r)
        out.update(scoo_rel_ptrs=s[0], scoo_row=s[1], scoo_col=s[2], scoo_eid=s[3])
        c = R.coo_to_separate_csr(src, dst, rel, eid, n, r)
        out.update(scsr_rel_ptrs=c[0], scsr_row_ptrs=c[1], scsr_col=c[2], scsr_eid=c[3])
        np.savez_compressed(os.path.join(os.path.dirname(__file__), f"converters_{name}.npz"), **out)
        print(name, "ok")


if __name__ == "__main__":
    main()
