"""Bit-exact parity of the GPU relation sort / CSR construction (through the
torch_hrt ops, i.e. through the C ABI) against oracle/converters_ref.py."""
import numpy as np
import pytest
import torch

from oracle import converters_ref as C

pytestmark = pytest.mark.gpu


def _rand_coo(n, e, r, seed):
    rng = np.random.default_rng(seed)
    return (rng.integers(0, n, e, dtype=np.int64), rng.integers(0, n, e, dtype=np.int64),
            rng.integers(0, r, e, dtype=np.int64), rng.permutation(e).astype(np.int64))


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _eq(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert np.array_equal(g.cpu().numpy(), w), "index arrays differ"


@pytest.mark.parametrize("n,e,r", [(1, 1, 1), (7, 0, 3), (100, 1000, 5), (5000, 70001, 300), (70000, 300000, 2)])
def test_converters_bit_exact(cuda, n, e, r):
    from het_b200.kernels import K
    src, dst, rel, eid = _rand_coo(n, e, r, seed=n + e)
    ptr, col, rl, ei = C.coo2csr(src, dst, rel, eid, n)
    # generic stable sort (COO -> CSR)
    got = K.hetb200_stable_sort_by_key(_t(src, cuda), n, [_t(dst, cuda), _t(rel, cuda), _t(eid, cuda)])
    _eq([got[3], got[0], got[1], got[2]], [ptr, col, rl, ei])
    # transpose_csr (argument / return order: row_ptrs, col, reltypes, eids)
    got = K.transpose_csr(_t(ptr, cuda), _t(col, cuda), _t(rl, cuda), _t(ei, cuda))
    _eq(got, C.transpose_csr(ptr, col, rl, ei))
    # the reference's only unit test: transposing twice preserves row_ptrs
    # (hrt/python/test/test_kernel_correctness.py:44)
    back = K.transpose_csr(*got)
    assert np.array_equal(back[0].cpu().numpy(), ptr)
    if e == 0:
        return
    got = K.convert_integrated_coo_to_separate_coo(_t(src, cuda), _t(dst, cuda), _t(rel, cuda), _t(eid, cuda), n, r)
    _eq(got, C.to_separate_coo(src, dst, rel, eid, n, r))
    got = K.convert_integrated_csr_to_separate_coo(_t(ptr, cuda), _t(col, cuda), _t(rl, cuda), _t(ei, cuda))
    _eq(got, C.csr_to_separate_coo(ptr, col, rl, ei))
    got = K.convert_integrated_coo_to_separate_csr(_t(src, cuda), _t(dst, cuda), _t(rel, cuda), _t(eid, cuda), n, r)
    _eq(got, C.to_separate_csr(src, dst, rel, eid, n, r))


def test_converters_accept_cpu_tensors(cuda):
    """The reference's converters take CPU tensors (DataConverters.inc.h asserts
    is_cpu); ours stage them through the GPU and return CPU tensors."""
    from het_b200.kernels import K
    src, dst, rel, eid = _rand_coo(50, 400, 4, seed=9)
    ptr, col, rl, ei = C.coo2csr(src, dst, rel, eid, 50)
    got = K.transpose_csr(*(torch.from_numpy(a) for a in (ptr, col, rl, ei)))
    assert all(not t.is_cuda for t in got)
    _eq(got, C.transpose_csr(ptr, col, rl, ei))


def test_graph_container_matches_oracle(cuda):
    from het_b200 import graphs
    from het_b200.graph import HetGraph
    coo = graphs.bgs_shaped(scale=0.05)
    g = HetGraph(coo, cuda)
    E, N, R = coo.num_edges, coo.num_nodes, coo.num_rels
    eid = np.arange(E, dtype=np.int64)
    rel_ptrs, row, col, e0 = C.to_separate_coo(coo.src, coo.dst, coo.etype, eid, N, R)
    row, col, e0 = C.sort_by_etype_eids(rel_ptrs, row, col, e0)
    optr, ocol, orel, oeid = C.coo2csr(coo.src, coo.dst, coo.etype, eid, N)
    iptr, icol, irel, ieid = C.coo2csr(coo.dst, coo.src, coo.etype, eid, N)
    ce, oeid, ieid = C.canonicalize_eids(e0, oeid, ieid)
    s = g.get_separate_coo_original()
    _eq([s["rel_ptrs"], s["row_indices"], s["col_indices"], s["eids"]], [rel_ptrs, row, col, ce])
    o = g.get_out_csr()
    _eq([o["row_ptrs"], o["col_indices"], o["rel_types"], o["eids"]], [optr, ocol, orel, oeid])
    i = g.get_in_csr()
    _eq([i["row_ptrs"], i["col_indices"], i["rel_types"], i["eids"]], [iptr, icol, irel, ieid])
