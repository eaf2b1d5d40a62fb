"""GPU: the allocation-free entry point hetb_hgt_enorm_to_unnormalized_bwd_static (workspace in a static
device array inside the library) against the existing hetb_hgt_enorm_to_unnormalized_bwd (workspace from the
caller), both called through the C ABI with raw device pointers, twice in a row on different inputs (the
array keeps the previous call's contents). Collected last (file name) so that a failure here hides nothing."""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "het_b200", "lib", "libhetb200.so")
STATIC_SCRATCH_FLOATS = 1 << 20   # HETB_STATIC_SCRATCH_FLOATS in include/hetb200.h
HETB_ETOOLARGE = -4

P, I64, I32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int


@pytest.fixture(scope="module")
def lib():
    assert os.path.exists(LIB), "libhetb200.so is not built (python -c 'import __graft_entry__ as g; g.build()')"
    lib = ctypes.CDLL(LIB)
    lib.hetb_hgt_enorm_to_unnormalized_bwd.restype = I32
    lib.hetb_hgt_enorm_to_unnormalized_bwd.argtypes = [P, P, P, P, I64, I64, I64, P, P, P, P, P, P, P, I32, P]
    lib.hetb_hgt_enorm_to_unnormalized_bwd_static.restype = I32
    lib.hetb_hgt_enorm_to_unnormalized_bwd_static.argtypes = [P, P, P, P, I64, I64, I64, P, P, P, P, P, P, I32, P]
    return lib


def _case(dev, n, e, r, H, seed):
    g = torch.Generator().manual_seed(seed)
    et = torch.sort(torch.randint(0, r, (e,), generator=g)).values        # positions sorted by relation
    rel_ptrs = torch.searchsorted(et, torch.arange(r + 1)).to(torch.int64)
    dst = torch.randint(0, n, (e,), generator=g)
    dst[: e // 8] = 0                                                      # a hub destination
    eids = torch.randperm(e, generator=g)
    attn = torch.randn(e, H, generator=g)
    mu = torch.rand(r, H, generator=g) + 0.5
    # a normalised score per (destination, head): softmax of attn * mu over each destination's edges
    x = attn[eids] * mu[et]
    mx = torch.full((n, H), -1e30).scatter_reduce(0, dst[:, None].expand(e, H), x, "amax")
    ex = torch.exp(x - mx[dst])
    norm = torch.empty(e, H)
    norm[eids] = ex / torch.zeros(n, H).index_add_(0, dst, ex)[dst]
    grad_norm = torch.randn(e, H, generator=g)
    to = lambda t: t.to(dev).contiguous()
    return dict(dst=to(dst), eids=to(eids), rel_ptrs=to(rel_ptrs), attn=to(attn), norm=to(norm),
                grad_norm=to(grad_norm), mu=to(mu), n=n, e=e, r=r, H=H)


def _run(lib, c, static):
    dev = c["attn"].device
    grad_attn = torch.full((c["e"], c["H"]), float("nan"), device=dev)
    grad_mu = torch.zeros(c["r"], c["H"], device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    head = (c["dst"].data_ptr(), c["eids"].data_ptr(), None, c["rel_ptrs"].data_ptr(), c["r"], c["n"], c["e"],
            c["attn"].data_ptr(), c["norm"].data_ptr(), c["grad_norm"].data_ptr(), c["mu"].data_ptr())
    if static:
        rc = lib.hetb_hgt_enorm_to_unnormalized_bwd_static(*head, grad_attn.data_ptr(), grad_mu.data_ptr(), c["H"], stream)
    else:
        scratch = torch.full((c["n"], c["H"]), float("nan"), device=dev)   # the entry point zeroes it
        rc = lib.hetb_hgt_enorm_to_unnormalized_bwd(*head, scratch.data_ptr(), grad_attn.data_ptr(),
                                                    grad_mu.data_ptr(), c["H"], stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    return grad_attn, grad_mu


def test_static_scratch_entry_point_matches_existing_one_twice(cuda, lib):
    from oracle.testing import close
    # two calls in a row, the second smaller than the first: stale sums of call 1 lie beyond and inside its range
    for n, e, r, H, seed in ((5000, 60000, 7, 4, 1), (1200, 9000, 3, 4, 2)):
        c = _case(cuda, n, e, r, H, seed)
        want_attn, want_mu = _run(lib, c, static=False)
        got_attn, got_mu = _run(lib, c, static=True)
        assert torch.isfinite(got_attn).all()
        # same kernels, atomics in another order: the tolerance the other GPU op tests use
        close(got_attn, want_attn.double().cpu(), f"grad_attn (n={n})", rtol=1e-4, atol=1e-5)
        close(got_mu, want_mu.double().cpu(), f"grad_mu (n={n})", rtol=1e-4, atol=2e-5)


def test_static_scratch_entry_point_rejects_oversize(cuda, lib):
    H = 4
    n = STATIC_SCRATCH_FLOATS // H + 1
    z = torch.zeros(8, device=cuda)
    zi = torch.zeros(8, dtype=torch.int64, device=cuda)
    rc = lib.hetb_hgt_enorm_to_unnormalized_bwd_static(
        zi.data_ptr(), zi.data_ptr(), None, zi.data_ptr(), 1, n, 2, z.data_ptr(), z.data_ptr(), z.data_ptr(),
        z.data_ptr(), z.data_ptr(), z.data_ptr(), H, torch.cuda.current_stream().cuda_stream)
    assert rc == HETB_ETOOLARGE
    torch.cuda.synchronize()
