"""Host logic of the walk plans (het_b200/plan.py), checked on CPU tensors.

The kernels in csrc/edge_plan.cu trust the plan: a row that is not flagged as split is STORED by
exactly one lane group, a flagged one is pre-zeroed and reduced, partial slots of the forward walk
are exactly the (item, head|tail) pairs the merge kernel visits. These tests emulate the walkers
with Python loops on small random graphs (hubs, empty rows, long runs) using the packed plan the
kernels would read, and compare the result with a direct evaluation."""
import numpy as np
import pytest
import torch

from het_b200 import plan as P


def _unpack(wp):
    m = wp.meta
    names = P._ORDER
    counts = dict(num_positions=m[0], num_items=m[1], num_nz_rows=m[2], num_fix_rows=m[3], num_empty_rows=m[4],
                  num_pslots=m[5], num_fix_runs=m[6])
    sizes = dict(item_ptr=m[1] + 1 if m[7] >= 0 else 0, item_row=m[1], nz_rows=m[2], nz_ptr=m[2] + 1, fix_rows=m[3],
                 empty_rows=m[4], pslot=m[5], pslot_row=m[5], fix_runs=m[6], idx_a=m[0], idx_b=m[0], idx_c=m[0],
                 idx_d=m[0], row_rel=None)
    blob = wp.blob.numpy()
    out = dict(counts)
    for i, n in enumerate(names):
        off = m[7 + i]
        if off < 0:
            out[n] = None
        elif sizes[n] is None:
            out[n] = blob[off:]
        else:
            out[n] = blob[off:off + sizes[n]]
    return out


def _random_graph(seed, n_dst, n_src, e, r, hub):
    g = np.random.default_rng(seed)
    src = g.integers(0, n_src, e)
    dst = g.integers(0, n_dst, e)
    if hub:
        dst[: e // 2] = 3
        src[e // 3: e // 3 + 100] = 5      # a hub source with a long run
    rel = g.integers(0, r, e)
    if hub:
        rel[e // 3: e // 3 + 100] = 1
    return src, dst, rel


def _compact(rel, node, n):
    key = rel.astype(np.int64) * n + node
    uniq, inv = np.unique(key, return_inverse=True)
    return uniq // n, inv  # relation of each compact row, compact row of each edge


def _plans(src, dst, rel, n_dst, n_src, item):
    urow_rel, inv_row = _compact(rel, src, n_src)
    ucol_rel, inv_col = _compact(rel, dst, n_dst)
    e = src.size
    base = np.lexsort((np.arange(e), rel))                 # the (rel, eid) order of the separate COO
    f = base[np.argsort(dst[base], kind="stable")]         # forward positions: (dst, rel, eid)
    b = base[np.argsort(src[base], kind="stable")]         # backward positions: (src, rel, eid)
    fpos = np.empty(e, dtype=np.int64)
    fpos[f] = np.arange(e)
    T = torch.from_numpy
    fwd = P.pack(P.forward_arrays(T(dst[f]), n_dst, T(src[f]), T(inv_row[f]), T(inv_col[f]), T(ucol_rel), item), "cpu")
    bwd = P.pack(P.backward_arrays(T(src[b]), n_src, T(dst[b]), T(inv_row[b]), T(inv_col[b]), T(fpos[b]),
                                   T(urow_rel), item), "cpu")
    return fwd, bwd, f, b, inv_row, inv_col, urow_rel, ucol_rel


CASES = [(0, 40, 40, 500, 3, False, 32), (1, 64, 64, 3000, 4, True, 32), (2, 10, 1000, 700, 2, True, 8),
         (3, 300, 7, 900, 5, False, 16), (4, 5, 5, 31, 2, False, 32), (5, 5, 5, 32, 1, False, 32),
         (6, 50, 50, 1, 1, False, 32), (7, 20, 20, 2000, 1, True, 4)]


@pytest.mark.parametrize("seed,n_dst,n_src,e,r,hub,item", CASES)
def test_forward_plan_emulated_walk_matches_softmax_aggregation(seed, n_dst, n_src, e, r, hub, item):
    src, dst, rel = _random_graph(seed, n_dst, n_src, e, r, hub)
    fwd, _, f, _, inv_row, inv_col, urow_rel, ucol_rel = _plans(src, dst, rel, n_dst, n_src, item)
    pl = _unpack(fwd)
    g = np.random.default_rng(seed + 100)
    x = g.normal(size=e) * 30.0          # logits by edge id, large enough that a naive exp overflows fp32
    v = g.normal(size=(e, 3))
    ip = pl["item_ptr"]
    assert ip[0] == 0 and ip[-1] == e and np.all(np.diff(ip) > 0) and np.diff(ip).max() < 2 * item
    nz_rows = pl["nz_rows"] & 0x7fffffff
    nz_ptr = pl["nz_ptr"]
    # every position's destination is the row the plan says it is in
    for j in range(pl["num_nz_rows"]):
        assert np.all(dst[f[nz_ptr[j]:nz_ptr[j + 1]]] == nz_rows[j])
    assert np.array_equal(pl["idx_a"], src[f]) and np.array_equal(pl["idx_b"], inv_row[f])
    assert np.array_equal(pl["idx_c"], inv_col[f])
    assert np.array_equal(pl["row_rel"][: ucol_rel.size], ucol_rel)

    NEG = -3.0e38
    ret = np.full((n_dst, 3), np.nan)
    lse = np.full(n_dst, np.nan)
    ssum = np.full(n_dst, np.nan)
    for d in pl["fix_rows"]:
        ret[d], lse[d], ssum[d] = 0.0, NEG, 0.0
    for d in pl["empty_rows"]:
        ret[d], lse[d] = 0.0, 0.0
    slots = {}
    stores = np.zeros(n_dst, dtype=int)
    for gi in range(pl["num_items"]):
        p0, p1 = ip[gi], ip[gi + 1]
        j = pl["item_row"][gi]
        d, row_start, row_end = nz_rows[j], nz_ptr[j], nz_ptr[j + 1]
        assert row_start <= p0 < row_end
        m, s, acc = NEG, 0.0, np.zeros(3)

        def flush():
            if row_start >= p0 and row_end <= p1:
                assert not (pl["nz_rows"][j] < 0), "a row flagged split was completed inside one item"
                ret[d] = acc / s
                lse[d] = m + np.log(s)
                stores[d] += 1
            else:
                assert pl["nz_rows"][j] < 0, "a split row is not flagged"
                slot = 2 * gi + (0 if row_start < p0 else 1)
                assert slot not in slots
                slots[slot] = (d, m, s, acc.copy())
                lse[d] = max(lse[d], m)

        for p in range(p0, p1):
            if p == row_end:
                flush()
                j += 1
                d, row_start, row_end = nz_rows[j], p, nz_ptr[j + 1]
                m, s, acc = NEG, 0.0, np.zeros(3)
            eid = f[p]
            mn = max(m, x[eid])
            sc, w = np.exp(m - mn), np.exp(x[eid] - mn)
            acc = acc * sc + w * v[eid]
            s = s * sc + w
            m = mn
        flush()
    # the merge kernel visits exactly the valid slots
    assert sorted(slots) == sorted(pl["pslot"].tolist())
    for sl, d in zip(pl["pslot"], pl["pslot_row"]):
        dd, m, s, acc = slots[int(sl)]
        assert dd == d
        w = np.exp(m - lse[d])
        ret[d] += w * acc
        ssum[d] += w * s
    for d in pl["fix_rows"]:
        ret[d] /= ssum[d]
        lse[d] += np.log(ssum[d])
    has = np.bincount(dst, minlength=n_dst) > 0
    assert np.all(stores[has & ~np.isin(np.arange(n_dst), pl["fix_rows"])] == 1)
    # direct evaluation
    want = np.zeros((n_dst, 3))
    want_lse = np.zeros(n_dst)
    for d in range(n_dst):
        idx = np.nonzero(dst == d)[0]
        if idx.size:
            mx = x[idx].max()
            w = np.exp(x[idx] - mx)
            want[d] = (w[:, None] * v[idx]).sum(0) / w.sum()
            want_lse[d] = mx + np.log(w.sum())
    assert not np.isnan(ret).any() and not np.isnan(lse).any()
    np.testing.assert_allclose(ret, want, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(lse[has], want_lse[has], rtol=1e-9, atol=1e-9)

    # second walk of the two-pass HGT backward (plan_hgt_grad_left_kernel) over the same items: a compact
    # (relation, destination) row is stored by exactly one lane group unless an item boundary cuts it, in
    # which case it is listed in fix_runs (pre-zeroed) and its pieces reduce
    Uc = ucol_rel.size
    val = g.normal(size=e)
    gl = np.full(Uc, np.nan)
    for uc in pl["fix_runs"]:
        gl[uc] = 0.0
    fix = set(int(t) for t in pl["fix_runs"])
    stores_l = np.zeros(Uc, dtype=int)
    left_of = pl["idx_c"]
    for gi in range(pl["num_items"]):
        p0, p1 = ip[gi], ip[gi + 1]
        cur = left_of[p0]
        open_head = p0 > 0 and left_of[p0 - 1] == cur
        acc = 0.0

        def flush_l(open_tail):
            if open_head or open_tail:
                assert int(cur) in fix, "a cut run is not listed"
                gl[cur] += acc
            else:
                assert np.isnan(gl[cur]) and int(cur) not in fix
                gl[cur] = acc
                stores_l[cur] += 1

        for p in range(p0, p1):
            if left_of[p] != cur:
                flush_l(False)
                open_head = False
                cur, acc = left_of[p], 0.0
            acc += val[f[p]]
        flush_l(p1 < e and left_of[p1] == cur)
    want_l = np.zeros(Uc)
    np.add.at(want_l, inv_col, val)
    assert not np.isnan(gl).any()
    np.testing.assert_allclose(gl, want_l, rtol=1e-9, atol=1e-9)
    runlen_c = np.bincount(inv_col, minlength=Uc)
    assert all(runlen_c[t] > 1 for t in fix)


@pytest.mark.parametrize("seed,n_dst,n_src,e,r,hub,item", CASES)
def test_backward_plan_emulated_walk_owns_every_row_and_run_once(seed, n_dst, n_src, e, r, hub, item):
    src, dst, rel = _random_graph(seed, n_dst, n_src, e, r, hub)
    _, bwd, f, b, inv_row, inv_col, urow_rel, ucol_rel = _plans(src, dst, rel, n_dst, n_src, item)
    pl = _unpack(bwd)
    ip = pl["item_ptr"]
    assert ip[0] == 0 and ip[-1] == e and np.all(np.diff(ip) > 0) and np.diff(ip).max() < 2 * item
    nz_rows, nz_ptr = pl["nz_rows"] & 0x7fffffff, pl["nz_ptr"]
    relw = pl["row_rel"][: urow_rel.size]
    assert np.array_equal(relw & 0x7fffffff, urow_rel)
    assert np.array_equal(pl["idx_a"], dst[b]) and np.array_equal(pl["idx_b"], inv_row[b])
    assert np.array_equal(pl["idx_c"], inv_col[b])
    fpos = np.empty(e, dtype=np.int64)
    fpos[f] = np.arange(e)
    assert np.array_equal(pl["idx_d"], fpos[b])           # forward position of the same edge
    g = np.random.default_rng(seed + 7)
    val = g.normal(size=e)                                # one scalar contribution per edge (by edge id)
    U = urow_rel.size
    grad_row = np.full(n_src, np.nan)                     # per-source accumulator (grad_kv)
    grad_run = np.full(U, np.nan)                         # per-run accumulator (grad_msg)
    for s_ in pl["fix_rows"]:
        grad_row[s_] = 0.0
    for u in pl["fix_runs"]:
        grad_run[u] = 0.0
    row_stores, run_stores = np.zeros(n_src, int), np.zeros(U, int)
    for gi in range(pl["num_items"]):
        p0, p1 = ip[gi], ip[gi + 1]
        j = pl["item_row"][gi]
        word, srow, row_end = pl["nz_rows"][j], nz_rows[j], nz_ptr[j + 1]
        assert nz_ptr[j] <= p0 < row_end
        acc_k, cur_u, acc_m = 0.0, -1, 0.0

        def flush_run():
            if cur_u < 0:
                return
            if relw[cur_u] < 0:
                grad_run[cur_u] += acc_m
            else:
                assert np.isnan(grad_run[cur_u]), "an unsplit run was stored twice"
                grad_run[cur_u] = acc_m
                run_stores[cur_u] += 1

        def flush_row():
            if word < 0:
                grad_row[srow] += acc_k
            else:
                assert np.isnan(grad_row[srow]), "an unsplit row was stored twice"
                grad_row[srow] = acc_k
                row_stores[srow] += 1

        for p in range(p0, p1):
            if p == row_end:
                flush_run(); cur_u = -1
                flush_row()
                j += 1
                word, srow, row_end = pl["nz_rows"][j], nz_rows[j], nz_ptr[j + 1]
                acc_k = 0.0
            if pl["idx_b"][p] != cur_u:
                flush_run()
                cur_u, acc_m = pl["idx_b"][p], 0.0
            assert src[b[p]] == srow
            acc_k += val[b[p]]
            acc_m += 2.0 * val[b[p]]
        flush_run()
        flush_row()
    want_row = np.zeros(n_src)
    np.add.at(want_row, src, val)
    want_run = np.zeros(U)
    np.add.at(want_run, inv_row, 2.0 * val)
    assert not np.isnan(grad_row).any() and not np.isnan(grad_run).any()
    np.testing.assert_allclose(grad_row, want_row, rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(grad_run, want_run, rtol=1e-9, atol=1e-9)
    # units no longer than an item are never split
    outdeg = np.bincount(src, minlength=n_src)
    runlen = np.bincount(inv_row, minlength=U)
    for u in pl["fix_runs"]:
        assert runlen[u] > item
    for s_ in pl["fix_rows"]:
        assert outdeg[s_] == 0 or outdeg[s_] > 1       # an empty row, or a row with an interior boundary


def test_plan_cache_round_trip_and_key(tmp_path):
    """On-disk plan cache (het_b200/plan.py: graph_key / save_plans / load_plans): what is loaded is what was
    saved; another graph, another item size or a truncated file is a miss, never a wrong plan."""
    src, dst, rel = _random_graph(3, 60, 50, 900, 4, True)
    fwd, bwd, f, b, *_ = _plans(src, dst, rel, 60, 50, 16)
    T = torch.from_numpy
    key = P.graph_key(60, 50, 4, T(src), T(dst), T(rel))
    assert key == P.graph_key(60, 50, 4, T(src), T(dst), T(rel))
    src2 = src.copy()
    src2[17] = (src2[17] + 1) % 50
    assert key != P.graph_key(60, 50, 4, T(src2), T(dst), T(rel))
    assert key != P.graph_key(60, 50, 4, T(src), T(dst), T(rel), item=8)
    path = str(tmp_path / "plan.pt")
    assert P.load_plans(path, key, "cpu") is None
    P.save_plans(path, key, fwd, bwd, T(f))
    got = P.load_plans(path, key, "cpu")
    assert got is not None
    for a, w in ((got[0], fwd), (got[1], bwd)):
        assert torch.equal(a.blob, w.blob) and a.meta == w.meta and a.num_items == w.num_items
        assert a.num_positions == w.num_positions
    assert torch.equal(got[2], T(f))
    assert P.load_plans(path, "0" * 40, "cpu") is None       # a plan of another graph
    with open(path, "r+b") as fh:
        fh.truncate(100)
    assert P.load_plans(path, key, "cpu") is None            # a damaged file
