"""Walk plans: the per-graph work decomposition and int32 index copies the plan-based fused edge
kernels read (csrc/edge_plan.cu, include/hetb200.h: hetb_walk_plan; SURVEY.md §8(b)
"hetb_graph_plan", §8(d) "i = 4").

Built once per graph, on the device the graph lives on, from the relation-sorted separate COO and
the compact-as-of-node lists (the reference's formats: MyHyb/MyHyb.h:1099-1150,
python/utils/coo_sorters.py:156-170, python/utils_lite/mydgl_graph_methods.py:11-157). The two
builders below are pure tensor functions (they run on CPU tensors too, which is how
tests/test_plan_cpu.py checks them against a brute-force restatement); the sorts that feed them are
the library's own stable radix sort (``K.hetb200_stable_sort_by_key``) on the GPU.

* forward plan: positions sorted by (destination, relation[, edge id]); rows = destinations.
* backward plan: positions sorted by (source, relation[, edge id]); rows = sources, runs = compact
  (relation, source) rows.

Work items are ~``ITEM`` consecutive positions. A candidate boundary k·ITEM that falls inside a
unit (forward: a destination row; backward: a run) no longer than ITEM is moved back to the unit's
start, so such units are never split: only power-law hubs are, and those are listed
(``fix_rows`` / ``fix_runs`` / the forward partial slots) instead of discovered at run time. A forward
plan also lists the compact (relation, destination) rows cut by an item boundary (``fix_runs``): the
second walk of the two-pass HGT backward owns every other such row outright.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

ITEM = 32
META_LEN = 21
_SPLIT = -(1 << 31)  # adding this to a non-negative int32 sets bit 31


def _items(unit_ptr: torch.Tensor, num_pos: int, item: int) -> torch.Tensor:
    """Item boundaries [G+1] (int64, strictly increasing, 0 .. num_pos) for units starting at
    ``unit_ptr`` ([U+1], unit_ptr[-1] == num_pos, every unit non-empty)."""
    dev = unit_ptr.device
    ends = torch.tensor([0, num_pos], dtype=torch.int64, device=dev)
    if num_pos <= item:
        return torch.unique(ends)
    k = torch.arange(item, num_pos, item, dtype=torch.int64, device=dev)
    u = torch.searchsorted(unit_ptr, k, right=True) - 1
    start = unit_ptr[u]
    b = torch.where(unit_ptr[u + 1] - start <= item, start, k)
    return torch.unique(torch.cat([ends, b]))


def _nonempty(keys_sorted: torch.Tensor, num_rows: int):
    counts = torch.bincount(keys_sorted, minlength=num_rows)
    nz = torch.nonzero(counts, as_tuple=False).view(-1)
    ptr = torch.zeros(nz.numel() + 1, dtype=torch.int64, device=keys_sorted.device)
    torch.cumsum(counts[nz], 0, out=ptr[1:])
    empty = torch.nonzero(counts == 0, as_tuple=False).view(-1)
    return nz, ptr, empty


def _split_units(unit_ptr, item_ptr):
    """indices of the units that contain an interior item boundary which is not their start"""
    ib = item_ptr[1:-1]
    if ib.numel() == 0:
        return ib
    u = torch.searchsorted(unit_ptr, ib, right=True) - 1
    return torch.unique(u[unit_ptr[u] != ib])


def forward_arrays(dst_sorted, num_dst, idx_src, idx_urow, idx_ucol, ucol_rel, item=ITEM):
    """Arrays of a forward plan. ``dst_sorted`` [E] non-decreasing destination of every position;
    ``idx_*`` [E] position-aligned (source node, compact (rel, src) row, compact (rel, dst) row);
    ``ucol_rel`` [Uc] relation of every compact (rel, dst) row. Returns a dict of int64 tensors."""
    E = int(dst_sorted.numel())
    nz, nz_ptr, empty = _nonempty(dst_sorted, num_dst)
    item_ptr = _items(nz_ptr, E, item)
    G = item_ptr.numel() - 1
    item_row = torch.searchsorted(nz_ptr, item_ptr[:-1], right=True) - 1 if E > 0 else item_ptr[:0]
    split = _split_units(nz_ptr, item_ptr)
    g = torch.arange(G, dtype=torch.int64, device=dst_sorted.device)
    if E > 0:
        head = nz_ptr[item_row] < item_ptr[:-1]
        rt = torch.searchsorted(nz_ptr, item_ptr[1:] - 1, right=True) - 1
        tail = (nz_ptr[rt + 1] > item_ptr[1:]) & (nz_ptr[rt] >= item_ptr[:-1])
        pslot = torch.cat([2 * g[head], 2 * g[tail] + 1])
        pslot_row = torch.cat([nz[item_row[head]], nz[rt[tail]]])
    else:
        pslot = pslot_row = g[:0]
    nz_word = nz.clone()
    nz_word[split] += _SPLIT
    # compact (relation, destination) rows cut by an item boundary (inside split rows only): the second walk of
    # the two-pass HGT backward pre-zeroes them and reduces their pieces
    if E > 0:
        change = torch.ones(E, dtype=torch.bool, device=dst_sorted.device)
        change[1:] = idx_ucol[1:] != idx_ucol[:-1]
        run_start = torch.nonzero(change, as_tuple=False).view(-1)
        run_ptr = torch.cat([run_start, torch.tensor([E], dtype=torch.int64, device=dst_sorted.device)])
        split_runs = idx_ucol[run_start][_split_units(run_ptr, item_ptr)]
    else:
        split_runs = item_ptr[:0]
    return dict(num_positions=E, item_ptr=item_ptr, item_row=item_row, nz_rows=nz_word, nz_ptr=nz_ptr,
                fix_rows=nz[split], empty_rows=empty, pslot=pslot, pslot_row=pslot_row, fix_runs=split_runs,
                idx_a=idx_src, idx_b=idx_urow, idx_c=idx_ucol, idx_d=None, row_rel=ucol_rel)


def backward_arrays(src_sorted, num_src, idx_dst, idx_urow, idx_ucol, idx_fwdpos, urow_rel, item=ITEM):
    """Arrays of a backward plan. ``src_sorted`` [E] non-decreasing source of every position; the
    positions of one source are sorted by relation, i.e. grouped by ``idx_urow`` (the compact
    (rel, src) row = the run); ``idx_fwdpos`` [E] forward-plan position of the same edge."""
    E = int(src_sorted.numel())
    dev = src_sorted.device
    nz, nz_ptr, empty = _nonempty(src_sorted, num_src)
    if E > 0:
        change = torch.ones(E, dtype=torch.bool, device=dev)
        change[1:] = idx_urow[1:] != idx_urow[:-1]
        run_start = torch.nonzero(change, as_tuple=False).view(-1)
        run_ptr = torch.cat([run_start, torch.tensor([E], dtype=torch.int64, device=dev)])
        run_id = idx_urow[run_start]
    else:
        run_ptr = torch.zeros(1, dtype=torch.int64, device=dev)
        run_id = run_ptr[:0]
    item_ptr = _items(run_ptr, E, item)
    item_row = torch.searchsorted(nz_ptr, item_ptr[:-1], right=True) - 1 if E > 0 else item_ptr[:0]
    split_rows = _split_units(nz_ptr, item_ptr)
    split_runs = run_id[_split_units(run_ptr, item_ptr)]
    nz_word = nz.clone()
    nz_word[split_rows] += _SPLIT
    rel_word = urow_rel.clone().to(torch.int64)
    rel_word[split_runs] += _SPLIT
    return dict(num_positions=E, item_ptr=item_ptr, item_row=item_row, nz_rows=nz_word, nz_ptr=nz_ptr,
                fix_rows=torch.cat([empty, nz[split_rows]]), empty_rows=None, pslot=None, pslot_row=None,
                fix_runs=split_runs, idx_a=idx_dst, idx_b=idx_urow, idx_c=idx_ucol, idx_d=idx_fwdpos,
                row_rel=rel_word)


_ORDER = ["item_ptr", "item_row", "nz_rows", "nz_ptr", "fix_rows", "empty_rows", "pslot", "pslot_row", "fix_runs",
          "idx_a", "idx_b", "idx_c", "idx_d", "row_rel"]


@dataclass
class WalkPlan:
    """``blob``: every array as int32, back to back (16-byte aligned); ``meta``: the counts and the
    element offset of every array (-1 = absent), in hetb_walk_plan's field order — the two arguments
    the plan ops take (csrc/torch_bindings.cpp: unpack_plan)."""
    blob: torch.Tensor
    meta: list
    num_items: int
    num_positions: int
    arrays: dict = field(default_factory=dict, repr=False)

    @property
    def index_bytes_per_position(self) -> float:
        return self.blob.numel() * 4 / max(1, self.num_positions)


def pack(arrays: dict, device, keep_arrays=False) -> WalkPlan:
    lim = (1 << 31) - 1
    parts, offs, pos = [], {}, 0

    for name in _ORDER:
        t = arrays.get(name)
        if t is None:
            offs[name] = -1
            continue
        t = t.to(torch.int64)
        if t.numel() and name not in ("nz_rows", "row_rel"):
            assert int(t.max()) <= lim and int(t.min()) >= 0, f"plan array {name} does not fit int32"
        offs[name] = pos
        n = t.numel()
        pad = (-n) % 4
        parts.append(t.to(torch.int32))
        if pad:
            parts.append(torch.zeros(pad, dtype=torch.int32, device=t.device))
        pos += n + pad
    if not parts:
        parts = [torch.zeros(4, dtype=torch.int32)]
    blob = torch.cat([p.to(device) for p in parts]).contiguous()

    def cnt(name):
        t = arrays.get(name)
        return 0 if t is None else int(t.numel())

    G = max(0, cnt("item_ptr") - 1)
    meta = [int(arrays["num_positions"]), G, cnt("nz_rows") , cnt("fix_rows"), cnt("empty_rows"), cnt("pslot"),
            cnt("fix_runs")] + [offs[n] for n in _ORDER]
    assert len(meta) == META_LEN
    return WalkPlan(blob, meta, G, int(arrays["num_positions"]), arrays if keep_arrays else {})


# ---------------------------------------------------------------------------
# On-disk plan cache: a plan depends only on the graph, costs two radix sorts and a
# dozen tensor passes to build (0.4 s at mag size) and is 0.6 GB there; a trainer that is restarted on the
# same graph can load it instead. Opt-in: HetGraph.get_edge_plan() uses it when the environment variable
# HETB_PLAN_CACHE names a directory.
# ---------------------------------------------------------------------------
PLAN_FORMAT = 3  # bump when forward_arrays / backward_arrays / pack change what they emit


def graph_key(num_dst, num_src, num_rels, *index_tensors, item=ITEM) -> str:
    """sha1 over the plan format, the item size, the graph's sizes and the bytes of its index arrays (the
    relation-sorted separate COO: rel_ptrs, row_indices, col_indices, eids)."""
    import hashlib
    h = hashlib.sha1()
    h.update(repr((PLAN_FORMAT, int(item), int(num_dst), int(num_src), int(num_rels))).encode())
    for t in index_tensors:
        a = t.detach().to("cpu", torch.int64).contiguous().numpy()
        h.update(repr(a.shape).encode())
        h.update(a.tobytes())
    return h.hexdigest()


def save_plans(path, key, fwd: WalkPlan, bwd: WalkPlan, fwd_eids) -> None:
    """Atomic write (temp file + rename): a killed process never leaves a truncated cache entry."""
    import os
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save({"format": PLAN_FORMAT, "key": key,
                "fwd": (fwd.blob.cpu(), list(fwd.meta), fwd.num_items, fwd.num_positions),
                "bwd": (bwd.blob.cpu(), list(bwd.meta), bwd.num_items, bwd.num_positions),
                "fwd_eids": fwd_eids.cpu()}, tmp)
    os.replace(tmp, path)


def load_plans(path, key, device):
    """(fwd, bwd, fwd_eids) or None when the file is missing, of another format or of another graph."""
    import os
    if not os.path.exists(path):
        return None
    try:
        d = torch.load(path, map_location="cpu")
    except Exception:
        return None
    if d.get("format") != PLAN_FORMAT or d.get("key") != key:
        return None

    def plan(t):
        blob, meta, items, pos = t
        if len(meta) != META_LEN or blob.dtype != torch.int32:
            return None
        return WalkPlan(blob.to(device).contiguous(), [int(x) for x in meta], int(items), int(pos))

    fwd, bwd = plan(d["fwd"]), plan(d["bwd"])
    if fwd is None or bwd is None:
        return None
    return fwd, bwd, d["fwd_eids"].to(device)
