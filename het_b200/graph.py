"""Device-resident heterograph container with the reference's accessor names.

Host-side mirror of ``MyDGLGraph`` (hrt/python/utils/mydgl_graph.py:293-420): the
models and backend functions ask it for dicts of int64 index tensors
(``get_separate_coo_original``, ``get_in_csr``, ``get_out_csr``,
``get_original_node_type_offsets`` ...). Unlike the reference, every format is
built ON THE GPU by the hand-written stable radix sort behind
``K.hetb200_stable_sort_by_key`` (csrc/convert.cu) instead of CPU
vector-of-vector bucketing (MyHyb/MyHyb.h:1099-1150) and Python dict remaps
(mydgl_graph.py:765-824).

Orderings (all bit-exact against oracle/converters_ref.py):
  separate COO : edges sorted by (relation, referential eid)        [a16]
  out-CSR      : stable by source;  in-CSR : stable by destination  [a17]
  eids         : canonicalised so separate-COO eids are 0..E-1      [a18]
"""
from __future__ import annotations

import torch

from .graphs import HeteroCOO
from .kernels import K


class HetGraph:
    """``num_nodes`` rows index destinations (in-CSR rows, op outputs); sources index
    ``num_src_nodes`` rows (out-CSR rows, gathered inputs). On one GPU both are N. A
    destination-range shard (het_b200/parallel.py) has local destination ids and sources
    that index the all-gathered feature matrix."""

    def __init__(self, coo: HeteroCOO, device="cuda"):
        self._build(coo.name, torch.from_numpy(coo.src), torch.from_numpy(coo.dst),
                    torch.from_numpy(coo.etype), coo.num_nodes, coo.num_nodes, coo.num_rels,
                    torch.from_numpy(coo.node_type_offsets), device)
        dev = torch.device(device)
        self.rel_src_type = torch.from_numpy(coo.rel_src_type).to(dev)
        self.rel_dst_type = torch.from_numpy(coo.rel_dst_type).to(dev)

    @classmethod
    def from_arrays(cls, name, src, dst, rel, num_src_nodes, num_dst_nodes, num_rels,
                    node_type_offsets, device="cuda"):
        g = cls.__new__(cls)
        g._build(name, src, dst, rel, num_src_nodes, num_dst_nodes, num_rels, node_type_offsets, device)
        return g

    def _build(self, name, src, dst, rel, num_src_nodes, num_dst_nodes, num_rels, node_type_offsets, device):
        dev = torch.device(device)
        self.name = name
        self.num_nodes = int(num_dst_nodes)
        self.num_src_nodes = int(num_src_nodes)
        self.num_rels = int(num_rels)
        self.num_edges = int(src.numel())
        src, dst, rel = src.to(dev), dst.to(dev), rel.to(dev)
        eid = torch.arange(self.num_edges, dtype=torch.int64, device=dev)
        self.node_type_offsets = node_type_offsets.to(dev)

        # relation sort: input is in eid order, so a stable sort by relation
        # yields (relation, eid) order — the net effect of ToSeparateCOO +
        # sort_coo_by_etype_eids (python/utils/coo_sorters.py:156-170).
        row, col, old_eid, rel_ptrs = K.hetb200_stable_sort_by_key(rel, self.num_rels, [src, dst, eid])
        # canonicalize_eids("separate_coo"): separate-COO position becomes the edge id
        remap = torch.empty_like(old_eid)
        remap[old_eid] = eid
        self.separate_coo = {
            "rel_ptrs": rel_ptrs, "row_indices": row, "col_indices": col, "eids": eid,
        }
        new_eid = remap  # edge id of original edge i
        self.orig_to_canonical_eid = remap
        o_col, o_rel, o_eid, o_ptr = K.hetb200_stable_sort_by_key(src, self.num_src_nodes, [dst, rel, new_eid])
        self.out_csr = {"row_ptrs": o_ptr, "col_indices": o_col, "rel_types": o_rel, "eids": o_eid}
        i_col, i_rel, i_eid, i_ptr = K.hetb200_stable_sort_by_key(dst, self.num_nodes, [src, rel, new_eid])
        self.in_csr = {"row_ptrs": i_ptr, "col_indices": i_col, "rel_types": i_rel, "eids": i_eid}
        self._empty = torch.empty(0, dtype=torch.int64, device=dev)

    # ---- accessor names of the reference's MyDGLGraph -------------------
    def get_separate_coo_original(self):
        return self.separate_coo

    def get_in_csr(self):
        return self.in_csr

    def get_out_csr(self):
        return self.out_csr

    def get_original_node_type_offsets(self):
        return self.node_type_offsets

    def get_relation_node_types(self):
        """(source node type, destination node type) of every relation, or None when the graph
        was built from bare arrays (canonical edge types, as DGL's heterograph keeps them)."""
        if getattr(self, "rel_src_type", None) is None:
            return None
        return self.rel_src_type, self.rel_dst_type

    def get_separate_unique_node_indices(self):
        """Double-sided compact-as-of-node lists (reference:
        generate_separate_unique_node_indices_for_each_etype,
        hrt/python/utils_lite/mydgl_graph_methods.py:72-157): per relation the sorted unique nodes
        that appear as source OR destination, back to back; the binary-search compact kinds
        (CompactAsOfNodeKind::Enabled) locate the row of a (relation, node) pair in them. Needs one
        id space for sources and destinations (not built on a destination shard)."""
        if getattr(self, "_unique_double_sided", None) is None:
            assert self.num_src_nodes == self.num_nodes, "double-sided lists need one node id space"
            coo = self.separate_coo
            dev = coo["rel_ptrs"].device
            counts = coo["rel_ptrs"][1:] - coo["rel_ptrs"][:-1]
            rel_of_edge = torch.repeat_interleave(torch.arange(self.num_rels, device=dev), counts)
            n = self.num_nodes
            uniq = torch.unique(torch.cat([rel_of_edge * n + coo["row_indices"], rel_of_edge * n + coo["col_indices"]]),
                                sorted=True)
            bounds = torch.arange(self.num_rels + 1, device=dev, dtype=torch.int64) * n
            self._unique_double_sided = {"rel_ptrs": torch.searchsorted(uniq, bounds).contiguous(),
                                         "node_indices": (uniq % n).contiguous()}
        return self._unique_double_sided

    def get_separate_unique_node_indices_single_sided(self):
        """Compact-as-of-node lists, one side at a time (reference:
        generate_separate_unique_node_indices_single_sided_for_each_etype,
        hrt/python/utils_lite/mydgl_graph_methods.py:11-157 — per relation
        ``torch.unique(row or col, return_inverse=True)``, concatenated with the running
        offsets). Built on the GPU from the relation-sorted COO in one pass per side:
        the key (relation, node) is already grouped by relation, so a sorted unique of
        ``rel * N + node`` yields the per-relation sorted node lists back to back.

        ``rel_ptrs_<side>``      [R+1]  segment of relation r inside the compact list
        ``node_indices_<side>``  [U]    node id of every unique (relation, node) pair
        ``inverse_idx_<side>``   [E]    compact row of edge id e ("edata_idx_to_inverse_idx";
                                         edge ids are separate-COO positions after canonicalisation)
        ``incsr_inverse_idx_<side>``, ``outcsr_inverse_idx_<side>``  [E]  the same per CSR position
        with side ``row`` = sources, ``col`` = destinations."""
        if getattr(self, "_unique_single_sided", None) is None:
            coo = self.separate_coo
            dev = coo["rel_ptrs"].device
            counts = coo["rel_ptrs"][1:] - coo["rel_ptrs"][:-1]
            rel_of_edge = torch.repeat_interleave(torch.arange(self.num_rels, device=dev), counts)
            out = {}
            for side, nodes, n in (("row", coo["row_indices"], self.num_src_nodes),
                                   ("col", coo["col_indices"], self.num_nodes)):
                key = rel_of_edge * n + nodes
                uniq, inverse = torch.unique(key, sorted=True, return_inverse=True)
                bounds = torch.arange(self.num_rels + 1, device=dev, dtype=torch.int64) * n
                out["rel_ptrs_" + side] = torch.searchsorted(uniq, bounds).contiguous()
                out["node_indices_" + side] = (uniq % n).contiguous()
                # eids are 0..E-1 in separate-COO order, so position == edge id
                out["inverse_idx_" + side] = inverse.contiguous()
                # the same map gathered into in-CSR / out-CSR position order (aligned with the
                # eids array of that format): what the in-CSR / out-CSR edge kernels read, one
                # dependent load level less than inverse_idx[eids[p]]
                out["incsr_inverse_idx_" + side] = inverse[self.in_csr["eids"]].contiguous()
                out["outcsr_inverse_idx_" + side] = inverse[self.out_csr["eids"]].contiguous()
            # the edges grouped by compact source row (a CSR over the (relation, source) rows): the
            # RGAT compact backward walks this order so a compact row's gradient is a register sum
            inv_row = out["inverse_idx_row"]
            order = torch.argsort(inv_row, stable=True)
            counts = torch.bincount(inv_row, minlength=int(out["node_indices_row"].numel()))
            ptrs = torch.zeros(counts.numel() + 1, dtype=torch.int64, device=inv_row.device)
            torch.cumsum(counts, 0, out=ptrs[1:])
            out["urow_ptrs"] = ptrs
            out["urow_eids"] = coo["eids"][order].contiguous()
            out["urow_dst"] = coo["col_indices"][order].contiguous()
            out["urow_inverse_idx_col"] = out["inverse_idx_col"][order].contiguous()
            self._unique_single_sided = out
        return self._unique_single_sided

    def get_edge_plan(self):
        """Walk plans of the plan-based fused edge passes (het_b200/plan.py; csrc/edge_plan.cu):
        ``fwd`` walks the positions sorted by (destination, relation, edge id), ``bwd`` by (source,
        relation, edge id). Both orders are stable sorts of the relation-sorted separate COO (whose
        position is the canonical edge id) by one endpoint, done with the library's radix sort; the
        compact (relation, node) row of every position rides along as a payload."""
        if getattr(self, "_edge_plan", None) is None:
            import os
            from . import plan as P
            coo = self.separate_coo
            dev = coo["rel_ptrs"].device
            cache_dir, cache_path, key = os.environ.get("HETB_PLAN_CACHE"), None, None
            if cache_dir:  # on-disk plan cache (opt-in): keyed by the graph's own index arrays
                key = P.graph_key(self.num_nodes, self.num_src_nodes, self.num_rels, coo["rel_ptrs"], coo["row_indices"],
                                  coo["col_indices"], coo["eids"])
                os.makedirs(cache_dir, exist_ok=True)
                cache_path = os.path.join(cache_dir, f"hetb_plan_{key}.pt")
                hit = P.load_plans(cache_path, key, dev)
                if hit is not None:
                    self._edge_plan = {"fwd": hit[0], "bwd": hit[1], "fwd_eids": hit[2], "scratch": {}, "from_cache": True}
                    return self._edge_plan
            u = self.get_separate_unique_node_indices_single_sided()
            E = self.num_edges
            assert max(E, self.num_nodes, self.num_src_nodes) < (1 << 31) - 1, "plan indices are int32"
            eid = coo["eids"]
            rels = torch.arange(self.num_rels, device=dev)
            ucol_rel = torch.repeat_interleave(rels, u["rel_ptrs_col"][1:] - u["rel_ptrs_col"][:-1])
            urow_rel = torch.repeat_interleave(rels, u["rel_ptrs_row"][1:] - u["rel_ptrs_row"][:-1])
            f_src, f_urow, f_ucol, f_eid, f_ptr = K.hetb200_stable_sort_by_key(
                coo["col_indices"], self.num_nodes, [coo["row_indices"], u["inverse_idx_row"], u["inverse_idx_col"], eid])
            dst_sorted = torch.repeat_interleave(torch.arange(self.num_nodes, device=dev), f_ptr[1:] - f_ptr[:-1])
            fwd_pos = torch.empty_like(eid)
            fwd_pos[f_eid] = torch.arange(E, device=dev)
            fwd = P.pack(P.forward_arrays(dst_sorted, self.num_nodes, f_src, f_urow, f_ucol, ucol_rel), dev)
            b_dst, b_urow, b_ucol, b_eid, b_ptr = K.hetb200_stable_sort_by_key(
                coo["row_indices"], self.num_src_nodes, [coo["col_indices"], u["inverse_idx_row"], u["inverse_idx_col"], eid])
            src_sorted = torch.repeat_interleave(torch.arange(self.num_src_nodes, device=dev), b_ptr[1:] - b_ptr[:-1])
            bwd = P.pack(P.backward_arrays(src_sorted, self.num_src_nodes, b_dst, b_urow, b_ucol, fwd_pos[b_eid],
                                           urow_rel), dev)
            self._edge_plan = {"fwd": fwd, "bwd": bwd, "fwd_eids": f_eid, "scratch": {}, "from_cache": False}
            if cache_path:
                P.save_plans(cache_path, key, fwd, bwd, f_eid)
        return self._edge_plan

    def plan_scratch(self, H, D):
        """Persistent scratch of the forward walks (partial slots of split rows, their sums) and the
        backward pre-pass, shared by every layer / call on this graph (stream-ordered reuse)."""
        pl = self.get_edge_plan()
        key = (int(H), int(D))
        if key not in pl["scratch"]:
            dev = self.separate_coo["rel_ptrs"].device
            G = max(1, pl["fwd"].num_items)
            f32 = dict(dtype=torch.float32, device=dev)
            pl["scratch"][key] = {
                "part_m": torch.empty(2 * G * H, **f32), "part_s": torch.empty(2 * G * H, **f32),
                "part_acc": torch.empty(2 * G * D, **f32), "ssum": torch.empty(self.num_nodes * H, **f32),
                "gdot": torch.empty(2 * self.num_nodes * H, **f32),  # (lse, <gout, out>) pairs per destination
            }
        return pl["scratch"][key]

    def get_num_nodes(self):
        return self.num_nodes

    def get_num_edges(self):
        return self.num_edges

    def get_num_rels(self):
        return self.num_rels

    def get_num_ntypes(self):
        return int(self.node_type_offsets.numel() - 1)

    def index_bytes(self) -> int:
        tot = 0
        for d in (self.separate_coo, self.in_csr, self.out_csr):
            tot += sum(t.numel() * 8 for t in d.values())
        return tot
