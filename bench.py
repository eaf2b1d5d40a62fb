#!/usr/bin/env python
"""Benchmark of the HET/Hector hot path on B200 — one JSON line on stdout (rank 0).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl hetb200|reference]
                    [--workload rgat_bgs|hgt_mag|rgat_wikikg2|rgat_aifb] [--no-graph]

A "step" is one full-graph training epoch (forward, loss, backward, Adam) exactly as the
reference times it (hrt/python/RGNNUtils/RGNNUtils.py:291-345). Default workload at N=1 is
BASELINE.json configs[1]: RGAT 1 layer, hidden 64, 1 head, on the synthetic BGS-shaped
heterograph. For N>1 the graph grows with N (N x BGS-shaped, "scaling": "weak") and is
sharded by destination range with an NCCL all-gather of node features per layer.

`value`  : edges/s with everything resident in HBM (CUDA events, max over ranks).
`e2e`    : the same metric through het_b200.train.FullGraphTrainer.run_from_host — node
           features and labels copied from pinned host memory every step, loss read back every
           step; the copy of step i+1 overlaps the kernels of step i (serial_ms_per_step is the
           same thing one step_from_host call at a time).
`roofline`: the kernel with the largest share of the step, timed alone inside this run.
`cpu_baseline`: the repo's DGL baseline model restated in plain torch (oracle/models_ref.py),
           timed on this box's host cores on a bounded sample.
`--impl reference`: that CPU baseline as its own arm (the reference's CUDA build and DGL are
           not installable here — DESIGN.md).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (model, graph shape, heads, layers, BASELINE.json config it corresponds to)
    "rgcn_aifb": ("rgcn", "aifb", 1, 2, "configs[0], with backward + Adam"),
    "rgat_aifb": ("rgat", "aifb", 1, 1, "configs[0] graph, RGAT"),
    "rgat_bgs": ("rgat", "bgs", 1, 1, "configs[1]"),
    "hgt_mag": ("hgt", "mag", 4, 2, "configs[2]"),
    "rgat_wikikg2": ("rgat", "wikikg2", 1, 1, "configs[3]"),
}
METRIC = "rgnn_fullgraph_fwd_bwd_edges_per_s"
UNIT = "edges/s"


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="hetb200", choices=["hetb200", "reference"])
    ap.add_argument("--workload", default="hgt_mag", choices=sorted(WORKLOADS),
                    help="default hgt_mag = BASELINE.json configs[2] on one GPU and configs[4] (the same graph, "
                         "sharded by destination range: strong scaling) on N > 1")
    ap.add_argument("--scale", type=float, default=None,
                    help="graph scale; default 1.0 (the named config, strong scaling over --gpus); "
                         "--scale N with --gpus N grows the graph with the GPU count (weak scaling)")
    ap.add_argument("--no-graph", action="store_true", help="do not capture the step in a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary lines (rgat_bgs, bf16 table storage)")
    ap.add_argument("--no-compact", dest="compact", action="store_false",
                    help="turn off the reference's compact_as_of_node_flag (default on: X.W_r is computed once "
                         "per unique (relation, node) pair instead of once per edge; same outputs)")
    ap.add_argument("--bf16-tables", action="store_true",
                    help="HGT: store the gathered tables (k, W_att.q, W_msg.v) as bf16 (fp32 arithmetic, outputs and "
                         "gradients; parity bar 2e-2). Off by default: the headline is the fp32 path.")
    ap.add_argument("--no-fused", dest="fused", action="store_false",
                    help="HGT: run inner product, edge softmax and aggregation as three ops instead of one pass")
    return ap.parse_args(argv)


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------
# CPU arm: the repo's DGL baseline models, restated (oracle/models_ref.py)
# ---------------------------------------------------------------------------
def cpu_dgl_baseline(wl, coo, steps, warmup, budget_s):
    """Times fwd + loss + bwd + Adam of the DGL-baseline restatement in fp32 on all host
    cores. Returns (edges_per_s, ms_per_step, cores, sample description, steps done)."""
    import torch
    import torch.nn.functional as Fn
    from oracle import models_ref as M

    kind, shape, heads, layers, _ = WORKLOADS[wl]
    note = ""
    # HeteroGraphConv keeps one projected copy of the source node type per relation for the backward:
    # R x N x 256 B. The wikikg2 shape (535 relations over one node type of 2.5 M nodes) would need 340 GB
    # of host memory, so the CPU arm runs the same shape at a bounded scale (the sample is stated).
    need = coo.num_rels * coo.num_nodes * 256
    if need > 24e9:
        from het_b200 import graphs
        full_e = coo.num_edges
        scale = 1.0
        while need * scale > 8e9:
            scale /= 2
        coo = graphs.SHAPES[shape](scale=scale * (coo.num_nodes / graphs.SHAPES[shape](scale=1 / 1024).num_nodes / 1024))
        note = f" [bounded sample: the {shape} shape scaled to E={coo.num_edges} of {full_e}: the full graph needs {need / 1e9:.0f} GB of host memory in this formulation]"
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n, R, D = coo.num_nodes, coo.num_rels, 64
    src, dst, et = (torch.from_numpy(a) for a in (coo.src, coo.dst, coo.etype))
    offs = coo.node_type_offsets
    ranges = [(int(offs[s]), int(offs[s + 1]), int(offs[d]), int(offs[d + 1]))
              for s, d in zip(coo.rel_src_type, coo.rel_dst_type)]
    g = torch.Generator().manual_seed(0)
    x = torch.nn.Parameter(torch.rand(n, D, generator=g) * 2 - 1)
    labels = torch.randint(0, D, (n,), generator=g)
    F = D // heads

    def P(*shape):
        return torch.nn.Parameter(torch.randn(*shape, generator=g) * 0.1)

    if kind == "rgcn":
        ps = [dict(w=P(R, D, D), b=P(D)) for _ in range(layers)]
        norm = 1.0 / torch.bincount(dst, minlength=n).clamp(min=1).to(torch.float32)[dst]

        def fwd():
            h = x
            for i, p in enumerate(ps):
                h = M.dgl_rgcn_layer(src, dst, et, h, p["w"], norm) + p["b"]
                if i + 1 < len(ps):
                    h = torch.relu(h)
            return h
    elif kind == "rgat":
        ps = [dict(fc=P(R, D, D), al=P(R, heads, F), ar=P(R, heads, F)) for _ in range(layers)]

        def fwd():
            h = x
            for p in ps:
                h = M.dgl_rgat_layer(src, dst, et, h, p["fc"], p["al"], p["ar"], 0.2, R, ranges)
            return h
    else:
        T = coo.num_ntypes
        ps = [dict(k=P(T, 1, D, D), q=P(T, 1, D, D), v=P(T, 1, D, D), a=P(T, 1, D, D), pri=P(R, heads),
                   att=P(R, heads, F, F), msg=P(R, heads, F, F), skip=P(T, 1, 1, 1)) for _ in range(layers)]
        offs_t = torch.from_numpy(offs)

        def fwd():
            h = x
            for p in ps:
                h = M.dgl_hgt_layer(src, dst, et, offs_t, coo.rel_dst_type, h, p["k"], p["q"], p["v"], p["a"],
                                    p["pri"], p["att"], p["msg"], p["skip"], R, ranges)
            return h
    params = [x] + [t for p in ps for t in p.values()]
    opt = torch.optim.Adam(params, lr=1e-2)

    def step():
        opt.zero_grad(set_to_none=True)
        loss = Fn.nll_loss(Fn.log_softmax(fwd(), dim=-1), labels)
        loss.backward()
        opt.step()
        return float(loss.detach())

    t_begin = time.perf_counter()
    for _ in range(warmup):
        step()
        if time.perf_counter() - t_begin > budget_s / 2:
            break
    done, t0 = 0, time.perf_counter()
    while done < steps:
        step()
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = (time.perf_counter() - t0) / done
    sample = (f"{done} full-graph step(s) of the DGL-baseline {kind.upper()} restatement (fp32, {layers} layer(s), "
              f"{heads} head(s)) on {coo.name}: N={n} R={R} E={coo.num_edges}" + note)
    return coo.num_edges * layers / dt, dt * 1e3, cores, sample, done


# ---------------------------------------------------------------------------
# Per-kernel roofline legs (timed alone, L2 flushed between launches)
# ---------------------------------------------------------------------------
def _time_kernel(torch, fn, flush, reps=10, pre=None):
    """Average device time (ms) of fn() over `reps` launches, each preceded by an L2 flush
    (a 256 MB write, larger than the 126 MB L2). CUDA events on torch's current stream,
    which is the stream the torch_hrt ops launch on. `pre` (untimed) restores an input fn() consumes."""
    for _ in range(3):
        if pre:
            pre()
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(reps):
        if pre:
            pre()
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / reps


def kernel_rooflines(trainer, peak_gbs, peak_src):
    """Times the hot kernels of one RGAT / HGT layer in isolation on the trainer's graph and
    returns roofline dicts. `achieved` = algorithmic (compulsory) bytes / time, with the
    byte formulas of SURVEY.md §8(d) / DESIGN.md §4 (s = 4 B fp32, i = 8 B int64 indices)."""
    import torch
    from het_b200.kernels import K
    g, dev = trainer.g, trainer.device
    E, N, Ns, R = g.num_edges, g.num_nodes, g.num_src_nodes, g.num_rels
    coo, ic, oc = g.get_separate_coo_original(), g.get_in_csr(), g.get_out_csr()
    layer = trainer.model.layers[0]
    H = layer.num_heads
    D = 64
    F = D // H
    flush = torch.zeros(64 * 1024 * 1024, device=dev)
    uniq_src = int(torch.unique(coo["row_indices"]).numel())
    out = []

    def leg(name, fn, nbytes, note):
        ms = _time_kernel(torch, fn, flush)
        gbs = nbytes / (ms * 1e-3) / 1e9
        out.append({"kernel": name, "bound": "hbm", "achieved": round(gbs, 1), "peak": peak_gbs, "unit": "GB/s",
                    "frac": round(gbs / peak_gbs, 4), "peak_source": peak_src, "ms": round(ms, 4),
                    "algorithmic_bytes": int(nbytes), "traffic": None, "note": note})

    x = torch.rand(Ns, D, device=dev) * 2 - 1
    eidx = torch.empty(0, dtype=torch.int64, device=dev)
    if trainer.kind == "rgat":
        W = layer.conv_weights.detach().contiguous()
        WT = W.transpose(2, 3).contiguous()
        feat = torch.empty(E, H, F, device=dev)
        d_row = {"separate_coo_rel_ptrs": coo["rel_ptrs"], "separate_coo_node_indices": coo["row_indices"],
                 "separate_coo_eids": coo["eids"]}
        leg("segment_mm_fwd(gather src, X.W_r)", lambda: K.rgnn_relational_matmul(d_row, 0, W, x, feat, True),
            uniq_src * D * 4 + E * D * 4 + R * D * D * 4 + E * 8,
            "M_unique_in*Din*s + M*Dout*s + G*Din*Dout*s + M*i")
        el, er = torch.randn(E, H, 1, device=dev), torch.randn(E, H, 1, device=dev)
        ssum, ex, ret = torch.empty(N, H, device=dev), torch.empty(E, H, 1, device=dev), torch.empty(N, H, F, device=dev)
        leg("rgat_edge_softmax_aggregate_fwd(in-CSR)",
            lambda: K.relational_fused_gat_csr(ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], eidx,
                                               eidx, feat, el, er, ssum, ex, ret, 0.2),
            E * D * 4 + 2 * E * H * 4 + E * 8 + (N + 1) * 8 + E * H * 4 + N * H * 4 + N * D * 4,
            "feat + el,er + eids + row_ptr ; exp + sum + ret")
        gout = torch.randn(N, H, F, device=dev)
        gf, gel, ger = torch.empty_like(feat), torch.empty_like(el), torch.empty_like(er)
        leg("rgat_edge_bwd(out-CSR)",
            lambda: K.backward_relational_fused_gat_csr(oc["row_ptrs"], oc["col_indices"], oc["eids"], oc["rel_types"],
                                                        eidx, eidx, feat, el, er, ssum, ex, ret, gout, gf, gel, ger, 0.2),
            E * D * 4 + 3 * E * H * 4 + N * H * 4 + 2 * N * D * 4 + 2 * E * 8 + E * D * 4 + 2 * E * H * 4,
            "feat,el,er,exp,sum,ret,gout,idx ; grad_feat,grad_el,grad_er")
        gx, gw = torch.zeros_like(x), torch.zeros_like(W)
        leg("segment_mm_bwd(dX scatter + dW)",
            lambda: K.backward_rgnn_relational_matmul(d_row, 0, WT, x, gf, gx, gw, True),
            E * D * 4 + 2 * uniq_src * D * 4 + 2 * R * D * D * 4 + E * 8,
            "dY + X_unique + W ; dX_unique + dW ; gather list")
    else:
        Watt = layer.relation_att.detach().contiguous()
        q = torch.randn(N, H, F, device=dev)
        k = torch.randn(Ns, H, F, device=dev)
        wq = torch.empty(E, H, F, device=dev)
        d_col = {"separate_coo_rel_ptrs": coo["rel_ptrs"], "separate_coo_node_indices": coo["col_indices"],
                 "separate_coo_eids": coo["eids"]}
        uniq_dst = int(torch.unique(coo["col_indices"]).numel())
        leg("segment_mm_fwd(gather dst, q.W_att)", lambda: K.rgnn_relational_matmul(d_col, 0, Watt, q, wq, False),
            uniq_dst * D * 4 + E * D * 4 + R * H * F * F * 4 + E * 8, "M_unique_in*D*s + M*D*s + W + M*i")
        score = torch.empty(E, H, device=dev)
        leg("inner_product_fwd", lambda: K.rgnn_inner_product_right_node_separatecoo(
            {}, 0, coo["rel_ptrs"], coo["eids"], coo["row_indices"], coo["col_indices"], wq, k, score),
            E * D * 4 + uniq_src * D * 4 + 2 * E * 8 + E * H * 4, "left + right_unique + idx ; score")
        mu = torch.ones(R, H, device=dev) / 4
        ssum, t, norm = torch.empty(N, H, device=dev), torch.empty(E, H, device=dev), torch.empty(E, H, device=dev)
        leg("hgt_edge_softmax_fwd(in-CSR)", lambda: K.hgt_full_graph_edge_softmax_ops_csr(
            ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], score, mu, ssum, t, norm),
            E * H * 4 + 2 * E * 8 + (N + 1) * 8 + 2 * E * H * 4 + N * H * 4, "a + eid,rel + row_ptr ; t + norm + sum")
        new_h = torch.empty(N, H, F, device=dev)
        leg("hgt_aggregate_fwd(in-CSR)", lambda: K.hgt_full_graph_message_mean_aggregation_csr(
            ic["row_ptrs"], ic["col_indices"], ic["rel_types"], ic["eids"], wq, t, ssum, mu, new_h),
            E * D * 4 + E * H * 4 + E * 8 + (N + 1) * 8 + N * H * 4 + N * D * 4, "msg + t + eid + row_ptr + sum ; out")
    del flush
    return out


def kernel_rooflines_compact(trainer, peak_gbs, peak_src):
    """Same as kernel_rooflines for the compact-as-of-node path (the default): the per-edge
    [E, 64] tensors are replaced by one row per unique (relation, node) pair, U_row / U_col rows,
    read by the edge kernels through the direct-indexing maps. Compulsory bytes count every
    compact / node row once (unique-row granularity), every per-edge scalar and index once."""
    import torch
    from het_b200 import backend as B
    from het_b200.kernels import K
    g, dev = trainer.g, trainer.device
    E, N, Ns, R = g.num_edges, g.num_nodes, g.num_src_nodes, g.num_rels
    coo, ic, oc = g.get_separate_coo_original(), g.get_in_csr(), g.get_out_csr()
    u = g.get_separate_unique_node_indices_single_sided()
    Ur, Uc = int(u["node_indices_row"].numel()), int(u["node_indices_col"].numel())
    layer = trainer.model.layers[0]
    H, D = getattr(layer, "num_heads", 1), 64
    F = D // H
    flush = torch.zeros(64 * 1024 * 1024, device=dev)
    uniq_src = int(torch.unique(coo["row_indices"]).numel())
    uniq_dst = int(torch.unique(coo["col_indices"]).numel())
    out = []

    def leg(name, fn, nbytes, note, pre=None):
        ms = _time_kernel(torch, fn, flush, pre=pre)
        gbs = nbytes / (ms * 1e-3) / 1e9
        out.append({"kernel": name, "bound": "hbm", "achieved": round(gbs, 1), "peak": peak_gbs, "unit": "GB/s",
                    "frac": round(gbs / peak_gbs, 4), "peak_source": peak_src, "ms": round(ms, 4),
                    "algorithmic_bytes": int(nbytes), "traffic": None, "note": note})

    x = torch.rand(Ns, D, device=dev) * 2 - 1
    d_row = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"], "unique_srcs_and_dests_node_indices": u["node_indices_row"]}
    d_col = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_col"], "unique_srcs_and_dests_node_indices": u["node_indices_col"]}
    ir, icx = u["inverse_idx_row"], u["inverse_idx_col"]                      # by edge id (COO ops)
    iir, iic = u["incsr_inverse_idx_row"], u["incsr_inverse_idx_col"]        # aligned with the in-CSR
    oir, oic = u["outcsr_inverse_idx_row"], u["outcsr_inverse_idx_col"]      # aligned with the out-CSR
    if trainer.kind == "rgcn":
        W = layer.weight.detach().unsqueeze(1).contiguous()
        WT = W.transpose(2, 3).contiguous()
        feat = torch.empty(Ur, 1, D, device=dev)
        leg("segment_mm_fwd(unique (rel,src) rows, X.W_r)", lambda: K.rgnn_relational_matmul(d_row, 1, W, x, feat, True),
            uniq_src * D * 4 + Ur * D * 4 + R * D * D * 4 + Ur * 8, "X_unique*D*s + U*D*s + R*D*D*s + U*i")
        ret = torch.empty(N, D, device=dev)
        norm = trainer.edge_norm
        leg("rgcn_aggregate_fwd(in-CSR, compact)",
            lambda: K.hetb200_rgcn_node_aggregation_csr_compact(ic["row_ptrs"], ic["eids"], iir, feat, norm, ret),
            Ur * D * 4 + E * 4 + 2 * E * 8 + (N + 1) * 8 + N * D * 4, "feat_c + norm + eid,inv_row + row_ptr ; ret")
        gout = torch.randn(N, D, device=dev)
        gf = torch.zeros_like(feat)
        leg("rgcn_aggregate_bwd(out-CSR, compact)",
            lambda: K.hetb200_backward_hgt_message_aggregation_csr_compact(oc["col_indices"], oc["eids"], oir, norm, gout, gf),
            N * D * 4 + E * 4 + 3 * E * 8 + Ur * D * 4, "gout + norm + dst,eid,inv_row ; grad_feat_c")
        gx, gw = torch.zeros_like(x), torch.zeros_like(W)
        leg("segment_mm_bwd(unique rows: dX scatter + dW)",
            lambda: K.backward_rgnn_relational_matmul(d_row, 1, WT, x, gf, gx, gw, True),
            Ur * D * 4 + 2 * uniq_src * D * 4 + 2 * R * D * D * 4 + Ur * 8, "dY_c + X_unique + W ; dX_unique + dW ; list")
        return out
    if trainer.kind == "rgat":
        W = layer.conv_weights.detach().contiguous()
        WT = W.transpose(2, 3).contiguous()
        feat = torch.empty(Ur, H, F, device=dev)
        leg("segment_mm_fwd(unique (rel,src) rows, X.W_r)", lambda: K.rgnn_relational_matmul(d_row, 1, W, x, feat, True),
            uniq_src * D * 4 + Ur * D * 4 + R * D * D * 4 + Ur * 8, "X_unique*D*s + U*D*s + R*D*D*s + U*i")
        el, er = torch.randn(Ur, H, 1, device=dev), torch.randn(Uc, H, 1, device=dev)
        ssum, ex, ret = torch.empty(N, H, device=dev), torch.empty(E, H, 1, device=dev), torch.empty(N, H, F, device=dev)
        if B.USE_EDGE_PLAN:
            pl, sc = g.get_edge_plan(), g.plan_scratch(H, D)
            el2, er2, lse = el.view(Ur, H).contiguous(), er.view(Uc, H).contiguous(), torch.empty(N, H, device=dev)
            leg("rgat_fused_fwd(plan walk: logits + online softmax + aggregate)",
                lambda: K.hetb200_rgat_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, feat, el2, er2, 0.2, lse, ret,
                                                      sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"]),
                Ur * D * 4 + (Ur + Uc) * H * 4 + 2 * E * 4 + pl["fwd"].blob.numel() * 4 - 3 * E * 4 + N * H * 4 + N * D * 4,
                "feat_c + el_c,er_c + int32 (urow,ucol) per position + items/rows ; lse + ret (nothing per edge)")
            gout = torch.randn(N, H, F, device=dev)
            gf, gel, ger = torch.empty_like(feat), torch.empty_like(el2), torch.zeros_like(er2)
            leg("rgat_fused_bwd(plan walk over compact source rows)",
                lambda: K.hetb200_rgat_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, feat, el2, er2, lse, ret, gout,
                                                      sc["gdot"], gf, gel, ger, 0.2),
                2 * Ur * D * 4 + 2 * (Ur + Uc) * H * 4 + 3 * N * H * 4 + 2 * N * D * 4 + 3 * E * 4
                + (pl["bwd"].blob.numel() * 4 - 4 * E * 4),
                "feat_c,el_c,er_c,lse,gdotr,ret,gout + int32 (dst,urow,ucol) per position ; grad_feat_c,grad_el_c,grad_er_c")
            gf = gf.view(Ur, H, F)
        else:
          leg("rgat_edge_softmax_aggregate_fwd(in-CSR, compact)",
            lambda: K.hetb200_relational_fused_gat_csr_compact(ic["row_ptrs"], ic["eids"], iir, iic, feat, el, er, ssum, ex,
                                                               ret, 0.2),
            Ur * D * 4 + (Ur + Uc) * H * 4 + 3 * E * 8 + (N + 1) * 8 + E * H * 4 + N * H * 4 + N * D * 4,
            "feat_c + el_c,er_c + eid,inv_row,inv_col + row_ptr ; exp + sum + ret")
          gout = torch.randn(N, H, F, device=dev)
          gf, gel, ger = torch.zeros_like(feat), torch.zeros_like(el), torch.zeros_like(er)
          leg("rgat_edge_bwd(compact-row walker)",
            lambda: K.hetb200_backward_relational_fused_gat_compact_rows(
                u["urow_ptrs"], u["urow_dst"], u["urow_eids"], u["urow_inverse_idx_col"], feat, el, er, ssum, ex, ret,
                gout, gf, gel, ger, 0.2),
            2 * Ur * D * 4 + 2 * (Ur + Uc) * H * 4 + E * H * 4 + 3 * N * H * 4 + 2 * N * D * 4 + 3 * E * 8 + (Ur + 1) * 8,
            "feat_c,el_c,er_c,exp,sum,ret,gout,gdotr,dst,eid,er_idx,urow_ptr ; grad_feat_c,grad_el_c,grad_er_c,gdotr")
        gx, gw = torch.zeros_like(x), torch.zeros_like(W)
        leg("segment_mm_bwd(unique rows: dX scatter + dW)",
            lambda: K.backward_rgnn_relational_matmul(d_row, 1, WT, x, gf, gx, gw, True),
            Ur * D * 4 + 2 * uniq_src * D * 4 + 2 * R * D * D * 4 + Ur * 8, "dY_c + X_unique + W ; dX_unique + dW ; list")
    else:
        Watt = layer.relation_att.detach().contiguous()
        q = torch.randn(N, H, F, device=dev)
        k = torch.randn(Ns, H, F, device=dev)
        wq = torch.empty(Uc, H, F, device=dev)
        leg("segment_mm_fwd(unique (rel,dst) rows, q.W_att)", lambda: K.rgnn_relational_matmul(d_col, 1, Watt, q, wq, False),
            uniq_dst * D * 4 + Uc * D * 4 + R * H * F * F * 4 + Uc * 8, "Q_unique*D*s + U*D*s + W + U*i")
        mu = torch.ones(R, H, device=dev) / 4
        msg = torch.randn(Ur, H, F, device=dev)
        new_h = torch.empty(N, H, F, device=dev)
        gout = torch.randn(N, H, F, device=dev)
        ssum = torch.empty(N, H, device=dev)
        if getattr(layer, "fused_attention_aggregation_flag", False) and B.USE_EDGE_PLAN:
            pl, sc = g.get_edge_plan(), g.plan_scratch(H, D)
            attn, lse = torch.empty(0, H, device=dev), torch.empty(N, H, device=dev)   # no saved scores (default)
            fixed_f = pl["fwd"].blob.numel() * 4 - 3 * E * 4   # items, rows, lists: everything but the 3 per-position arrays
            leg("hgt_fused_fwd(plan walk: score + online softmax + aggregate)",
                lambda: K.hetb200_hgt_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, wq, k, msg, mu, attn, lse, new_h,
                                                     sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"]),
                Uc * D * 4 + uniq_src * D * 4 + Ur * D * 4 + 3 * E * 4 + fixed_f + N * H * 4 + N * D * 4,
                "left_c + k_unique + msg_c + int32 (src,urow,ucol) per position + items/rows ; lse + out "
                "(nothing per edge is written: the backward recomputes the score)")
            gl, gk, gm, gmu = torch.zeros_like(wq), torch.empty_like(k), torch.empty_like(msg), torch.zeros(R, H, device=dev)
            fixed_b = pl["bwd"].blob.numel() * 4 - 4 * E * 4
            leg("hgt_fused_bwd(plan walk over (source, relation) runs)",
                lambda: K.hetb200_hgt_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, wq, k, msg, mu, attn, lse, new_h, gout,
                                                     sc["gdot"], gl, gk, gm, gmu),
                Uc * D * 4 + uniq_src * D * 4 + Ur * D * 4 + 3 * E * 4 + fixed_b + 3 * N * H * 4 + 2 * N * D * 4
                + Uc * D * 4 + uniq_src * D * 4 + Ur * D * 4,
                "left_c,k,msg_c + int32 (dst,urow,ucol) per position + lse,out,gout + (lse,gdoto) pairs ; "
                "grad_left_c + grad_k + grad_msg_c (score recomputed, nothing per edge is read)")
        elif getattr(layer, "fused_attention_aggregation_flag", False):
            attn, t = torch.empty(E, H, device=dev), torch.empty(E, H, device=dev)
            leg("hgt_fused_score_softmax_aggregate_fwd(in-CSR, compact)",
                lambda: K.hetb200_hgt_fused_attention_aggregation_csr(
                    ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], iic, iir, wq, k, msg, mu, attn, t,
                    ssum, new_h),
                Uc * D * 4 + uniq_src * D * 4 + Ur * D * 4 + 5 * E * 8 + (N + 1) * 8 + 2 * E * H * 4 + N * H * 4 + N * D * 4,
                "left_c + k_unique + msg_c + eid,rel,src,inv_col,inv_row + row_ptr ; score + exp + sum + out")
            gl, gk, gm, gmu = torch.zeros_like(wq), torch.zeros_like(k), torch.zeros_like(msg), torch.zeros(R, H, device=dev)
            leg("hgt_fused_edge_bwd(out-CSR, compact)",
                lambda: K.hetb200_backward_hgt_fused_attention_aggregation_csr(
                    oc["row_ptrs"], oc["col_indices"], oc["eids"], oc["rel_types"], oic, oir, wq, k, msg, mu, attn, t,
                    ssum, new_h, gout, gl, gk, gm, gmu),
                Uc * D * 4 + uniq_src * D * 4 + Ur * D * 4 + 5 * E * 8 + (Ns + 1) * 8 + 2 * E * H * 4 + N * H * 4
                + 2 * N * D * 4 + Uc * D * 4 + uniq_src * D * 4 + Ur * D * 4,
                "left_c,k,msg_c,idx,score,exp,sum,out,gout ; grad_left_c + grad_k + grad_msg_c")
        else:
            score = torch.empty(E, H, device=dev)
            dinv = {"edata_idx_to_inverse_idx": icx}
            leg("inner_product_fwd(compact left)", lambda: K.rgnn_inner_product_right_node_separatecoo(
                dinv, 2, coo["rel_ptrs"], coo["eids"], coo["row_indices"], coo["col_indices"], wq, k, score),
                Uc * D * 4 + uniq_src * D * 4 + 3 * E * 8 + E * H * 4, "left_c + right_unique + eid,row,inv ; score")
            mu = torch.ones(R, H, device=dev) / 4
            ssum, t, norm = torch.empty(N, H, device=dev), torch.empty(E, H, device=dev), torch.empty(E, H, device=dev)
            leg("hgt_edge_softmax_fwd(in-CSR)", lambda: K.hgt_full_graph_edge_softmax_ops_csr(
                ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], score, mu, ssum, t, norm),
                E * H * 4 + 2 * E * 8 + (N + 1) * 8 + 2 * E * H * 4 + N * H * 4, "a + eid,rel + row_ptr ; t + norm + sum")
            msg = torch.randn(Ur, H, F, device=dev)
            new_h = torch.empty(N, H, F, device=dev)
            leg("hgt_aggregate_fwd(in-CSR, compact msg)", lambda: K.hetb200_hgt_message_aggregation_csr_compact(
                ic["row_ptrs"], ic["eids"], iir, msg, t, ssum, new_h),
                Ur * D * 4 + E * H * 4 + 2 * E * 8 + (N + 1) * 8 + N * H * 4 + N * D * 4, "msg_c + t + eid,inv + row_ptr + sum ; out")
            gout = torch.randn(N, H, F, device=dev)
            gs = torch.randn(E, H, device=dev)
            gl, gr = torch.zeros_like(wq), torch.zeros_like(k)
            leg("inner_product_bwd(compact left)", lambda: K.backward_inner_product_right_node_separatecoo(
                dinv, 2, coo["rel_ptrs"], coo["eids"], coo["row_indices"], coo["col_indices"], wq, k, gs, gl, gr),
                2 * Uc * D * 4 + 2 * uniq_src * D * 4 + 3 * E * 8 + E * H * 4,
                "left_c + right_unique + g + idx ; grad_left_c + grad_right_unique")
            ga, gmu = torch.empty(E, H, device=dev), torch.zeros(R, H, device=dev)
            leg("hgt_edge_softmax_bwd(out-CSR, compact msg)", lambda: K.hetb200_backward_hgt_edge_softmax_ops_csr_compact(
                oc["col_indices"], oc["eids"], oc["rel_types"], oir, msg, score, norm, new_h, gout, mu, ga, gmu),
                Ur * D * 4 + 2 * E * H * 4 + 2 * N * D * 4 + 4 * E * 8 + E * H * 4, "msg_c + a,norm + out,gout + idx ; grad_a")
            gm = torch.zeros_like(msg)
            leg("hgt_aggregate_bwd(out-CSR, compact msg)", lambda: K.hetb200_backward_hgt_message_aggregation_csr_compact(
                oc["col_indices"], oc["eids"], oir, norm, gout, gm),
                E * H * 4 + N * D * 4 + 3 * E * 8 + Ur * D * 4, "norm + gout + idx ; grad_msg_c")
    del flush
    return out


# ---------------------------------------------------------------------------
def _sha1(path):
    import hashlib
    try:
        with open(path, "rb") as f:
            return hashlib.sha1(f.read()).hexdigest()
    except OSError:
        return None


def _secondary_lines(args, dev, torch):
    """Named secondary measurements beside the headline (resident arm only, same timing rules):
    configs[1] (RGAT on the BGS-shaped graph) and the headline workload with bf16 table storage."""
    from het_b200 import graphs
    from het_b200.train import FullGraphTrainer
    out = {}
    for name, (kind, shape, heads, layers, kw) in {
            "rgat_bgs (configs[1]) f32": ("rgat", "bgs", 1, 1, {}),
            "hgt_mag (configs[2]) bf16 storage of the gathered tables, fp32 arithmetic (parity bar 2e-2)":
                ("hgt", "mag", 4, 2, {"bf16_tables": True})}.items():
        coo = graphs.SHAPES[shape](scale=1.0)
        tr = FullGraphTrainer(kind, coo, dev, heads=heads, layers=layers, cuda_graph=not args.no_graph,
                              compact_as_of_node=True, fused=True, **kw)
        for _ in range(max(3, args.warmup)):
            tr.step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            tr.step()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / args.steps
        out[name] = {"ms_per_step": ms, "value": coo.num_edges * layers / (ms * 1e-3), "unit": UNIT,
                     "edges_per_step": coo.num_edges * layers}
        del tr
        torch.cuda.empty_cache()
    return out


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _make_graph(args):
    from het_b200 import graphs
    kind, shape, heads, layers, cfg = WORKLOADS[args.workload]
    scale = args.scale if args.scale is not None else 1.0
    return graphs.SHAPES[shape](scale=scale), scale


def _scaling_label(args, scale):
    """"weak" when the graph grows with the GPU count (--scale == --gpus > 1), else "strong" (total work fixed)."""
    return "weak" if (args.gpus > 1 and abs(scale - args.gpus) < 1e-9) else "strong"


def run_reference(args, rank):
    """The reference arm: the repo's DGL baseline (restated, oracle/models_ref.py) on the host
    cores. Rank 0 only; other ranks exit without work."""
    if rank != 0:
        return
    coo, scale = _make_graph(args)
    kind, shape, heads, layers, cfg = WORKLOADS[args.workload]
    val, ms, cores, sample, done = cpu_dgl_baseline(args.workload, coo, args.steps, args.warmup, budget_s=150.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": _scaling_label(args, scale),
        "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload} ({cfg}): {kind.upper()} {layers} layer(s), hidden 64, {heads} head(s), "
                               f"{coo.name} x{scale:g} N={coo.num_nodes} R={coo.num_rels} E={coo.num_edges}",
                   "timing": "time.perf_counter around fwd+loss+bwd+Adam, all host cores"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def _teardown(trainer, world):
    """Multi-GPU exit. The captured step graphs hold NCCL work, so they go before the communicator;
    a communicator teardown that still blocks (seen with graph-captured collectives) must not turn a
    finished measurement into a hang, so a watchdog ends the process once the line is out."""
    if world <= 1:
        return
    import torch
    import torch.distributed as dist
    sys.stdout.flush()
    def _stuck():
        sys.stderr.write("bench.py: NCCL communicator teardown did not return within 60 s; exiting with code 3\n")
        sys.stderr.flush()
        os._exit(3)

    dog = threading.Timer(60.0, _stuck)
    dog.daemon = True
    dog.start()
    trainer._graphs.clear()
    trainer._loss.clear()
    torch.cuda.synchronize()
    dist.destroy_process_group()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback for the product path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert world == max(1, args.gpus), f"--gpus {args.gpus} but WORLD_SIZE={world}"

    from het_b200.kernels import launch_count
    from het_b200.train import FullGraphTrainer

    kind, shape, heads, layers, cfg = WORKLOADS[args.workload]
    coo, scale = _make_graph(args)
    # whole-step CUDA-graph capture; with NCCL collectives inside only when asked for (HETB_GRAPH_MULTI=1)
    use_graph = (not args.no_graph) and (world == 1 or os.environ.get("HETB_GRAPH_MULTI", "1") == "1")
    trainer = FullGraphTrainer(kind, coo, dev, heads=heads, layers=layers, rank=rank, world=world,
                               cuda_graph=use_graph, compact_as_of_node=args.compact, fused=args.fused,
                               bf16_tables=args.bf16_tables)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident arm -----------------------------------------------------
    l0 = launch_count()
    trainer._zero("embed")
    eager_loss = None
    if use_graph:  # count this library's launches per step once, eagerly (graph replays bypass the host counter)
        trainer.cuda_graph = False
        el = trainer.step().clone()
        if world > 1:
            dist.all_reduce(el)
        eager_loss = float(el)
        trainer.cuda_graph = True
    launches_per_step = launch_count() - l0
    for _ in range(max(3, args.warmup)):
        trainer.step()
    if not use_graph:
        l1 = launch_count(); trainer.step(); launches_per_step = launch_count() - l1
    barrier()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        loss = trainer.step()
    ev1.record()
    barrier()
    ms_total = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_step = float(ms_total.item()) / args.steps
    clocks = sampler.stop() if sampler else None
    if world > 1:   # each rank holds its shard's share of the (sum / N_total) loss
        loss = loss.clone(); dist.all_reduce(loss)
    final_loss = float(loss.item())
    value = coo.num_edges * layers / (ms_step * 1e-3)

    # ---- end-to-end arm (host buffers, H2D + D2H inside the timed region) --
    gen = torch.Generator().manual_seed(7)
    x_host = (torch.rand(trainer.n_local, 64, generator=gen) * 2 - 1).pin_memory()
    y_host = torch.randint(0, 64, (trainer.n_local,), generator=gen).pin_memory()
    x_host2, y_host2 = x_host.clone().pin_memory(), y_host.clone().pin_memory()
    batches = [(x_host, y_host), (x_host2, y_host2)]
    for _ in range(max(3, args.warmup)):
        trainer.step_from_host(x_host, y_host)
    # (i) one step at a time: copy in, step, read the loss, all serialised
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        e2e_loss = trainer.step_from_host(x_host, y_host)
    e1.record()
    barrier()
    e_ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
    e2e_serial_ms_step = float(e_ms.item()) / args.steps
    # (ii) the epoch loop a user runs (FullGraphTrainer.run_from_host): same copies and loss reads
    # per step, the next step's H2D overlapping this step's kernels. This is the headline e2e.
    trainer.run_from_host(batches[i & 1] for i in range(max(3, args.warmup)))
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_losses = trainer.run_from_host(batches[i & 1] for i in range(args.steps))
    e1.record()
    barrier()
    assert len(e2e_losses) == args.steps
    e2e_loss = e2e_losses[-1]
    e_ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
    e2e_ms_step = float(e_ms.item()) / args.steps
    e2e_value = coo.num_edges * layers / (e2e_ms_step * 1e-3)

    # ---- exchange cost, collective by collective (N > 1): the same calls the step makes, timed alone ----
    comm = None
    if world > 1:
        plan = trainer.plan
        part = torch.randn(plan.max_rows, 64, device=dev)
        full = torch.empty(plan.max_rows * world, 64, device=dev)
        nparam = sum(p.numel() for p in trainer.model.parameters())
        flat = torch.zeros(nparam, device=dev)

        def t_coll(fn, reps=10):
            for _ in range(3):
                fn()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            barrier()
            t = torch.tensor([a.elapsed_time(b) / reps], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        if plan.exchange == "halo":   # the all-to-all of the referenced rows and its reverse
            sb = torch.randn(sum(plan.send_counts), 64, device=dev)
            rb = torch.empty(sum(plan.recv_counts), 64, device=dev)
            ag = t_coll(lambda: dist.all_to_all_single(rb, sb, plan.recv_counts, plan.send_counts))
            rs = t_coll(lambda: dist.all_to_all_single(sb, rb, plan.send_counts, plan.recv_counts))
            nbytes = (sum(plan.recv_counts) - plan.recv_counts[rank]) * 64 * 4
            del sb, rb
        else:
            ag = t_coll(lambda: dist.all_gather_into_tensor(full, part))
            rs = t_coll(lambda: dist.reduce_scatter_tensor(part, full))
            nbytes = plan.max_rows * (world - 1) * 64 * 4
        ar = t_coll(lambda: dist.all_reduce(flat))
        comm = {"exchange": plan.exchange, "forward_exchange_ms": ag, "backward_exchange_ms": rs, "weight_all_reduce_ms": ar,
                "rows_read_of_all_rows": round(plan.num_gathered_rows / (plan.max_rows * world), 4),
                "rows_per_rank": int(plan.max_rows), "bytes_received_per_rank_per_collective": int(nbytes),
                "forward_GBps_in": nbytes / (ag * 1e-3) / 1e9, "backward_GBps": nbytes / (rs * 1e-3) / 1e9,
                "collectives_per_step": {"forward_exchange": layers, "backward_exchange": layers, "all_reduce": 1},
                "comm_ms_per_step_if_fully_exposed": layers * (ag + rs) + ar,
                "edges_per_rank": [int(v) for v in plan.edges_per_rank]}
        del part, full, flat

    if rank != 0:
        _teardown(trainer, world)
        return

    peak, peak_src = _peaks()
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": _scaling_label(args, scale), "vs_baseline": None,
        "dtype": "f32 arithmetic, bf16 storage of the gathered tables" if args.bf16_tables else "f32",
        "data": "synthetic",
        "config": {
            "workload": f"{args.workload} ({cfg}): {kind.upper()} {layers} layer(s) fwd+loss+bwd+Adam, hidden 64, "
                        f"{heads} head(s), {coo.name} x{scale:g} N={coo.num_nodes} R={coo.num_rels} E={coo.num_edges}",
            "parallelism": "single GPU" if world == 1 else f"destination shards x{world} (rank p owns slice p of every node "
                           "type's id range), per layer one NCCL " + ("all-to-all of the layer-input rows each shard's "
                           "in-edges reference (halo) and its reverse in backward" if trainer.plan.exchange == "halo" else
                           "all-gather of the layer input and a reduce-scatter in backward") + ", all-reduce of weight grads",
            "cuda_graph": use_graph, "compact_as_of_node": bool(args.compact), "fused_edge_pass": bool(args.fused and args.compact),
            "l2": "inputs larger than L2: per-edge tensors of one step total "
                  f"{coo.num_edges // world * 64 * 4 * 6 / 1e6:.0f} MB per GPU vs 126 MB L2 (no flush between steps)",
            "edges_per_step": coo.num_edges * layers,
        },
        "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms_step,
                "api": "FullGraphTrainer.run_from_host (double-buffered H2D on a copy stream, loss read one step late)",
                "serial_ms_per_step": e2e_serial_ms_step,
                "h2d_bytes_per_step": trainer.h2d_bytes_per_step() * world,
                "d2h_bytes_per_step": trainer.d2h_bytes_per_step() * world, "loss": e2e_loss},
        "gpu_launches": int(launches_per_step * args.steps),
        "gpu_launches_per_step": int(launches_per_step),
        "clocks": clocks, "loss": final_loss, "eager_first_loss": eager_loss,
    }
    if comm is not None:
        line["comm"] = comm
    if not args.no_roofline:
        legs = (kernel_rooflines_compact if (args.compact or kind == "rgcn") else kernel_rooflines)(trainer, peak, peak_src)
        # measured DRAM traffic per launch (one ncu --set full capture, profiles/ncu_traffic.json), only for
        # the configuration it was captured on; "dram_frac" = that traffic / this run's kernel time / peak
        # The file is stamped with the sha1 of the kernel sources it was captured from; a stale stamp
        # (kernels edited since the capture) drops the traffic figures instead of reporting old ones.
        traffic, traffic_note = {}, None
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                tj = json.load(f)
            stale = [fn for fn, h in tj.get("source_sha1", {}).items() if _sha1(os.path.join(ROOT, fn)) != h]
            if stale or not tj.get("source_sha1"):
                traffic_note = f"profiles/ncu_traffic.json is stale for {stale or 'all sources (no stamp)'}: traffic omitted"
            elif scale == 1.0 and args.compact and args.fused and world == 1 and not args.bf16_tables:
                traffic = tj.get(args.workload, {})
        except (OSError, ValueError):
            traffic_note = "profiles/ncu_traffic.json missing"
        for d in legs:
            if d["kernel"] in traffic:
                d["traffic"] = int(traffic[d["kernel"]])
                d["dram_frac"] = round(d["traffic"] / (d["ms"] * 1e-3) / 1e9 / peak, 4)
        top = max(legs, key=lambda d: d["ms"])
        line["roofline"] = {k: top.get(k) for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "dram_frac",
                                                     "kernel", "ms", "algorithmic_bytes", "peak_source")}
        if traffic_note:
            line["roofline"]["traffic_note"] = traffic_note
        line["kernels"] = legs
    # step-level roofline: SURVEY.md 8(d) "fused layer" compulsory bytes (layer input and output once, int32
    # src / rel per edge, row pointers, the saved score per edge and head, log-sum-exp per node and head;
    # backward = forward + the output gradient read and the input gradient written) over the step time
    H_, D_ = heads, 64
    n_l, e_l = trainer.n_local, trainer.num_edges_local
    fwd_b = 2 * n_l * D_ * 4 + e_l * 8 + (n_l + 1) * 4 + e_l * H_ * 4 + n_l * H_ * 4
    bwd_b = fwd_b + 2 * n_l * D_ * 4
    step_bytes = layers * (fwd_b + bwd_b)
    line["step_roofline"] = {"bound": "hbm", "algorithmic_bytes_per_gpu": int(step_bytes),
                             "achieved": round(step_bytes / (ms_step * 1e-3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                             "frac": round(step_bytes / (ms_step * 1e-3) / 1e9 / peak, 4),
                             "note": "SURVEY 8(d) fused-layer compulsory bytes x layers x (fwd + bwd) / ms_per_step"}
    if world == 1 and args.workload == "hgt_mag" and not args.no_extra and not args.bf16_tables:
        line["extra"] = _secondary_lines(args, dev, torch)
    if world == 1 and not args.no_cpu_baseline:
        trainer = None
        torch.cuda.empty_cache()
        big = coo.num_edges > 5_000_000  # one full-graph CPU step at mag / wikikg2 size is ~20 s: no warm-up step
        v, ms, cores, sample, done = cpu_dgl_baseline(args.workload, coo, 1 if big else 3, 0 if big else 1, budget_s=25.0)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                                "ms_per_step": ms}
    print(json.dumps(line), flush=True)
    _teardown(trainer, world)


if __name__ == "__main__":
    main()
