"""Times the plan-based fused edge kernels alone on a BASELINE graph (L2 flushed before every launch).
usage (GPU box): python tools/bench_edge_plan.py [hgt|rgat] [scale] [bf16]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200 import graphs
from het_b200.graph import HetGraph
from het_b200.kernels import K

kind = sys.argv[1] if len(sys.argv) > 1 else "hgt"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
dt = torch.bfloat16 if (len(sys.argv) > 3 and sys.argv[3] == "bf16") else torch.float32
dev = torch.device("cuda")
coo = graphs.mag_shaped(scale) if kind == "hgt" else graphs.bgs_shaped(scale)
g = HetGraph(coo, dev)
u = g.get_separate_unique_node_indices_single_sided()
pl = g.get_edge_plan()
H = 4 if kind == "hgt" else 1
F, D = 64 // H, 64
E, N, Ns, R = g.num_edges, g.num_nodes, g.num_src_nodes, g.num_rels
Ur, Uc = u["node_indices_row"].numel(), u["node_indices_col"].numel()
sc = g.plan_scratch(H, D)
flush = torch.zeros(64 * 1024 * 1024, device=dev)


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


torch.manual_seed(0)
lse, out = torch.empty(N, H, device=dev), torch.empty(N, H, F, device=dev)
gout = torch.randn(N, H, F, device=dev)
if kind == "hgt":
    left = torch.randn(Uc, H, F, device=dev).to(dt); k = torch.randn(Ns, H, F, device=dev).to(dt)
    msg = torch.randn(Ur, H, F, device=dev).to(dt); mu = torch.ones(R, H, device=dev) / 4
    attn = torch.empty(E, H, device=dev)
    f = lambda: K.hetb200_hgt_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu, attn, lse, out,
                                             sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"])
    gl, gk, gm, gmu = torch.zeros(Uc, H, F, device=dev), torch.empty(Ns, H, F, device=dev), torch.empty(Ur, H, F, device=dev), torch.zeros(R, H, device=dev)
    no_attn = torch.empty(0, H, device=dev)   # default: the backward recomputes the score
    b = lambda: K.hetb200_hgt_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, left, k, msg, mu, no_attn, lse, out, gout,
                                             sc["gdot"], gl, gk, gm, gmu)
    b_saved = lambda: K.hetb200_hgt_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, left, k, msg, mu, attn, lse, out, gout,
                                                   sc["gdot"], gl, gk, gm, gmu)
    f_nosave = lambda: K.hetb200_hgt_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu, no_attn, lse, out,
                                                    sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"])
else:
    feat = torch.randn(Ur, H, F, device=dev).to(dt)
    el, er = torch.randn(Ur, H, device=dev), torch.randn(Uc, H, device=dev)
    f = lambda: K.hetb200_rgat_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, feat, el, er, 0.2, lse, out,
                                              sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"])
    gf, gel, ger = torch.empty(Ur, H, F, device=dev), torch.empty(Ur, H, device=dev), torch.zeros(Uc, H, device=dev)
    b = lambda: K.hetb200_rgat_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, feat, el, er, lse, out, gout, sc["gdot"],
                                              gf, gel, ger, 0.2)
tf = timeit(f)
tb = timeit(b)
extra = ""
if kind == "hgt":   # the default backward: two walks (grad_left owned by a destination-major walk); per-kernel split via torch.profiler
    gl2 = torch.empty(Uc, H, F, device=dev)
    attn0 = attn.clone()

    def b2():
        K.hetb200_hgt_fused_bwd_plan2(pl["bwd"].blob, pl["bwd"].meta, pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu,
                                      attn, lse, out, gout, sc["gdot"], gl2, gk, gm, gmu)
    ts = []
    for i in range(13):
        attn.copy_(attn0)
        flush.add_(1.0)
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a_.record(); b2(); b_.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(a_.elapsed_time(b_))
    ts.sort()
    from torch.profiler import ProfilerActivity, profile
    attn.copy_(attn0)
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        b2()
        torch.cuda.synchronize()
    parts = ", ".join(f"{e.key.split('<')[0].replace('void hetb::', '')} {e.device_time_total:.0f} us"
                      for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:4])
    extra = (f" | fwd without saved scores {timeit(f_nosave):.3f} ms, bwd reading saved scores {timeit(b_saved):.3f} ms"
             f" | two-pass bwd {ts[len(ts) // 2]:.3f} ms ({parts}); fwd fix_runs {pl['fwd'].meta[6]}")
    f()  # restore the scores the timing loops above consumed
    attn0 = attn.clone()
    gmu.zero_(); gl.zero_(); b()
    gmu.zero_(); attn.copy_(attn0); b2()
    print("grad_left two-pass vs single walk: max |diff| / max |ref| =",
          float((gl2 - gl).abs().max() / gl.abs().max()), flush=True)
print(f"{kind} x{scale:g} {str(dt)[6:]}: E={E} items fwd/bwd {pl['fwd'].num_items}/{pl['bwd'].num_items} "
      f"pslots {pl['fwd'].meta[5]} fix_rows {pl['fwd'].meta[3]} | fwd {tf:.3f} ms  bwd {tb:.3f} ms "
      f"{extra} [{os.environ.get('HETB_VARIANT', 'default')}]", flush=True)
