"""Kernel table of ONE full-graph training step (forward + loss + backward + Adam) of the bench workload,
run eagerly (no CUDA graph) under torch.profiler's CUDA activity: live, warm kernel times whose sum can be
compared with bench.py's ms_per_step. Usage (GPU box): python tools/profile_step.py [hgt_mag|rgat_bgs|...] [bf16]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from het_b200.train import FullGraphTrainer

workload = sys.argv[1] if len(sys.argv) > 1 else "hgt_mag"
bf16 = len(sys.argv) > 2 and sys.argv[2] == "bf16"
kind, shape, heads, layers, cfg = bench.WORKLOADS[workload]
args = bench.parse_args(["--workload", workload])
coo, scale = bench._make_graph(args)
dev = torch.device("cuda:0")
trainer = FullGraphTrainer(kind, coo, dev, heads=heads, layers=layers, cuda_graph=False, compact_as_of_node=True,
                           fused=True, bf16_tables=bf16)
for _ in range(3):
    trainer.step()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    trainer.step()
b.record()
torch.cuda.synchronize()
print(f"{workload}{' bf16 tables' if bf16 else ''}: eager step {a.elapsed_time(b) / 5:.3f} ms", flush=True)

from torch.profiler import ProfilerActivity, profile

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        trainer.step()
    torch.cuda.synchronize()
ka = prof.key_averages()
rows = sorted(((k.device_time_total / 3, k.count // 3, k.key) for k in ka), reverse=True)
tot = sum(r[0] for r in rows)
print(f"sum of kernel times {tot / 1e3:.3f} ms per step, {sum(r[1] for r in rows)} launches")
for t, n, name in rows[:40]:
    print(f"{t:10.1f} us {n:4d}x {100 * t / tot:5.1f}%  {name[:110]}")
