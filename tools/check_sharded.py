"""Multi-GPU parity: destination-range sharded training step (NCCL) vs the single-GPU step.
Run:  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
          --master-port 29511 tools/check_sharded.py [rgat|hgt] [compact]
Every rank also runs the unsharded step on its own GPU and compares its shard of the
logits, the loss and every (all-reduced) weight gradient. rtol 1e-4 (fp32 summation order)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as Fn

from het_b200 import graphs, parallel
from het_b200.train import FullGraphTrainer

kind = sys.argv[1] if len(sys.argv) > 1 else "rgat"
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)

coo = graphs.bgs_shaped(scale=0.2) if kind == "rgat" else graphs.mag_shaped(scale=0.01)
heads, layers = (1, 1) if kind == "rgat" else (4, 2)
# "compact": the sharded arm runs the bench defaults (compact-as-of-node + fused HGT edge pass); the
# single-GPU arm stays on the per-edge path, so the comparison crosses both implementations
compact = len(sys.argv) > 2 and sys.argv[2] == "compact"
sh = FullGraphTrainer(kind, coo, dev, heads=heads, layers=layers, rank=rank, world=world, cuda_graph=False,
                      compact_as_of_node=compact, fused=compact)
full = FullGraphTrainer(kind, coo, dev, heads=heads, layers=layers, rank=0, world=1, cuda_graph=False)
# magnitude-preserving weights (same on both arms and on every rank): logits and every gradient O(0.1-1)
from oracle.testing import unit_gain_init  # noqa: E402
unit_gain_init(sh.model, seed=11)
unit_gain_init(full.model, seed=11)


def close(a, b, what, rtol=1e-4, atol=2e-5):
    """relative to max|reference| (no absolute floor); prints the reference magnitude beside the error"""
    a, b = a.double(), b.double()
    scale = b.abs().max()
    err = (a - b).abs()
    ok = bool((err <= atol * scale + rtol * b.abs()).all()) and bool(torch.isfinite(a).all())
    print(f"[rank {rank}] {what}: max err {err.max():.3e} vs max|ref| {scale:.3e} {'ok' if ok else 'FAIL'}", flush=True)
    return ok


ok = True
# forward
own = sh.owned.to(dev)
logits_sh = sh.forward(sh.embed())
logits_full = full.forward(full.embed())
ok &= close(logits_sh, logits_full[own], "logits shard")
# backward
loss_sh = Fn.nll_loss(Fn.log_softmax(logits_sh, -1), sh.labels, reduction="sum") / coo.num_nodes
loss_sh.backward()
parallel.all_reduce_grads(list(sh.model.parameters()))
loss_full = Fn.nll_loss(Fn.log_softmax(logits_full, -1), full.labels, reduction="sum") / coo.num_nodes
loss_full.backward()
tot = loss_sh.detach().clone()
dist.all_reduce(tot)
ok &= close(tot, loss_full.detach(), "loss")
for (n1, p1), (n2, p2) in zip(sh.model.named_parameters(), full.model.named_parameters()):
    ok &= close(p1.grad, p2.grad, f"grad {n1}", rtol=2e-4, atol=3e-5)
ok &= close(sh.embed.embeds.grad, full.embed.embeds.grad[own], "grad embedding shard", rtol=2e-4, atol=3e-5)
# a few optimizer steps through the public API (eager + captured)
for _ in range(3):
    l_sh = sh.step()
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("SHARDED PARITY", "OK" if int(flag) == 1 else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag) == 1 else 1)
