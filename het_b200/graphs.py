"""Deterministic synthetic heterographs shaped like the reference's datasets.

There is no network for DGL/OGB downloads (the reference's loaders,
hrt/python/utils/graphiler_datasets_loader.py, need them), so the benchmark
shapes of BASELINE.json are generated: node types occupy contiguous id ranges
(``node_type_offsets``, as hrt/python/utils/mydglgraph_converters.py:341 builds
them), every relation connects one (src type, dst type) pair, relation sizes are
Zipf-distributed where the real dataset has many relations, and destination
in-degrees are power-law (hubs) so heavy CSR rows are exercised.

Output is an *integrated COO* in random edge order: ``src, dst, etype`` int64
numpy arrays; edge i has referential id i.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class HeteroCOO:
    name: str
    num_nodes: int
    num_rels: int
    src: np.ndarray
    dst: np.ndarray
    etype: np.ndarray
    node_type_offsets: np.ndarray  # [T+1]
    rel_src_type: np.ndarray  # [R]
    rel_dst_type: np.ndarray  # [R]
    meta: dict = field(default_factory=dict)

    @property
    def num_edges(self) -> int:
        return int(self.src.size)

    @property
    def num_ntypes(self) -> int:
        return int(self.node_type_offsets.size - 1)


def _zipf_sizes(rng, total: int, n: int, s: float) -> np.ndarray:
    w = 1.0 / np.arange(1, n + 1, dtype=np.float64) ** s
    sizes = np.maximum(1, np.floor(w / w.sum() * total)).astype(np.int64)
    sizes[0] += total - sizes.sum()  # exact total; relation 0 is the largest
    assert sizes.min() >= 1 and sizes.sum() == total
    return sizes[rng.permutation(n)]


def _powerlaw_pick(rng, lo: int, hi: int, count: int, skew: float) -> np.ndarray:
    """`count` node ids in [lo, hi) with P(rank k) ~ k^-skew over a random ranking."""
    n = hi - lo
    if skew <= 0:
        return rng.integers(lo, hi, size=count, dtype=np.int64)
    cdf = np.cumsum(1.0 / np.arange(1, n + 1, dtype=np.float64) ** skew)
    cdf /= cdf[-1]
    ranks = np.searchsorted(cdf, rng.random(count), side="left")
    np.minimum(ranks, n - 1, out=ranks)
    perm = rng.permutation(n).astype(np.int64)
    return lo + perm[ranks]


def make_heterograph(name, ntype_sizes, rel_types, rel_sizes, seed, dst_skew=0.8) -> HeteroCOO:
    rng = np.random.default_rng(seed)
    offs = np.zeros(len(ntype_sizes) + 1, dtype=np.int64)
    np.cumsum(np.asarray(ntype_sizes, dtype=np.int64), out=offs[1:])
    srcs, dsts, ets = [], [], []
    for r, ((ts, td), m) in enumerate(zip(rel_types, rel_sizes)):
        srcs.append(rng.integers(offs[ts], offs[ts + 1], size=int(m), dtype=np.int64))
        dsts.append(_powerlaw_pick(rng, int(offs[td]), int(offs[td + 1]), int(m), dst_skew))
        ets.append(np.full(int(m), r, dtype=np.int64))
    src, dst, et = np.concatenate(srcs), np.concatenate(dsts), np.concatenate(ets)
    o = rng.permutation(src.size)
    rt = np.asarray(rel_types, dtype=np.int64).reshape(-1, 2)
    return HeteroCOO(name, int(offs[-1]), len(rel_sizes), src[o], dst[o], et[o], offs,
                     rt[:, 0].copy(), rt[:, 1].copy(), {"seed": seed, "dst_skew": dst_skew})


def _random_typed(name, n_nodes, n_types, n_rels, n_edges, zipf_s, seed, scale=1.0):
    rng = np.random.default_rng(seed + 1000)
    n_nodes = max(n_types, int(n_nodes * scale))
    n_edges = max(n_rels, int(n_edges * scale))
    cuts = np.sort(rng.choice(np.arange(1, n_nodes), size=n_types - 1, replace=False)) if n_types > 1 else []
    sizes = np.diff(np.concatenate([[0], cuts, [n_nodes]])).astype(np.int64)
    rel_types = [(int(rng.integers(n_types)), int(rng.integers(n_types))) for _ in range(n_rels)]
    rel_sizes = _zipf_sizes(rng, n_edges, n_rels, zipf_s)
    return make_heterograph(name, sizes, rel_types, rel_sizes, seed)


def aifb_shaped(scale=1.0):  # config A: ~7.3K nodes, 90 relations, ~49K edges
    return _random_typed("aifb_shaped", 7_300, 7, 90, 49_000, 1.0, seed=0, scale=scale)


def bgs_shaped(scale=1.0):  # config B: ~94K nodes, 207 relations, ~0.67M edges
    return _random_typed("bgs_shaped", 94_000, 27, 207, 670_000, 1.0, seed=1, scale=scale)


def wikikg2_shaped(scale=1.0):  # config W: ~2.5M nodes, 535 relations, ~17M edges
    return _random_typed("wikikg2_shaped", 2_500_000, 1, 535, 17_000_000, 1.2, seed=3, scale=scale)


def mag_shaped(scale=1.0):
    """config C: ogbn-mag shape — author, paper, institution, field_of_study;
    writes, cites, affiliated_with, has_topic and their reverses (~1.9M nodes,
    ~21M edges at scale 1)."""
    nt = [max(8, int(x * scale)) for x in (1_135_000, 736_000, 8_700, 60_000)]
    A, P, I, Fd = 0, 1, 2, 3
    fwd = [((A, P), 3_600_000), ((P, P), 2_700_000), ((A, I), 500_000), ((P, Fd), 3_750_000)]
    rel_types, rel_sizes = [], []
    for (s, d), m in fwd:
        rel_types += [(s, d), (d, s)]
        rel_sizes += [max(1, int(m * scale))] * 2
    return make_heterograph("mag_shaped", nt, rel_types, rel_sizes, seed=2)


SHAPES = {"aifb": aifb_shaped, "bgs": bgs_shaped, "mag": mag_shaped, "wikikg2": wikikg2_shaped}
