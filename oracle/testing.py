"""TEST INFRASTRUCTURE ONLY — the comparator and the test initialisation shared by tests/,
__graft_entry__.smoke() and tools/check_sharded.py."""
from __future__ import annotations

import torch

RTOL, ATOL = 1e-4, 1e-5


def close(got, want, what, rtol=RTOL, atol=ATOL):
    """|got - want| <= atol * max|want| + rtol * |want| elementwise. The scale is the reference
    tensor's own magnitude (no floor): a tensor of tiny values is held to a proportionally tiny
    error. An all-zero reference demands an all-zero result."""
    g = got.detach().double().cpu().reshape(-1)
    w = want.detach().double().cpu().reshape(-1)
    assert g.shape == w.shape, f"{what}: shape {got.shape} vs {want.shape}"
    assert bool(torch.isfinite(g).all()), f"{what}: non-finite values in the result"
    if w.numel() == 0:
        return
    scale = w.abs().max()
    err = (g - w).abs()
    bad = err > (atol * scale + rtol * w.abs())
    assert not bool(bad.any()), (f"{what}: max err {err.max():.3e} vs max|ref| {scale:.3e} "
                                 f"(rel {float(err.max() / scale) if scale > 0 else float('inf'):.2e}), "
                                 f"{int(bad.sum())} of {w.numel()} bad")


def unit_gain_init(model, seed=0):
    """Every typed linear preserves magnitude (U(+-sqrt(3 / fan_in)) per [in, out] matrix), relation_pri
    in 0.5..1.5, skip in -1..1, biases +-0.1: layer outputs and EVERY parameter gradient come out
    O(0.1 .. 1), for 2 layers as well. The layers' own xavier init on the 4-D parameter tensors gives
    outputs ~1e-3 and attention-path gradients ~1e-5, against which a wrong kernel cannot be seen."""
    gen = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            leaf = name.rsplit(".", 1)[-1]
            if leaf in ("relation_pri",):
                p.copy_(torch.rand(p.shape, generator=gen) + 0.5)
            elif leaf in ("skip",):
                p.copy_(torch.rand(p.shape, generator=gen) * 2 - 1)
            elif leaf in ("h_bias",):
                p.copy_(torch.rand(p.shape, generator=gen) * 0.2 - 0.1)
            elif leaf in ("attn_l", "attn_r"):       # [R, H, F]: a row-dot over F
                b = (3.0 / p.shape[-1]) ** 0.5
                p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) * b)
            elif p.dim() >= 2:                        # [..., in, out]
                b = (3.0 / p.shape[-2]) ** 0.5
                p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) * b)
