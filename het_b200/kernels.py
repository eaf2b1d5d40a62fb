"""Loads the in-tree ``libtorch_hrt.so`` and exposes ``K = torch.ops.torch_hrt``.

Mirror of the reference's ``hrt/python/kernels/__init__.py:5-17`` (load the
library, call ``build_debug_info`` once, bind ``K``). There is no fallback: if
the library is missing or does not load, importing this module raises.
"""
from __future__ import annotations

import os
from pathlib import Path

import torch

_LIB_DIR = Path(__file__).resolve().parent / "lib"
TORCH_LIB = _LIB_DIR / "libtorch_hrt.so"
KERNEL_LIB = _LIB_DIR / "libhetb200.so"

if not TORCH_LIB.exists() or not KERNEL_LIB.exists():
    raise ImportError(
        f"hetb200 native libraries not found in {_LIB_DIR}; run `python -m het_b200.build` "
        "(there is no Python/CPU fallback for the torch_hrt ops)"
    )

torch.ops.load_library(str(TORCH_LIB))
K = torch.ops.torch_hrt

if os.environ.get("HETB200_PRINT_BUILD_INFO", "0") == "1":
    K.build_debug_info()


def launch_count() -> int:
    """Kernels launched by libhetb200 since process start."""
    return int(K.hetb200_launch_count())
