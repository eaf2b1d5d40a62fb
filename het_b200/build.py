"""In-tree build of the two shared libraries:

* ``het_b200/lib/libhetb200.so``   — CUDA kernels + C ABI (include/hetb200.h), nvcc, sm_100a only
* ``het_b200/lib/libtorch_hrt.so`` — torch op registration shim (``torch.ops.torch_hrt.*``)

Replaces the reference's CMake build (hrt/CMakeLists.txt:274-294), keeping its one
idea worth keeping: ONE library that registers every op.
Run ``python -m het_b200.build`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB = PKG / "lib"
OBJ = PKG / "build"
INCLUDE = ROOT / "include"

CUDA_SOURCES = ["abi.cu", "edge_ops.cu", "edge_plan.cu", "edge_coo.cu", "segment_mm.cu", "segment_mm_tc.cu", "segment_mm_tc_tma.cu", "segment_mm_tc_wide.cu", "segment_mm_gradw_tc.cu", "convert.cu", "loss.cu", "optim.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    f"-I{INCLUDE}",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed:\n  " + " ".join(map(str, cmd)) + "\n" + r.stdout + r.stderr)


def _stale(out: Path, deps: list[Path]) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build_kernels(force: bool = False) -> Path:
    LIB.mkdir(exist_ok=True)
    OBJ.mkdir(exist_ok=True)
    nvcc = _nvcc()
    headers = [CSRC / "common.cuh", INCLUDE / "hetb200.h"] + sorted(CSRC.glob("*.cuh"))
    jobs = []
    objs = []
    for src in CUDA_SOURCES:
        s = CSRC / src
        o = OBJ / (s.stem + ".o")
        objs.append(o)
        if force or _stale(o, [s] + headers):
            jobs.append([nvcc, *NVCC_FLAGS, *os.environ.get("HETB_NVCC_EXTRA", "").split(), "-c", str(s), "-o", str(o)])
    with ThreadPoolExecutor(max_workers=4) as ex:
        list(ex.map(_run, jobs))
    out = LIB / "libhetb200.so"
    if force or jobs or _stale(out, objs):
        _run([nvcc, "-shared", "-o", str(out), *map(str, objs), "-lcudart"])
    return out


def build_torch_shim(force: bool = False) -> Path:
    import torch
    from torch.utils import cpp_extension as ce

    LIB.mkdir(exist_ok=True)
    src = CSRC / "torch_bindings.cpp"
    out = LIB / "libtorch_hrt.so"
    kern = LIB / "libhetb200.so"
    if not (force or _stale(out, [src, INCLUDE / "hetb200.h"])):
        return out
    torch_lib = Path(torch.__file__).parent / "lib"
    inc = [f"-I{p}" for p in ce.include_paths(device_type="cuda")] + [f"-I{INCLUDE}"]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = [
        "g++", "-O2", "-std=c++17", "-fPIC", "-shared",
        f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-DTORCH_API_INCLUDE_EXTENSION_H",
        *inc, str(src), "-o", str(out),
        f"-L{LIB}", "-lhetb200", f"-L{torch_lib}",
        "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
        "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{torch_lib}",
    ]
    assert kern.exists(), "build kernels first"
    _run(cmd)
    return out


def build_all(force: bool = False) -> None:
    build_kernels(force)
    build_torch_shim(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print("built:", *(p.name for p in sorted(LIB.glob("*.so"))))
