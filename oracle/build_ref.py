"""TEST INFRASTRUCTURE ONLY — builds the checker libraries under ``oracle/_ref/``.

``oracle/_ref/`` is git-ignored (never in history) but travels to the GPU box
with the gpurun snapshot, like the product's own ``.so`` files.

* ``libhet_ref_converters.so`` — the REFERENCE's CPU converters
  (/root/reference/hrt/include/MyHyb/MyHyb.h) behind oracle/ref_shim.cpp, g++.
* ``libhet_ref_kernels.so``    — the REFERENCE's CUDA kernel templates
  (hrt/include/DGLHackKernel/{GAT,HGT,RGNN}/*.cu.h) behind oracle/ref_kernels_shim.cu,
  instantiated exactly as the reference's op wrappers instantiate them and compiled
  by nvcc for sm_100a. Runs only on the GPU box; lets the ``-m gpu`` tests compare
  the product against the reference kernels themselves.

Both are compiled from the reference sources *where they lie* (nothing is copied)
and only when /root/reference exists, i.e. in the build container; the GPU box
uses the prebuilt files. The reference's own CMake build is not run (needs glog,
g++-9, cusp vs CUDA 12.9 thrust: see DESIGN.md).
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
REF = Path(os.environ.get("HET_REFERENCE", "/root/reference"))
CUDA_INC = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda")) / "include"


def _run(cmd):
    r = subprocess.run(list(map(str, cmd)), capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed:\n  " + " ".join(map(str, cmd)) + "\n" + r.stdout[-4000:] + r.stderr[-4000:])


def _stale(out: Path, deps) -> bool:
    return (not out.exists()) or any(Path(d).stat().st_mtime > out.stat().st_mtime for d in deps)


def reference_present() -> bool:
    return (REF / "hrt" / "include" / "MyHyb" / "MyHyb.h").exists()


def build_converters(force=False) -> Path | None:
    out = OUT / "libhet_ref_converters.so"
    if not reference_present():
        return out if out.exists() else None
    OUT.mkdir(exist_ok=True)
    src = HERE / "ref_shim.cpp"
    if force or _stale(out, [src, HERE / "stubs" / "cusp" / "csr_matrix.h"]):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w",
              "-DTHRUST_DEVICE_SYSTEM=THRUST_DEVICE_SYSTEM_CPP",
              "-DTHRUST_HOST_SYSTEM=THRUST_HOST_SYSTEM_CPP",
              f"-I{HERE / 'stubs'}", f"-I{REF / 'hrt' / 'include'}", f"-I{CUDA_INC}",
              src, "-o", out])
    return out


def build_kernels(force=False) -> Path | None:
    out = OUT / "libhet_ref_kernels.so"
    src = HERE / "ref_kernels_shim.cu"
    if not reference_present() or not src.exists():
        return out if out.exists() else None
    OUT.mkdir(exist_ok=True)
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if force or _stale(out, [src]):
        inc = REF / "hrt" / "include"
        _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-w",
              "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr",
              # macros.h only defines its SGEMM tile macros for CUDA_ARCHS 80/86/89; 80 selects
              # the A100 tile shape (32x32x8, 64 threads) the reference was tuned with.
              "-DCUDA_ARCHS=80",
              f"-I{inc}", f"-I{inc / 'DGLHackKernel'}", f"-I{inc / 'DGLHackKernel' / 'RGNN'}",
              src, "-o", out, "-lcudart"])
    return out


def build_all(force=False):
    return build_converters(force), build_kernels(force)


if __name__ == "__main__":
    print(*build_all(), sep="\n")
