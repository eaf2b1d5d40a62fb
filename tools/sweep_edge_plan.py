"""Rebuilds csrc/edge_plan.cu with each -D variant given on the command line and times the kernels
(tools/bench_edge_plan.py) — run on the GPU box (nvcc is in the image).
usage: python tools/sweep_edge_plan.py "" "-DHETB_PLAN_FWD_U=2" "-DHETB_PLAN_MINB=2 -DHETB_PLAN_FWD_U=6" ..."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = os.path.join(root, "het_b200", "csrc", "edge_plan.cu")
for var in sys.argv[1:] + [""]:  # ends on the default build
    os.utime(src)
    env = dict(os.environ, HETB_NVCC_EXTRA=var, HETB_VARIANT=var or "default")
    subprocess.run([sys.executable, "-m", "het_b200.build"], cwd=root, env=env, check=True, capture_output=True)
    for args in (["hgt", "1.0"],):
        subprocess.run([sys.executable, os.path.join(root, "tools", "bench_edge_plan.py")] + args, cwd=root, env=env)
