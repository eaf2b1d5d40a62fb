"""Writes profiles/ncu_traffic.json from `ncu --set full` captures: per bench leg the DRAM bytes
(dram__bytes_read.sum + dram__bytes_write.sum) of one launch of its dominant kernel, stamped with the sha1 of
the kernel sources (bench.py drops the figures when the stamp no longer matches).
usage: python tools/make_ncu_traffic.py <workload> <file.ncu-rep> [<workload> <file.ncu-rep> ...]"""
import csv, hashlib, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LEGS = {  # kernel-name substring -> bench.py leg name
    "hgt_mag": {"plan_fwd_walk_kernel": "hgt_fused_fwd(plan walk: score + online softmax + aggregate)",
                "plan_fwd_walk8_kernel": "hgt_fused_fwd(plan walk: score + online softmax + aggregate)",
                "plan_hgt_bwd_walk_kernel": "hgt_fused_bwd(plan walk over (source, relation) runs)"},
    "rgat_bgs": {"plan_fwd_walk_kernel": "rgat_fused_fwd(plan walk: logits + online softmax + aggregate)",
                 "plan_fwd_walk8_kernel": "rgat_fused_fwd(plan walk: logits + online softmax + aggregate)",
                 "plan_rgat_bwd_walk_kernel": "rgat_fused_bwd(plan walk over compact source rows)"},
}
SOURCES = ["het_b200/csrc/edge_plan.cu", "het_b200/csrc/common.cuh"]


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


out = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu --set full, B200, scale 1, default path, "
                   "cold cache); written by tools/make_ncu_traffic.py, read by bench.py into roofline.traffic",
       "source_sha1": {f: hashlib.sha1(open(os.path.join(ROOT, f), "rb").read()).hexdigest() for f in SOURCES}}
for wl, rep in zip(sys.argv[1::2], sys.argv[2::2]):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    out.setdefault(wl, {})   # several reports per workload merge (later ones win per leg)
    per_kernel = {}  # kernel substring -> max over its captured launches
    for r in rows[2:]:
        for sub in LEGS[wl]:
            if sub in r[col["Kernel Name"]]:
                b = sum(to_bytes(r[col[k]], units[col[k]]) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
                per_kernel[sub] = int(max(per_kernel.get(sub, 0), b))
    legs = {}
    for sub, b in per_kernel.items():  # a leg that runs two kernels gets their sum
        legs[LEGS[wl][sub]] = legs.get(LEGS[wl][sub], 0) + b
    out[wl].update(legs)
json.dump(out, open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
