"""Summarise an .ncu-rep (ncu --set full): per kernel duration, DRAM bytes, DRAM / L2 / tensor
throughput and top stall reasons. usage: python tools/ncu_summary.py file.ncu-rep > summary.txt"""
import csv, io, subprocess, sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 % of peak"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1 % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe % active"),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue slots % busy"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    print("==", r[col["Kernel Name"]][:110])
    for k, nm in KEYS:
        if k in col:
            print(f"   {nm:24s} {r[col[k]]:>16s} {units[col[k]]}")
    stalls = [(float(r[i] or 0), h) for h, i in col.items()
              if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for v, h in sorted(stalls, reverse=True)[:4]:
        print(f"   stall {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]:22s} {v:8.2f} warps/issue")
