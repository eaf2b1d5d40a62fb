"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel total time, count, share.
usage: python tools/summarize_launches.py launches.csv [skip_first_n_launches]"""
import collections, csv, sys

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lines = [l for l in open(path) if not l.startswith("==")]
rows = list(csv.DictReader(lines))[skip:]
agg = collections.OrderedDict()
for r in rows:
    name = r["Kernel Name"].split("(")[0][-70:]
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot / 1e3:.3f} ms total (cold-cache, serialised: compare shares)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"{v[1]:11.1f} us {v[0]:5d}x {100 * v[1] / tot:5.1f}%  avg {v[1] / v[0]:9.1f} us  {k}")
