"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python tools/summarise_launches.py launches.csv > summary.md"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H, data = rows[hdr], rows[hdr + 1:]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in data:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] == "ns" else v
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
ours = sum(v[1] for k, v in agg.items() if k.startswith("b2::"))
print(f"total {tot / 1e3:.2f} ms over {sum(v[0] for v in agg.values())} launches; b2:: kernels {100 * ours / tot:.1f} % of GPU time\n")
print("| kernel | launches | total us | us/launch | share |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"| `{k[:100]}` | {v[0]} | {v[1]:.0f} | {v[1] / v[0]:.2f} | {100 * v[1] / tot:.1f} % |")
