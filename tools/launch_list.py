#!/usr/bin/env python
"""Markdown summary of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv python bench.py ...`).

    python tools/launch_list.py gpurun_out/X_launches.csv "title" > profiles/rNN_launches_bench.md
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
names = rows[hdr]
seq = []
for r in rows[hdr + 1:]:
    if len(r) < len(names):
        continue
    d = dict(zip(names, r))
    v = float(d["Metric Value"].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(d["Metric Unit"], 1e-6)
    seq.append((d["Kernel Name"].replace("gdy::", "").split("(")[0], d["Grid Size"], d["Block Size"], v))
agg = collections.OrderedDict()
for k, g, b, v in seq:
    a = agg.setdefault(k, [0, 0.0, g, b])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"# {sys.argv[2] if len(sys.argv) > 2 else 'Launch list'}\n")
print("| kernel | launches | total ms | share | grid | block |\n|---|---|---|---|---|---|")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"| `{k}` | {a[0]} | {a[1]:.3f} | {100 * a[1] / tot:.2f} % | {a[2]} | {a[3]} |")
mean = lambda key: (lambda v: sum(v) / max(len(v), 1))([v for k, _, _, v in seq if key in k])
fw, rd, fl = mean("k_force_warp"), mean("k_reduce_partials"), mean("FillFunctor")
print(f"\nOne timed step = L2 flush (`FillFunctor`, {fl:.3f} ms, outside the CUDA-event bracket) + `k_force_warp` ({fw:.3f} ms) + "
      f"`k_reduce_partials` ({rd:.3f} ms): the force kernel is {100 * fw / (fw + rd):.2f} % of the event-bracketed step.")
