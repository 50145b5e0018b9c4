"""A/B of the persistent kernel's per-step time between two builds of the library (GDY_LIB selects the build).

    GDY_LIB=tools/_old.so python tools/ab_persist_latency.py; python tools/ab_persist_latency.py
Prints one JSON line: microseconds per block step with every step inside the persistent kernel (n <= nact_max).
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

out = {"lib": os.environ.get("GDY_LIB", "in-tree")}
for n, cap in ((1024, 1024), (262144, -1)):
    m, r, v = plummer(n, seed=default_seed(n))
    h = Hermite4B200(m, r, v)
    h.set_persist(cap)
    h.init_acc_jrk()
    at = h.init_dt(0.0)
    best = None
    for rep in range(3):
        st, tn, ti = h.run(at, 1e9, max_steps=4000, trace_cap=0)
        at = st.atime
        us = st.wall_ms * 1e3 / st.iterations
        best = us if best is None else min(best, us)
    out[f"n{n}_us_per_step"] = round(best, 2)
    out[f"n{n}_persist"] = h.persist_stats()
    h.close()
print(json.dumps(out))
