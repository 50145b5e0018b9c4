#!/usr/bin/env python
"""Per-block-step wall time of gdy_run by active-set size (median microseconds per nact bin)."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=16384)
ap.add_argument("--steps", type=int, default=3000)
ap.add_argument("--spectrum", default="equal")
ap.add_argument("--dump", default="")
a = ap.parse_args()
m, r, v = plummer(a.n, seed=default_seed(a.n), mass_spectrum=a.spectrum)
h = Hermite4B200(m, r, v)
h.init_acc_jrk()
at = h.init_dt(0.0)
st, tn, ti, tu = h.run(at, 1e9, max_steps=200, trace_cap=200, trace_us=True)
st, tn, ti, tu = h.run(st.atime, 1e9, max_steps=a.steps, trace_cap=a.steps, itime0=st.itime, trace_us=True)
print(f"N={a.n} steps={st.iterations} sum_nact={st.nsteps} wall={st.wall_ms:.1f} ms "
      f"mean={st.wall_ms * 1e3 / st.iterations:.1f} us/step  pairs/s={st.interactions / st.wall_ms / 1e6:.1f}e9")
if a.dump:
    np.savez(a.dump, nact=tn, us=tu, itime=ti)
pl, ps = h.persist_stats()
prof = h.persist_profile()
print(f"  persistent kernel: {pl} launches, {ps} block steps ({ps / max(pl, 1):.1f} per launch)")
if any(prof) and os.environ.get("GDY_PERSIST_V", "2") != "1":
    names = ["load", "A.publish", "B.predict+scan", "P1.wait_i", "P1.pairs+publish", "P2", "P3.wait", "P3.lists", "exit",
             "-", "-", "-", "xchg.wait_total", "xchg.count", "-", "-"]
    cnt = prof[14] or ps   # block steps the timers ran on (GDY_PERSIST_PROF_NACT restricts them by active-set size)
    print(f"  ns/step by phase (CTA 0, v2, {cnt} profiled block steps): " + "  ".join(
        f"{nm} {v / max(cnt, 1):.0f}" for nm, v in zip(names, prof) if v and nm != "-" and nm != "xchg.count"))
elif any(prof):
    names = ["P1a", "P1b", "bar1", "P2", "bar2", "P3", "exit", "-", "b.stage", "b.sync", "b.pairs", "b.foldsync", "-", "-", "-", "-"]
    print("  ns/step by phase (CTA 0): " + "  ".join(f"{nm} {v / max(ps, 1):.0f}" for nm, v in zip(names, prof) if v))
    for nm, base in (("CTA 0", 16), ("CTA G/2", 80)):
        ring = prof[base:base + 64]
        tags, ts = ring[0::2], ring[1::2]
        k = max([i for i, t in enumerate(tags) if t] + [0]) + 1
        if k > 1:
            print(f"  stamps of one step, {nm}: " + " ".join(f"[{tags[i]}]+{ts[i] - ts[0]}" for i in range(k)))
edges = [0, 1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 4096, 16384, 1 << 30]
for lo, hi in zip(edges[:-1], edges[1:]):
    sel = (tn > lo) & (tn <= hi)
    if sel.any():
        print(f"  nact in ({lo:6d},{hi:10d}]  steps {sel.sum():6d}  median {np.median(tu[sel]):9.1f} us  "
              f"share of wall {100 * tu[sel].sum() / tu.sum():5.1f}%")
