#!/usr/bin/env python
"""Per-block-step wall time of the HOOK loop (the reference's own loop body, one C-ABI call per hook) by active-set size,
beside gdy_run on the same steps.  GDY_LIB selects another build of the library."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=16384)
ap.add_argument("--steps", type=int, default=2000)
a = ap.parse_args()
m, r, v = plummer(a.n, seed=default_seed(a.n))
h = Hermite4B200(m, r, v)
h.init_acc_jrk()
at = h.init_dt(0.0)
nacts, us = np.zeros(a.steps, dtype=np.int64), np.zeros(a.steps)
h.sync()
for s in range(a.steps):
    t0 = time.perf_counter()
    nact = h.find_particles_to_move(at)
    h.save_old_acc_jrk(nact); h.predicted_pos_vel(at); h.update_acc_jrk(nact); h.correction_pos_vel(at, nact)
    at = h.next_integration_time()
    us[s] = (time.perf_counter() - t0) * 1e6
    nacts[s] = nact
h2 = Hermite4B200(m, r, v)
h2.init_acc_jrk()
at2 = h2.init_dt(0.0)
st, tn, ti, tu = h2.run(at2, 1e9, max_steps=a.steps, trace_cap=a.steps, trace_us=True)
assert np.array_equal(tn, nacts) and st.atime == at, "hook loop and gdy_run disagree on the schedule"
print(f"N={a.n} steps={a.steps} lib={os.environ.get('GDY_LIB', 'production')}: hook loop {us.sum() * 1e-3:.1f} ms, gdy_run {tu.sum() * 1e-3:.1f} ms, same schedule")
edges = [0, 2, 8, 32, 128, 512, 4096, 1 << 30]
for lo, hi in zip(edges[:-1], edges[1:]):
    sel = (nacts > lo) & (nacts <= hi)
    if sel.any():
        print(f"  nact in ({lo:5d},{hi:10d}]  steps {sel.sum():5d}  hook loop median {np.median(us[sel]):8.1f} us   gdy_run median {np.median(tu[sel]):8.1f} us")
