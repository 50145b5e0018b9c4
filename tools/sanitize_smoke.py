#!/usr/bin/env python
"""Small run that touches every kernel family, meant for `compute-sanitizer --tool memcheck python tools/sanitize_smoke.py`:
N not a multiple of 64, init sweep, every force variant on ragged active lists, hook loop, fused run with the persistent
kernel, energy, k-th neighbour.  Prints 'sanitize_smoke ok' at the end; results are checked against the oracle."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer  # noqa: E402
from oracle import oracle as O  # noqa: E402

n = 1000
m, r, v = plummer(n, seed=77)
h = Hermite4B200(m, r, v, softening=1e-4, eta=0.01, device=0)
o = O.PortOracle(m, r, v, softening=1e-4, eta=0.01)
h.init_acc_jrk(); o.init_acc_jrk()
fr = o.get_state("f6")["f6"]
rng = np.random.default_rng(3)
short = bool(os.environ.get("GDY_SANIT_SHORT"))   # racecheck is slow: one force variant per size, the step kernels in full
for nact in ((33,) if short else (1, 3, 33, 200, 999, 1000)):
    move = np.sort(rng.choice(n, size=nact, replace=False)).astype(np.uint32)
    for var in ((-1,) if short else range(-1, 16)):
        h.set_tuning(var, -1)
        f, _ = h.force_sweep(move)
        ea = (np.linalg.norm(f[:, :3] - fr[move, :3], axis=1) / np.linalg.norm(fr[move, :3], axis=1)).max()
        assert ea < 1e-12, (nact, var, ea)
h.set_tuning(-1, -1)
at, ar = h.init_dt(0.0), o.init_dt(0.0)
assert at == ar
for persist in (0, -1):
    h.set_persist(persist)
    st, tn, _ = h.run(at, 1e9, max_steps=60, trace_cap=60)
    at = st.atime
sr, tnr, _ = o.run(ar, 1e9, max_steps=120, trace_cap=120)
assert at == sr.atime
e = h.get_energy(); er = o.get_energy()
assert abs(e[0] - er[0]) < 1e-10 * abs(er[0])
h.neighbour_distance(6)
h.close()
print("sanitize_smoke ok")
