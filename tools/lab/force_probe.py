#!/usr/bin/env python
"""Which j-positions does the warp kernel get wrong?  One massive j at a time, all i, against numpy."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gravidy_b200.hermite4 import Hermite4B200
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
rng = np.random.default_rng(1)
r = rng.standard_normal((n, 3)); v = rng.standard_normal((n, 3))
e2 = 1e-8
bad = []
h = Hermite4B200(np.zeros(n), r, v, softening=1e-4)
h.set_tuning(2, int(sys.argv[2]) if len(sys.argv) > 2 else 1)
r4 = np.zeros((n, 4)); v4 = np.zeros((n, 4)); r4[:, :3] = r; v4[:, :3] = v
for j in range(n):
    r4[:, 3] = 0.0; r4[j, 3] = 1.0
    h.upload_particles(r4, v4)
    h.init_acc_jrk()
    f = h.download_forces()
    d = r[j] - r; dv = v[j] - v
    r2 = (d * d).sum(1) + e2
    a = d / r2[:, None] ** 1.5
    err = np.abs(f[:, :3] - a).max() / np.abs(a).max()
    if err > 1e-10:
        k = int(np.argmax(np.abs(f[:, :3] - a).max(1)))
        bad.append((j, float(err), k, f[k, :3] / a[k]))
print("n", n, "bad j:", len(bad))
for b in bad[:40]:
    print(b)
print("bad j list:", [b[0] for b in bad])
