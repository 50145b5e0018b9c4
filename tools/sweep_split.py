#!/usr/bin/env python
"""j-split sweep of the TMA-ring force variants at awkward active-set sizes: how good is plan_force's wave model?"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=262144)
ap.add_argument("--variants", default="0,1,2")
ap.add_argument("--nacts", default="10460,18652,40000")
ap.add_argument("--splits", default="1,2,3,4,5,6,7,8,10,12,14,16,18,20,24,28,32,40,48,64")
ap.add_argument("--reps", type=int, default=4)
ap.add_argument("--world", type=int, default=1, help="logical shard 0 of WORLD (no communicator): the j-range of one GPU of a sharded run")
a = ap.parse_args()
m, r, v = plummer(a.n, seed=default_seed(a.n))
h = Hermite4B200(m, r, v, rank=0, world=a.world)
h.init_acc_jrk()
rng = np.random.default_rng(0)
nj = a.n // a.world
print("n,nact,variant,jsplit,us,gpairs_per_s")
for nact in [int(x) for x in a.nacts.split(",")]:
    move = np.sort(rng.choice(a.n, size=min(nact, a.n), replace=False)).astype(np.uint32)
    for var in [int(x) for x in a.variants.split(",")]:
        for js in [-1] + [int(x) for x in a.splits.split(",")]:
            h.set_tuning(var, js)
            ts = [h.force_sweep(move, want_forces=False)[1] for _ in range(a.reps)]
            us = float(np.median(ts[1:])) * 1e3
            print(f"{a.n},{len(move)},{var},{js},{us:.1f},{len(move) * nj / us / 1e3:.1f}", flush=True)
h.set_tuning(-1, -1)
