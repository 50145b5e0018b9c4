#!/usr/bin/env python
"""Tuning sweep: device time of the force path (kernel + j-split reduction) per kernel variant and
active-set size.  Writes a CSV; the heuristic in gdy_ctx.cu (pick_variant / plan_force) is set from it.

    python tools/sweep_force.py --n 262144 --out gpurun_out/sweep_256k.csv
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

NAMES = ["warp_i256", "warp_i128", "warp_i64", "warp_i32", "lane_i16", "lane_i8", "lane_i4", "lane_i2",
         "lane_i1", "lane_i1x2", "lane_i1x4", "warp_i512", "warp_i128h", "lane_i16x2", "lane_i16x4", "lane_i1x8"]
TI = [256, 128, 64, 32, 16, 8, 4, 2, 1, 2, 4, 512, 128, 32, 64, 8]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=262144)
    ap.add_argument("--out", default="gpurun_out/sweep.csv")
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--nacts", default="")
    a = ap.parse_args()
    n = a.n
    m, r, v = plummer(n, seed=default_seed(n))
    h = Hermite4B200(m, r, v)
    h.init_acc_jrk()
    rng = np.random.default_rng(0)
    nacts = [int(x) for x in a.nacts.split(",")] if a.nacts else \
        [1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512, 768, 1024, 1536, 2048, 3072,
         4096, 8192, 16384, 32768, 65536]
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    with open(a.out, "w") as f:
        f.write("n,nact,variant,name,jsplit,us,gpairs_per_s\n")
        for nact in nacts:
            if nact > n:
                continue
            move = np.sort(rng.choice(n, size=nact, replace=False)).astype(np.uint32)
            best = (None, 1e30)
            for var in [-1] + list(range(len(NAMES))):
                if var >= 0:
                    # skip hopeless combinations: i-tile far larger than nact or far too many CTAs
                    if TI[var] > 8 * nact and TI[var] > 32:
                        continue
                    if nact / TI[var] > 4096:
                        continue
                for js in ([-1] if var < 0 else [-1, 1]):
                    if js == 1 and nact < 20000:
                        continue
                    h.set_tuning(var, js)
                    ts = []
                    for _ in range(a.reps):
                        _, ms = h.force_sweep(move, want_forces=False)
                        ts.append(ms)
                    us = float(np.median(ts[1:])) * 1e3
                    name = "auto" if var < 0 else NAMES[var]
                    f.write(f"{n},{nact},{var},{name},{js},{us:.2f},{nact * n / us / 1e3:.2f}\n")
                    f.flush()
                    if var >= 0 and us < best[1]:
                        best = (name, us)
            print(f"nact={nact:6d} best={best[0]} {best[1]:.1f} us", flush=True)
    h.set_tuning(-1, -1)


if __name__ == "__main__":
    main()
