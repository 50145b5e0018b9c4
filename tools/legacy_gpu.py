#!/usr/bin/env python
"""Same-box data point: the reference's own CUDA kernels (k_init_acc_jrk, k_update + k_reduce), recompiled for
sm_100a from /root/reference (baseline/legacy_gpu_harness.cu -> baseline/_ref/liblegacy_gpu.so), against this
repo's force kernels on the same predicted particle set.  Kernel time only on both sides (CUDA events).

    python tools/legacy_gpu.py --n 65536 --nacts 64,1024,16384 --out gpurun_out/legacy_gpu.json
"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

LIB = os.path.join(ROOT, "baseline", "_ref", "liblegacy_gpu.so")


def load():
    lib = ctypes.CDLL(LIB)
    lib.legacy_gpu_forces.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                      ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.legacy_gpu_forces.restype = ctypes.c_int
    return lib


def legacy_forces(lib, pred7, e2, move=None, want=True, reps=3):
    n = len(pred7)
    nact = n if move is None else len(move)
    out = np.empty((nact, 6)) if want else None
    ms = np.zeros(2, dtype=np.float32)
    rc = lib.legacy_gpu_forces(n, e2, pred7.ctypes.data, nact, None if move is None else move.ctypes.data,
                               None if out is None else out.ctypes.data, reps, ms.ctypes.data)
    if rc:
        raise RuntimeError(f"legacy_gpu_forces failed ({rc})")
    return out, float(ms[0]), float(ms[1])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=65536)
    ap.add_argument("--nacts", default="32,256,2048,16384")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    lib = load()
    n = a.n
    m, r, v = plummer(n, seed=default_seed(n))
    h = Hermite4B200(m, r, v)
    h.init_acc_jrk()
    e2 = h.e2
    pred7 = np.ascontiguousarray(np.concatenate([r, v, m[:, None]], axis=1))
    rows = []

    def rel(a_, b_):
        return float((np.linalg.norm(a_ - b_, axis=1) / np.linalg.norm(b_, axis=1)).max())

    # full-active: k_init_acc_jrk against gdy_init_acc_jrk's force sweep
    ours_ms = min(h.force_sweep_full(1) for _ in range(a.reps))
    fo = h.download_forces()
    fl, best, mean = legacy_forces(lib, pred7, e2, None, True, a.reps)
    rows.append({"case": "init_acc_jrk (full-active)", "n": n, "nact": n, "legacy_ms": best, "ours_ms": ours_ms,
                 "legacy_gpairs_s": n * n / best / 1e6, "ours_gpairs_s": n * n / ours_ms / 1e6,
                 "speedup": best / ours_ms, "max_rel_diff_acc": rel(fl[:, :3], fo[:, :3]),
                 "max_rel_diff_jrk": rel(fl[:, 3:], fo[:, 3:])})
    rng = np.random.default_rng(1)
    for nact in [int(x) for x in a.nacts.split(",")]:
        move = np.sort(rng.choice(n, size=nact, replace=False)).astype(np.uint32)
        ts = [h.force_sweep(move, want_forces=False)[1] for _ in range(a.reps + 1)]
        ours_ms = float(min(ts[1:]))
        fo, _ = h.force_sweep(move)
        fl, best, mean = legacy_forces(lib, pred7, e2, move, True, a.reps)
        rows.append({"case": "update_acc_jrk (k_update + k_reduce)", "n": n, "nact": nact, "legacy_ms": best,
                     "ours_ms": ours_ms, "legacy_gpairs_s": nact * n / best / 1e6,
                     "ours_gpairs_s": nact * n / ours_ms / 1e6, "speedup": best / ours_ms,
                     "max_rel_diff_acc": rel(fl[:, :3], fo[:, :3]), "max_rel_diff_jrk": rel(fl[:, 3:], fo[:, 3:])})
    for row in rows:
        print(json.dumps(row), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
