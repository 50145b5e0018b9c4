"""Run the PRODUCT BINARY (gravidy_b200/host/gravidy-b200: the reference's command line, text input, .info/.log/snapshot
files) on a BASELINE-sized Plummer model and time it from the outside.

    python tools/host_program_run.py --particles 16384 --tend 1 --gpus 1 [--hooks] [--spectrum salpeter]
        [--save-log profiles/x.log] [--compare-log profiles/x.log]

Prints one JSON line: wall time of the whole process (parse + upload + nine energy evaluations + loop + snapshot
writes), the ElapsedTime column of the last log row (the loop alone, by the program's own clock), the last row's
Iter / Nsteps / Energy / CumE, and — with --compare-log — whether the schedule columns (IteTime, Iter, Nsteps) of every
row equal the other run's (e.g. the -g 1 log against a -g 8 run) and how far the energies are apart.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402

BIN = os.path.join(ROOT, "gravidy_b200", "host", "gravidy-b200")


def rows(text):
    return np.array([[float(t) for t in ln.split()] for ln in text.splitlines() if ln.strip() and not ln.startswith("#")])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=16384)
    ap.add_argument("--tend", type=float, default=1.0)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--spectrum", default="equal")
    ap.add_argument("--hooks", action="store_true")
    ap.add_argument("--save-log")
    ap.add_argument("--compare-log")
    ap.add_argument("--timeout", type=float, default=1500)
    a = ap.parse_args()
    n = a.particles
    m, r, v = plummer(n, seed=default_seed(n), mass_spectrum=a.spectrum)
    with tempfile.TemporaryDirectory() as d:
        t0 = time.perf_counter()
        with open(os.path.join(d, "input.dat"), "w") as f:
            ml, rl, vl = m.tolist(), r.tolist(), v.tolist()   # Python floats: repr() round-trips every bit
            for i in range(n):
                f.write(f"{i} {ml[i]!r} {rl[i][0]!r} {rl[i][1]!r} {rl[i][2]!r} {vl[i][0]!r} {vl[i][1]!r} {vl[i][2]!r}\n")
        t_write = time.perf_counter() - t0
        cmd = [BIN, "-i", "./input.dat", "-o", "./run", "-t", repr(a.tend), "-g", str(a.gpus)] + (["--hooks"] if a.hooks else [])
        t0 = time.perf_counter()
        p = subprocess.run(cmd, cwd=d, capture_output=True, text=True, timeout=a.timeout)
        wall = time.perf_counter() - t0
        out = {"cmd": " ".join(cmd[1:]), "n": n, "spectrum": a.spectrum, "gpus": a.gpus, "rc": p.returncode,
               "wall_s": wall, "input_write_s": t_write}
        if p.returncode != 0:
            out["stderr"] = p.stderr[-1500:]
            print(json.dumps(out))
            raise SystemExit(1)
        log = open(os.path.join(d, "run.out.log")).read()
        g = rows(log)
        last = g[-1]
        out.update({"log_rows": len(g), "itime": last[1], "iter": int(last[2]), "nsteps": int(last[3]), "energy": last[4],
                    "cum_rel_energy_error": last[6], "loop_elapsed_s": last[7], "gflops_column": last[8],
                    "outside_the_loop_s": wall - last[7], "success": "[SUCCESS]" in p.stderr})
        if a.save_log:
            os.makedirs(os.path.dirname(os.path.abspath(a.save_log)), exist_ok=True)
            with open(a.save_log, "w") as f:
                f.write(log)
        if a.compare_log:
            o = rows(open(a.compare_log).read())
            same = o.shape == g.shape and bool(np.array_equal(o[:, :4], g[:, :4]))
            out["compare"] = {"with": a.compare_log, "schedule_columns_identical": same,
                              "max_rel_energy_difference": float(np.max(np.abs(o[:, 4] - g[:, 4]) / np.abs(o[:, 4]))) if o.shape == g.shape else None}
        print(json.dumps(out))


if __name__ == "__main__":
    main()
