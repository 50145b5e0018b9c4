"""Long-run summaries of the UNMODIFIED reference (Hermite4CPU, oracle/_ref) for BASELINE.json's
configs 1-2: Plummer N = 4096 and N = 16384, equal mass, eps = 1e-4, eta = 0.01, t = 0 -> 1.

    python tests/golden/make_golden_long.py [n ...]        (N = 16384 needs ~5 min on 8 threads)

Writes tests/golden/ref_long_n<N>.json: block steps, sum of nact, energies at t = 0 and t = 1, the
cumulative relative energy error, and the nact/ITIME of the first 2000 block steps (before chaos
separates two faithful integrations).  The GPU suite integrates the same model and compares
(tests/test_gpu_long.py): the early schedule must be identical and the energy error at t = 1 in the
same decade (BASELINE.json north_star).
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402
from oracle.oracle import RefOracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
TRACE = 2000


def one(n, spectrum="equal", softening=1e-4, eta=0.01, t_end=1.0):
    m, r, v = plummer(n, seed=default_seed(n), mass_spectrum=spectrum)
    o = RefOracle(m, r, v, softening=softening, eta=eta)
    o.set_threads(os.cpu_count() or 1)
    o.init_acc_jrk()
    at = o.init_dt(0.0)
    e0 = o.get_energy()
    w0 = time.perf_counter()
    st, tn, ti = o.run(at, t_end, trace_cap=TRACE)
    wall = time.perf_counter() - w0
    e1 = o.get_energy()
    d = dict(n=n, seed=default_seed(n), spectrum=spectrum, softening=softening, eta=eta, t_end=t_end,
             block_steps=int(st.iterations), sum_nact=int(st.nsteps), pairs=int(st.interactions),
             t_reached=st.itime, e0=e0[0], e1=e1[0], ekin1=e1[1], epot1=e1[2],
             cum_rel_energy_error=abs((e1[0] - e0[0]) / e0[0]),
             first_atime=at, trace_nact=[int(x) for x in tn], trace_itime=[float(x) for x in ti],
             wall_s=wall, update_s=st.update_s, threads=os.cpu_count(),
             made_by="tests/golden/make_golden_long.py on the unmodified reference (oracle/_ref)")
    name = f"ref_long_n{n}" + ("" if spectrum == "equal" else "_" + spectrum) + ".json"
    with open(os.path.join(OUT, name), "w") as f:
        json.dump(d, f)
    print(name, d["block_steps"], d["sum_nact"], d["cum_rel_energy_error"], f"{wall:.1f}s", flush=True)


if __name__ == "__main__":
    ns = [int(x) for x in sys.argv[1:]] or [4096, 16384]
    for n in ns:
        one(n)
