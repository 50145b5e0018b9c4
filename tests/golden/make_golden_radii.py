"""Generate tests/golden/ref_run_n64_radii/run.out.radii: the Lagrange-radii file the unmodified reference driver writes
with -l for the run of tests/golden/ref_run_n64 (same particles, same options; oracle/_ref/libgravidy_ref.so), plus
ref_radii.npz with the t = 0 values in full precision (Lagrange radii, core radius, relaxation time, half-mass radius).

The reference overruns its six-entry layers_radii vector under -l; the harness resizes that public member before the
run (oracle/ref_harness.cpp: gref_pad_layers), the reference's code is not changed.

    python tests/golden/make_golden_radii.py        # build container only (needs /root/reference)
"""
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gravidy_b200.plummer import plummer  # noqa: E402
from oracle.oracle import RefOracle, build  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_run_n64_radii")

if __name__ == "__main__":
    build(force=True)
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(OUT)
    m, r, v = plummer(64, seed=3)
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as d:
        os.chdir(d)
        try:
            o = RefOracle(m, r, v, softening=1e-4, eta=0.01, integration_time=0.5, interval_time=0.125, workdir=".")
            o.set_threads(1)   # the reference's density-centre loop races with more
            lag0 = o.lagrange_radii()
            core = o.core_and_relaxation()
            o.close()
            o = RefOracle(m, r, v, softening=1e-4, eta=0.01, integration_time=0.5, interval_time=0.125, workdir=".")
            o.set_threads(1)
            o.set_print_options(print_all=False, print_lagrange=True)
            o.pad_layers()
            o.integration()
            o.close()
        finally:
            os.chdir(cwd)
        radii = [f for f in os.listdir(d) if f.endswith(".radii")]
        assert len(radii) == 1, os.listdir(d)
        shutil.copy(os.path.join(d, radii[0]), os.path.join(OUT, "run.out.radii"))
    np.savez(os.path.join(OUT, "ref_radii.npz"), lagrange_t0=lag0, r_core=core[0], t_rlx=core[1], r_hm=core[2])
    print(open(os.path.join(OUT, "run.out.radii")).read())
    print(lag0, core)
