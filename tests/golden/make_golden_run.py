"""Generate tests/golden/ref_run_n64/: the output FILES of the unmodified reference driver
(Hermite4CPU::integration through oracle/_ref/libgravidy_ref.so) for a tiny run with -a:
<out>.info, <out>.log, <out>.snapshot_NN, <out>.all.snap_NNNN.  The host program
gravidy_b200/host/gravidy-b200 must reproduce them (tests/test_gpu_host_program.py).

    python tests/golden/make_golden_run.py        # build container only (needs /root/reference)
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gravidy_b200.plummer import plummer  # noqa: E402
from oracle.oracle import RefOracle, build  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_run_n64")

if __name__ == "__main__":
    build(force=True)
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(OUT)
    m, r, v = plummer(64, seed=3)
    cwd = os.getcwd()
    os.chdir(OUT)   # relative names end up in the .info file
    try:
        o = RefOracle(m, r, v, softening=1e-4, eta=0.01, integration_time=0.5, interval_time=0.125,
                      workdir=".")
        o.set_print_options(print_all=True, print_lagrange=False)  # -l overruns layers_radii in the reference (heap corruption)
        o.set_threads(1)   # the reference's density-centre loop races with more
        o.integration()
        o.close()
    finally:
        os.chdir(cwd)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
