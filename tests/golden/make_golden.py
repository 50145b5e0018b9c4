"""Generate the golden fixtures under tests/golden/ from the UNMODIFIED reference.

Run in the build container only (needs /root/reference → oracle/_ref/libgravidy_ref.so):

    python tests/golden/make_golden.py

The reference ships no golden vectors, so these are outputs of the reference's own
Hermite4CPU object code (oracle/ref_harness.cpp) on seeded Plummer models from
gravidy_b200/plummer.py.  The fixtures travel to the GPU box; /root/reference does not.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gravidy_b200.plummer import plummer  # noqa: E402
from oracle.oracle import RefOracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def steps_fixture(name, n, seed, spectrum, softening, eta, nsteps):
    m, r, v = plummer(n, seed=seed, mass_spectrum=spectrum)
    o = RefOracle(m, r, v, softening=softening, eta=eta)
    o.set_threads(2)
    o.init_acc_jrk()
    atime = o.init_dt(0.0)
    init = o.get_state("f6", "t", "dt")
    e0 = o.get_energy()
    d = dict(m=m, r=r, v=v, softening=np.float32(softening), eta=np.float32(eta),
             e2=o.e2, init_f6=init["f6"], init_dt=init["dt"], init_atime=atime,
             e0=np.array(e0))
    # step-by-step with per-hook captures for the LAST step
    nacts, itimes = [], []
    for s in range(nsteps):
        itime = atime
        nact = o.find_particles_to_move(itime)
        if s == nsteps - 1:
            pre = o.get_state()
        o.save_old_acc_jrk(nact)
        o.predicted_pos_vel(itime)
        if s == nsteps - 1:
            d["last_p7"] = o.get_state("p7")["p7"]
        o.update_acc_jrk(nact)
        if s == nsteps - 1:
            d["last_f6"] = o.get_state("f6")["f6"]
        o.correction_pos_vel(itime, nact)
        atime = o.next_integration_time()
        nacts.append(nact); itimes.append(itime)
    post = o.get_state()
    for k, a in pre.items():
        d["pre_" + k] = a
    for k, a in post.items():
        d["post_" + k] = a
    d["nact"] = np.array(nacts, dtype=np.uint32)
    d["itime"] = np.array(itimes)
    d["final_atime"] = atime
    d["e1"] = np.array(o.get_energy())
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "n", n, "steps", nsteps, "sum nact", int(np.sum(nacts)))


def run_fixture(name, n, seed, softening, eta, t_end):
    m, r, v = plummer(n, seed=seed)
    o = RefOracle(m, r, v, softening=softening, eta=eta)
    o.set_threads(4)
    o.init_acc_jrk()
    atime = o.init_dt(0.0)
    e0 = o.get_energy()
    st, tn, ti = o.run(atime, t_end, trace_cap=1 << 20)
    e1 = o.get_energy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), n=n, seed=seed,
                        softening=np.float32(softening), eta=np.float32(eta), t_end=t_end,
                        iterations=st.iterations, nsteps=st.nsteps, interactions=st.interactions,
                        itime=st.itime, atime=st.atime, nact=tn, e0=np.array(e0), e1=np.array(e1))
    print(name, st.iterations, st.nsteps, e0[0], e1[0], abs((e1[0] - e0[0]) / e0[0]))


if __name__ == "__main__":
    # ragged N (not a multiple of any tile), equal masses, default softening
    steps_fixture("plummer_n97_steps", 97, 7, "equal", 1e-4, 0.01, 48)
    # Salpeter spectrum: wide dt hierarchy, tiny active blocks
    steps_fixture("salpeter_n200_steps", 200, 11, "salpeter", 1e-4, 0.01, 64)
    # BASELINE configs[0]: N=1024 equal mass, e2=1e-4 (eps=0.01), eta=0.01, to t=1
    run_fixture("plummer_n1024_t1_eps1e-2", 1024, 52, 0.01, 0.01, 1.0)
    # and with the default softening, to t=0.25 (keeps the generator fast)
    run_fixture("plummer_n1024_t025_eps1e-4", 1024, 52, 1e-4, 0.01, 0.25)
