"""CPU suite, part 1: the oracle itself.

The C restatement (oracle/hermite_oracle.c) is pinned against fixtures produced by the UNMODIFIED
reference object code (tests/golden/make_golden.py -> oracle/_ref/libgravidy_ref.so); where that
library is present (the build container, or prebuilt on the GPU box) the two are also compared live.
"""
import os

import numpy as np
import pytest

from gravidy_b200.plummer import plummer
from oracle import oracle as O

STEP_FIXTURES = ["plummer_n97_steps", "salpeter_n200_steps"]


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


@pytest.mark.parametrize("name", STEP_FIXTURES)
def test_port_init_matches_reference_fixture(golden_dir, name):
    g = _load(golden_dir, name)
    o = O.PortOracle(g["m"], g["r"], g["v"], softening=float(g["softening"]), eta=float(g["eta"]))
    assert o.e2 == float(g["e2"])
    o.init_acc_jrk()
    atime = o.init_dt(0.0)
    st = o.get_state("f6", "dt")
    assert np.array_equal(st["f6"], g["init_f6"])          # bit-exact: same order, no FMA
    assert np.array_equal(st["dt"], g["init_dt"])
    assert atime == float(g["init_atime"])
    e = o.get_energy()
    np.testing.assert_allclose(e, g["e0"], rtol=1e-13)


@pytest.mark.parametrize("name", STEP_FIXTURES)
def test_port_steps_match_reference_fixture(golden_dir, name):
    g = _load(golden_dir, name)
    o = O.PortOracle(g["m"], g["r"], g["v"], softening=float(g["softening"]), eta=float(g["eta"]))
    o.init_acc_jrk()
    atime = o.init_dt(0.0)
    nsteps = len(g["nact"])
    for s in range(nsteps):
        itime = atime
        nact = o.find_particles_to_move(itime)
        assert nact == g["nact"][s] and itime == g["itime"][s]
        if s == nsteps - 1:
            pre = o.get_state()
            for k, a in pre.items():
                assert np.array_equal(a, g["pre_" + k]), k
        o.save_old_acc_jrk(nact)
        o.predicted_pos_vel(itime)
        if s == nsteps - 1:
            assert np.array_equal(o.get_state("p7")["p7"], g["last_p7"])
        o.update_acc_jrk(nact)
        if s == nsteps - 1:
            assert np.array_equal(o.get_state("f6")["f6"], g["last_f6"])
        o.correction_pos_vel(itime, nact)
        atime = o.next_integration_time()
    post = o.get_state()
    for k, a in post.items():
        assert np.array_equal(a, g["post_" + k]), k
    assert atime == float(g["final_atime"])


def test_port_run_matches_reference_fixture(golden_dir):
    g = _load(golden_dir, "plummer_n1024_t025_eps1e-4")
    m, r, v = plummer(int(g["n"]), seed=int(g["seed"]))
    o = O.PortOracle(m, r, v, softening=float(g["softening"]), eta=float(g["eta"]))
    o.init_acc_jrk()
    atime = o.init_dt(0.0)
    e0 = o.get_energy()
    st, tn, _ = o.run(atime, float(g["t_end"]), trace_cap=1 << 16)
    assert st.iterations == int(g["iterations"])
    assert st.nsteps == int(g["nsteps"])
    assert st.interactions == int(g["interactions"])
    assert st.itime == float(g["itime"]) and st.atime == float(g["atime"])
    assert np.array_equal(tn, g["nact"])
    np.testing.assert_allclose(e0, g["e0"], rtol=1e-13)
    np.testing.assert_allclose(o.get_energy(), g["e1"], rtol=1e-12)


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref/libgravidy_ref.so not built here")
def test_port_equals_reference_live():
    m, r, v = plummer(150, seed=3, mass_spectrum="salpeter")
    a = O.RefOracle(m, r, v, softening=1e-3, eta=0.02)
    b = O.PortOracle(m, r, v, softening=1e-3, eta=0.02)
    assert a.e2 == b.e2 and a.eta == b.eta
    for o in (a, b):
        o.init_acc_jrk()
    ta, tb = a.init_dt(0.0), b.init_dt(0.0)
    assert ta == tb
    sa, _, _ = a.run(ta, 0.05)
    sb, _, _ = b.run(tb, 0.05)
    assert (sa.iterations, sa.nsteps, sa.atime) == (sb.iterations, sb.nsteps, sb.atime)
    A, B = a.get_state(), b.get_state()
    for k in A:
        assert np.array_equal(A[k], B[k]), k


def test_normalize_dt_edges():
    m, r, v = plummer(8, seed=1)
    o = O.PortOracle(m, r, v)
    dmin, dmax = 2.0 ** -23, 2.0 ** -3
    old = 2.0 ** -10
    assert o.normalize_dt(old / 8, old, 0.0) == dmin            # <= old/8 goes straight to the floor
    assert o.normalize_dt(old / 8 * 1.01, old, 0.0) == old / 8
    assert o.normalize_dt(old / 2 * 1.01, old, 0.0) == old / 2
    assert o.normalize_dt(old * 1.5, old, 0.0) == old
    assert o.normalize_dt(old * 3, old, 4 * old) == 2 * old      # t/(2 old) integer: may double
    assert o.normalize_dt(old * 3, old, 3 * old) == old          # not commensurate: stays
    assert o.normalize_dt(1.0, dmax, 0.0) == dmax                # clamp
    assert o.normalize_dt(float("nan"), old, 0.0) == old         # NaN falls through


def test_lagrange_radii_fixture_is_the_reference_output(golden_dir):
    """tests/golden/ref_run_n64_radii (the reference driver's -l file and the t = 0 values in full precision) against
    the reference compiled here: NbodyUtils::lagrange_radii / get_core_radius / get_half_mass_relaxation_time
    (reference: src/include/utils/NbodyUtils.cpp:113-160, 223-328).  Skips where oracle/_ref is not built."""
    from oracle import oracle as O
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    from gravidy_b200.plummer import plummer
    m, r, v = plummer(64, seed=3)
    o = O.RefOracle(m, r, v, softening=1e-4, eta=0.01)
    o.set_threads(1)
    lag = o.lagrange_radii()
    core = o.core_and_relaxation()
    o.close()
    g = np.load(os.path.join(golden_dir, "ref_run_n64_radii", "ref_radii.npz"))
    assert np.array_equal(lag, g["lagrange_t0"])
    assert core == (float(g["r_core"]), float(g["t_rlx"]), float(g["r_hm"]))
    assert (np.diff(lag) > 0).all()
    row0 = open(os.path.join(golden_dir, "ref_run_n64_radii", "run.out.radii")).readline().split()
    np.testing.assert_allclose([float(t) for t in row0[2:]], lag, rtol=0, atol=5.01e-7)

