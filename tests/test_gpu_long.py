"""GPU suite: BASELINE.json configs 1-2 end to end — Plummer N = 4096 and N = 16384, eps = 1e-4,
eta = 0.01, integrated to t = 1 NBU by the fused driver, against the summaries the UNMODIFIED reference
(Hermite4CPU) produced on the same models (tests/golden/ref_long_n*.json, made by
tests/golden/make_golden_long.py).  Bar of the north star: the early block schedule identical (before
chaos can separate two faithful integrations), the totals within 1 %, the cumulative relative energy
error at t = 1 within the same decade."""
import json
import math
import os

import numpy as np
import pytest

from gravidy_b200.hermite4 import Hermite4B200
from gravidy_b200.plummer import plummer

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("n", [4096, 16384])
def test_integration_to_t1_matches_reference_run(n):
    g = json.load(open(os.path.join(GOLD, f"ref_long_n{n}.json")))
    m, r, v = plummer(n, seed=g["seed"], mass_spectrum=g["spectrum"])
    h = Hermite4B200(m, r, v, softening=g["softening"], eta=g["eta"])
    h.init_acc_jrk()
    at = h.init_dt(0.0)
    assert at == g["first_atime"]
    e0 = h.get_energy()[0]
    assert abs(e0 - g["e0"]) < 1e-11 * abs(g["e0"])
    cap = len(g["trace_nact"])
    st, tn, ti = h.run(at, g["t_end"], trace_cap=cap)
    e1 = h.get_energy()[0]
    # the first 2000 block steps: same times, same active-set sizes
    k = min(cap, len(tn))
    assert np.array_equal(tn[:k], np.array(g["trace_nact"][:k], dtype=np.uint32))
    assert np.array_equal(ti[:k], np.array(g["trace_itime"][:k]))
    # totals (identical in practice; 1 % allows for a chaotic split late in the run)
    assert abs(int(st.iterations) - g["block_steps"]) <= 0.01 * g["block_steps"]
    assert abs(int(st.nsteps) - g["sum_nact"]) <= 0.01 * g["sum_nact"]
    assert st.itime == g["t_reached"]
    # cumulative relative energy error in the same decade as the reference's
    err = abs((e1 - e0) / e0)
    assert abs(math.log10(err) - math.log10(g["cum_rel_energy_error"])) < 1.0, (err, g["cum_rel_energy_error"])
    h.close()
