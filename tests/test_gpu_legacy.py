"""GPU suite: cross-check against the reference's OWN CUDA kernels (k_init_acc_jrk, k_update + k_reduce from
src/include/gpu/Hermite4GPU_kernels.cu, recompiled for sm_100a into baseline/_ref/liblegacy_gpu.so by
baseline/Makefile).  Same predicted particle set in, forces out; the two GPU implementations must agree to
the north star's 1e-12 (they differ in summation order, FMA contraction and the rsqrt refinement only)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
LIB = os.path.join(ROOT, "baseline", "_ref", "liblegacy_gpu.so")


@pytest.mark.skipif(not os.path.exists(LIB), reason="baseline/_ref/liblegacy_gpu.so not built (needs /root/reference)")
@pytest.mark.parametrize("spectrum", ["equal", "salpeter"])
def test_forces_agree_with_the_reference_gpu_kernels(spectrum):
    import legacy_gpu as L
    from gravidy_b200.hermite4 import Hermite4B200
    from gravidy_b200.plummer import plummer

    n = 4096
    m, r, v = plummer(n, seed=77, mass_spectrum=spectrum)
    h = Hermite4B200(m, r, v)
    h.init_acc_jrk()
    ours = h.download_forces()
    lib = L.load()
    pred7 = np.ascontiguousarray(np.concatenate([r, v, m[:, None]], axis=1))

    def rel(a, b):
        return (np.linalg.norm(a - b, axis=1) / np.linalg.norm(b, axis=1)).max()

    legacy, _, _ = L.legacy_forces(lib, pred7, h.e2, None, True, 1)
    assert rel(legacy[:, :3], ours[:, :3]) < 1e-12 and rel(legacy[:, 3:], ours[:, 3:]) < 1e-12
    rng = np.random.default_rng(3)
    for nact in (1, 33, 1000):
        move = np.sort(rng.choice(n, size=nact, replace=False)).astype(np.uint32)
        legacy, _, _ = L.legacy_forces(lib, pred7, h.e2, move, True, 1)
        got, _ = h.force_sweep(move)
        assert rel(legacy[:, :3], got[:, :3]) < 1e-12 and rel(legacy[:, 3:], got[:, 3:]) < 1e-12, nact
    h.close()
