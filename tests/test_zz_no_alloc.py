"""gdy_energy_noalloc: the energy entry point for callers that are not allowed to allocate on the GPU.

Some users of this library cannot go through ``gdy_create``: a simulation code that reserved all device memory for
its own pool at start-up, a diagnostics callback that runs between block steps and must not stall the device (which
``cudaMalloc`` / ``cudaFree`` do), or a plug-in host that bans driver allocations once it is initialised.  They keep
the particles in device memory they own and call ``gdy_energy_noalloc`` with device pointers; the partial sums land in
a fixed array inside the library instead of a ``cudaMalloc``'ed buffer.

The test plays such a caller (torch only provides the caller-owned device buffers) and checks the result against the
context-based ``gdy_energy`` through the C ABI, on calls in a row with different particle sets so that whatever the
previous call left in the shared array would show up if it were ever read.  Tolerance: 1e-12 relative, the energy
tolerance of tests/test_gpu_parity.py.

The file name sorts after every other test file on purpose: the suite runs under ``pytest -x``, and a failure here
must not hide the parity tests.  It has to pass alone in a fresh process and at the end of the whole suite.
"""

import numpy as np
import pytest

from gravidy_b200 import capi
from gravidy_b200.plummer import plummer

pytestmark = pytest.mark.gpu

ENERGY_RTOL = 1e-12


def _as_r4v4(m, r, v):
    n = len(m)
    r4 = np.zeros((n, 4)); r4[:, :3] = r; r4[:, 3] = m
    v4 = np.zeros((n, 4)); v4[:, :3] = v
    v4[:, 3] = 123.0   # the fourth velocity slot is unused by the reference's layout: it must not matter
    return r4, v4


def _noalloc_energy(lib, torch, m, r, v, stream=None):
    """What a caller without allocation rights does: its own device buffers in, two doubles out."""
    r4, v4 = _as_r4v4(m, r, v)
    d_r4 = torch.from_numpy(r4).cuda()
    d_v4 = torch.from_numpy(v4).cuda()
    d_out = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    rc = lib.gdy_energy_noalloc(d_r4.data_ptr(), d_v4.data_ptr(), len(m), d_out.data_ptr(),
                                stream.cuda_stream if stream is not None else None)
    assert rc == 0, lib.gdy_last_error(None).decode()
    if stream is not None:
        stream.synchronize()
    else:
        torch.cuda.synchronize()
    ek, ep = d_out.cpu().numpy()
    return float(ek + ep), float(ek), float(ep)


def _context_energy(m, r, v):
    from gravidy_b200.hermite4 import Hermite4B200
    h = Hermite4B200(m, r, v)
    try:
        return h.get_energy()
    finally:
        h.close()


def test_noalloc_energy_matches_context_energy_on_calls_in_a_row():
    torch = pytest.importorskip("torch")
    lib = capi.load()
    torch.cuda.set_device(0)
    # different sizes and mass spectra back to back: 5000 -> 777 leaves most of the first call's partial sums in the
    # shared array while the second call runs; 33000 uses j-splits > 1 chunks > 1; the last repeats the first
    cases = [(5000, 21, "salpeter"), (777, 5, "equal"), (33000, 8, "salpeter"), (5000, 21, "salpeter")]
    got = []
    for n, seed, spectrum in cases:
        m, r, v = plummer(n, seed=seed, mass_spectrum=spectrum)
        e_na = _noalloc_energy(lib, torch, m, r, v)
        e_ctx = _context_energy(m, r, v)
        assert np.all(np.isfinite(e_na)), (n, e_na)
        np.testing.assert_allclose(e_na, e_ctx, rtol=ENERGY_RTOL, err_msg=f"n={n}")
        got.append(e_na)
    assert got[0] == got[3]   # same input, same decomposition, same bits, whatever ran in between


def test_noalloc_energy_on_a_caller_stream_and_free_memory_untouched():
    torch = pytest.importorskip("torch")
    lib = capi.load()
    torch.cuda.set_device(0)
    m, r, v = plummer(4099, seed=3, mass_spectrum="salpeter")
    st = torch.cuda.Stream()
    first = _noalloc_energy(lib, torch, m, r, v, stream=st)   # set-up (address look-up, kernel load) happens here at the latest
    r4, v4 = _as_r4v4(m, r, v)
    d_r4, d_v4 = torch.from_numpy(r4).cuda(), torch.from_numpy(v4).cuda()
    d_out = torch.zeros(2, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        rc = lib.gdy_energy_noalloc(d_r4.data_ptr(), d_v4.data_ptr(), len(m), d_out.data_ptr(), st.cuda_stream)
        assert rc == 0, lib.gdy_last_error(None).decode()
    st.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "the device's free memory changed across gdy_energy_noalloc calls"
    ek, ep = d_out.cpu().numpy()
    assert (float(ek + ep), float(ek), float(ep)) == first
    np.testing.assert_allclose(first, _context_energy(m, r, v), rtol=ENERGY_RTOL)


def test_noalloc_energy_at_the_maximum_size():
    """n = GDY_NOALLOC_MAX_N is the largest set the fixed array is laid out for: it must be accepted and right."""
    torch = pytest.importorskip("torch")
    lib = capi.load()
    torch.cuda.set_device(0)
    m, r, v = plummer(capi.NOALLOC_MAX_N, seed=11, mass_spectrum="equal")
    e_na = _noalloc_energy(lib, torch, m, r, v)
    np.testing.assert_allclose(e_na, _context_energy(m, r, v), rtol=ENERGY_RTOL)


def test_noalloc_energy_refuses_what_does_not_fit():
    torch = pytest.importorskip("torch")
    lib = capi.load()
    torch.cuda.set_device(0)
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    rc = lib.gdy_energy_noalloc(d.data_ptr(), d.data_ptr(), capi.NOALLOC_MAX_N + 1, out.data_ptr(), None)
    assert rc == capi.GDY_ERR_INVALID
    assert b"GDY_NOALLOC_MAX_N" in lib.gdy_last_error(None)
    assert lib.gdy_energy_noalloc(d.data_ptr(), d.data_ptr(), 0, out.data_ptr(), None) == capi.GDY_ERR_INVALID
    assert lib.gdy_energy_noalloc(None, d.data_ptr(), 2, out.data_ptr(), None) == capi.GDY_ERR_INVALID
    torch.cuda.synchronize()
    assert torch.all(out == 0).item()   # a refused call launches nothing
