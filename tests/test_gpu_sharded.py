"""GPU suite: the j-sharded path.

* logical shards on ONE GPU (no communicator): the partial forces of the G shards, added on the host,
  equal the unsharded result — the same kernels and launch logic the multi-GPU run uses;
* two logical ranks on ONE GPU (GDY_PS_GRID = half the SMs each, same-device peer attach, no NCCL): the
  persistent kernel's NVLink-style flagged-slot exchange and the mid-block peer exchange run for real, two
  host threads, two cooperative kernels side by side; replicated state bit-identical on both ranks and the
  schedule of the one-GPU run.  Also the watchdog: a rank whose peer never shows up gets an error, not a hang;
* with >= 2 GPUs: two real shards in one process (two threads), NCCL allreduce inside
  update_acc_jrk / gdy_run; replicated state must stay BIT-identical on both ranks and follow the
  single-GPU schedule.
"""
import threading

import numpy as np
import pytest

from gravidy_b200.plummer import plummer

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def H():
    from gravidy_b200.hermite4 import Hermite4B200
    return Hermite4B200


def rel_err(a, b):
    da = np.linalg.norm(a[:, :3] - b[:, :3], axis=1) / np.linalg.norm(b[:, :3], axis=1)
    dj = np.linalg.norm(a[:, 3:] - b[:, 3:], axis=1) / np.linalg.norm(b[:, 3:], axis=1)
    return max(float(da.max()), float(dj.max()))


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("n", [200, 5000])
def test_logical_shards_sum_to_unsharded(H, n, world):
    m, r, v = plummer(n, seed=4, mass_spectrum="salpeter")
    full = H(m, r, v)
    full.init_acc_jrk()
    ref_all = full.download_forces()
    rng = np.random.default_rng(2)
    for nact in (1, 7, 100, n):
        move = np.sort(rng.choice(n, size=nact, replace=False)).astype(np.uint32)
        ref, _ = full.force_sweep(move)
        assert rel_err(ref, ref_all[move]) < 1e-12
        tot = np.zeros((nact, 6))
        for rank in range(world):
            h = H(m, r, v, rank=rank, world=world, nccl_uid=None)   # logical shard
            h.init_acc_jrk()                                        # partial forces of this shard
            part, _ = h.force_sweep(move)
            tot += part
            h.close()
        assert rel_err(tot, ref) < 1e-12, (n, world, nact)


def _two_gpu_available():
    import torch
    return torch.cuda.is_available() and torch.cuda.device_count() >= 2


@pytest.mark.skipif(not _two_gpu_available(), reason="needs 2 GPUs")
@pytest.mark.parametrize("peers,xmid", [(False, None), (True, "1"), (True, "0")])
def test_two_real_shards_nccl(H, peers, xmid, monkeypatch):
    """peers = False: every block step through the per-hook kernels + ncclAllReduce.  peers = True: small
    blocks through the persistent kernel with the fused NVLink exchange (peer memory); the blocks above its cap
    through the peer-memory exchange fused into the corrector (xmid = "1") or through ncclAllReduce ("0")."""
    if xmid is not None:
        monkeypatch.setenv("GDY_XMID", xmid)
    n, nsteps = 3000, 300 if peers else 60
    m, r, v = plummer(n, seed=8)
    single = H(m, r, v)
    e_init1 = single.get_energy()
    single.init_acc_jrk()
    at1 = single.init_dt(0.0)
    st1, tn1, ti1 = single.run(at1, 1e9, max_steps=nsteps, trace_cap=nsteps)
    s1 = single.get_state()

    uid = H.nccl_unique_id()
    out, err = [None, None], []
    handles = [None, None]
    gate = threading.Barrier(2)

    def work(rank):
        try:
            h = H(m, r, v, device=rank, rank=rank, world=2, nccl_uid=uid)
            if peers:
                handles[rank] = h.peer_export()
                gate.wait(timeout=120)
                h.peer_attach(handles)
                gate.wait(timeout=120)
            e_init = h.get_energy()      # same state as the one-GPU run -> must be the same BITS (fixed-order sums)
            h.init_acc_jrk()
            f0 = h.download_forces()
            at = h.init_dt(0.0)
            st, tn, ti = h.run(at, 1e9, max_steps=nsteps, trace_cap=nsteps)
            out[rank] = (f0, at, st.atime, tn, ti, h.get_state(), h.persist_stats(), h.get_energy(), e_init)
            if peers:
                gate.wait(timeout=120)   # nobody unmaps its buffer while a peer may still store into it
            h.close()
        except Exception as e:   # noqa: BLE001
            err.append(e)

    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
    assert not err, err
    a, b = out
    assert np.array_equal(a[0], b[0])                     # allreduce result bit-identical on both ranks
    assert a[1] == b[1] == at1
    assert np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])
    for k in a[5]:
        if k == "p7":
            continue   # a rank predicts its own j-slice plus the active particles only
        assert np.array_equal(a[5][k], b[5][k]), k        # replicated state stays bit-identical
    # ...and the predicted set agrees where both ranks hold a current prediction: each rank's slice
    # is at the last ITIME on that rank; check the two slices against the single-GPU prediction
    for rank in range(2):
        lo, hi = H.shard_range(n, rank, 2)
        assert np.array_equal(out[rank][5]["p7"][lo:hi, 6], s1["p7"][lo:hi, 6])
    if peers:
        assert a[6][1] > 0 and a[6] == b[6]                # the persistent kernel did steps, the same on both
    else:
        assert a[6][1] == 0
    # and follows the single-GPU run
    assert np.array_equal(a[3], tn1) and np.array_equal(a[4], ti1) and a[2] == st1.atime
    np.testing.assert_allclose(a[5]["r4"], s1["r4"], rtol=0, atol=1e-11)
    assert np.array_equal(a[5]["dt"], s1["dt"])
    # energy: i-chunks per rank + exchange of the chunk sums -> the same bits on both ranks (fixed-order summation)
    assert a[7] == b[7]
    assert a[8] == b[8] == e_init1                         # i-sharded energy: the bits of the one-GPU sum
    np.testing.assert_allclose(a[7], single.get_energy(), rtol=1e-11)


def _run_two_logical_ranks(H, m, r, v, nsteps, nact_max=None, only_rank=None, t_end=1e9):
    """Two contexts (rank 0/1 of 2) on device 0, peers attached through the same address space, one host thread
    each.  only_rank: the other rank initialises but never steps (watchdog test)."""
    out, err = [None, None], [None, None]
    handles = [None, None]
    gate = threading.Barrier(2)

    def work(rank):
        h = None
        try:
            h = H(m, r, v, device=0, rank=rank, world=2, nccl_uid=None)
            handles[rank] = h.peer_export()
            gate.wait(timeout=120)
            h.peer_attach(handles)
            if nact_max is not None:
                h.set_persist(nact_max)
            gate.wait(timeout=120)
            h.init_acc_jrk()             # <= 16384 particles: the world's sums through the peer exchange
            f0 = h.download_forces()
            at = h.init_dt(0.0)
            gate.wait(timeout=120)
            if only_rank is None or rank == only_rank:
                st, tn, ti = h.run(at, t_end, max_steps=nsteps, trace_cap=nsteps)
                out[rank] = (f0, at, st.atime, tn, ti, h.get_state(), h.persist_stats())
        except Exception as e:   # noqa: BLE001
            err[rank] = e
        finally:
            try:
                gate.wait(timeout=120)   # nobody frees its buffer while the peer may still store into it
            except threading.BrokenBarrierError:
                pass
            if h is not None:
                h.close()

    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=600)
    return out, err


@pytest.mark.parametrize("n,nsteps,nact_max", [(3000, 500, None), (3000, 300, 2048), (12000, 250, None), (640, 400, 64),
                                               (12000, 250, 2048)])   # > 16 owned particles per CTA: send-all-then-receive
def test_two_logical_ranks_one_device(H, monkeypatch, n, nsteps, nact_max):
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    monkeypatch.setenv("GDY_PS_GRID", str(sms // 2))
    m, r, v = plummer(n, seed=8)
    single = H(m, r, v)
    single.init_acc_jrk()
    f1 = single.download_forces()
    at1 = single.init_dt(0.0)
    st1, tn1, ti1 = single.run(at1, 1e9, max_steps=nsteps, trace_cap=nsteps)
    s1 = single.get_state()
    single.close()

    out, err = _run_two_logical_ranks(H, m, r, v, nsteps, nact_max)
    assert err == [None, None], err
    a, b = out
    assert np.array_equal(a[0], b[0])                      # exchanged sums: the same bits on both ranks
    assert rel_err(a[0], f1) < 1e-12
    assert a[1] == b[1] == at1
    assert np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])
    for k in a[5]:
        if k == "p7":
            continue   # a rank predicts its own j-slice plus the active particles only
        assert np.array_equal(a[5][k], b[5][k]), k         # replicated state stays bit-identical
    assert a[6][1] > 0 and a[6] == b[6]                    # the persistent kernel did steps, the same on both
    assert np.array_equal(a[3], tn1) and np.array_equal(a[4], ti1) and a[2] == st1.atime   # the one-GPU schedule
    np.testing.assert_allclose(a[5]["r4"], s1["r4"], rtol=0, atol=1e-11)
    assert np.array_equal(a[5]["dt"], s1["dt"]) and np.array_equal(a[5]["t"], s1["t"])


def test_missing_peer_is_an_error_not_a_hang(H, monkeypatch):
    """Rank 1 attaches but never steps: rank 0's exchange poll must give up (GDY_SPIN_TIMEOUT_MS) and gdy_run must
    return an error (the reference's convention is to fail loudly: src/include/gpu/Hermite4GPU.cuh:43-50)."""
    import time
    import torch
    from gravidy_b200 import capi
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    monkeypatch.setenv("GDY_PS_GRID", str(sms // 2))
    monkeypatch.setenv("GDY_SPIN_TIMEOUT_MS", "300")
    m, r, v = plummer(2000, seed=3)
    t0 = time.perf_counter()
    out, err = _run_two_logical_ranks(H, m, r, v, 50, only_rank=0)
    assert time.perf_counter() - t0 < 60
    assert err[1] is None and isinstance(err[0], capi.GdyError), err
    assert "watchdog" in str(err[0])


def test_missing_rank_at_create_is_an_error_not_a_hang():
    """Rank 0 of a world of two creates its context, rank 1 never does: ncclCommInitRank would block for ever;
    gdy_create_sharded gives up after GDY_NCCL_INIT_TIMEOUT_MS with GDY_ERR_NCCL.  In a process of its own: the helper
    thread that is left inside NCCL must not outlive anything but that process."""
    import os
    import subprocess
    import sys
    code = (
        "import os, sys, time\n"
        "sys.path.insert(0, %r)\n"
        "os.environ['GDY_NCCL_INIT_TIMEOUT_MS'] = '3000'\n"
        "import numpy as np\n"
        "from gravidy_b200 import capi\n"
        "from gravidy_b200.hermite4 import Hermite4B200\n"
        "from gravidy_b200.plummer import plummer\n"
        "m, r, v = plummer(512, seed=1)\n"
        "uid = Hermite4B200.nccl_unique_id()\n"
        "t0 = time.perf_counter()\n"
        "try:\n"
        "    Hermite4B200(m, r, v, device=0, rank=0, world=2, nccl_uid=uid)\n"
        "    print('CREATED')\n"
        "except capi.GdyError as e:\n"
        "    print('ERROR after %%.1f s: %%s' %% (time.perf_counter() - t0, e))\n"
        "sys.stdout.flush()\n"
        "os._exit(0)\n" % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert "ERROR after" in p.stdout and "waited 3000 ms for the other ranks" in p.stdout, (p.stdout, p.stderr[-2000:])
