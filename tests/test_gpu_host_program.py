"""GPU suite: the C++ host program (gravidy_b200/host/gravidy-b200) against the output FILES of the
unmodified reference driver (tests/golden/ref_run_n64, made by tests/golden/make_golden_run.py):
same command line, same .info / .log / .snapshot_NN / .all.snap_NNNN formats and contents, same block
schedule (Iter and Nsteps columns), in both loop modes (fused device driver and --hooks)."""
import os
import shutil
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "gravidy_b200", "host", "gravidy-b200")
REF = os.path.join(ROOT, "tests", "golden", "ref_run_n64")


@pytest.fixture(scope="module")
def binary():
    from gravidy_b200 import build as B
    from gravidy_b200.host import build as HB
    B.build()
    return HB.build()


def run_host(binary, tmp_path, extra):
    shutil.copy(os.path.join(REF, "input.dat"), tmp_path / "input.dat")
    p = subprocess.run([binary, "-i", "./input.dat", "-o", "./run", "-t", "0.5", "-a"] + extra,
                       cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert "[SUCCESS]" in p.stderr and "Finishing..." in p.stderr
    return p


def same_layout(a, b):
    """same number of lines, same line lengths, same positions of blanks (fixed-width columns)"""
    la, lb = a.splitlines(), b.splitlines()
    assert len(la) == len(lb)
    for x, y in zip(la, lb):
        assert len(x) == len(y), (x, y)
        assert [c == " " for c in x] == [c == " " for c in y], (x, y)


def numbers(text):
    return np.array([[float(t) for t in ln.split()] for ln in text.splitlines() if not ln.startswith("#")])


@pytest.mark.parametrize("mode", [[], ["--hooks"]])
def test_output_files_match_reference(binary, tmp_path, mode):
    run_host(binary, tmp_path, mode)
    ref_files = sorted(f for f in os.listdir(REF) if f.startswith("run.out"))
    got_files = sorted(f for f in os.listdir(tmp_path) if f.startswith("run.out"))
    assert got_files == ref_files
    for f in ref_files:
        ref = open(os.path.join(REF, f)).read()
        got = open(tmp_path / f).read()
        if f.endswith(".info") or f.endswith("snapshot_00"):
            assert got == ref, f                                     # byte-identical
        elif f.endswith(".log"):
            same_layout(got, ref)
            assert got.splitlines()[0] == ref.splitlines()[0]        # header row
            g, r = numbers(got), numbers(ref)
            assert np.array_equal(g[:, :4], r[:, :4])                # tag, IteTime, Iter, Nsteps: exact
            np.testing.assert_allclose(g[:, 4], r[:, 4], rtol=1e-9)  # Energy
            assert (g[1:, 6] < 1e-6).all() and (r[1:, 6] < 1e-6).all()   # CumE: same regime
        else:                                                        # snapshots and all-particle dumps
            same_layout(got, ref)
            assert got.splitlines()[0][:7] == ref.splitlines()[0][:7]
            np.testing.assert_allclose(numbers(got), numbers(ref), rtol=5e-6, atol=1e-12)


def test_fused_and_hook_modes_write_identical_files(binary, tmp_path):
    a, b = tmp_path / "a", tmp_path / "b"
    a.mkdir(); b.mkdir()
    run_host(binary, a, [])
    run_host(binary, b, ["--hooks"])
    for f in sorted(os.listdir(a)):
        if not f.startswith("run.out"):
            continue
        x, y = open(a / f).read(), open(b / f).read()
        if f.endswith(".log"):   # ElapsedTime and GFLOPS columns are wall-clock
            x = "\n".join(ln[:92] for ln in x.splitlines())
            y = "\n".join(ln[:92] for ln in y.splitlines())
        assert x == y, f


def test_lagrange_radii_and_resume(binary, tmp_path):
    shutil.copy(os.path.join(REF, "input.dat"), tmp_path / "input.dat")
    p = subprocess.run([binary, "-i", "./input.dat", "-t", "0.25", "-l"], cwd=tmp_path,
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    radii = [ln.split() for ln in open(tmp_path / "input.dat.out.radii").read().splitlines()]
    assert len(radii) == 3 and all(r[0] == "01" and len(r) == 8 for r in radii)
    vals = np.array([[float(t) for t in r[2:]] for r in radii])
    assert (np.diff(vals, axis=1) >= 0).all() and (vals > 0).all()   # radii grow with the mass fraction
    # against the file the unmodified reference driver writes with -l for the same run (tests/golden/
    # make_golden_radii.py; NbodyUtils::lagrange_radii, reference: src/include/utils/NbodyUtils.cpp:262-328): same
    # times, radii equal to the six printed decimals (one unit of the last digit for a rounding flip)
    ref = [ln.split() for ln in open(os.path.join(ROOT, "tests", "golden", "ref_run_n64_radii", "run.out.radii")).read().splitlines()]
    ref_vals = np.array([[float(t) for t in r[1:]] for r in ref[:3]])
    got_vals = np.array([[float(t) for t in r[1:]] for r in radii])
    assert np.array_equal(got_vals[:, 0], ref_vals[:, 0])
    np.testing.assert_allclose(got_vals[:, 1:], ref_vals[:, 1:], rtol=0, atol=1.01e-6)
    full = np.load(os.path.join(ROOT, "tests", "golden", "ref_run_n64_radii", "ref_radii.npz"))
    np.testing.assert_allclose(got_vals[0, 1:], full["lagrange_t0"], rtol=0, atol=5.01e-7)   # t = 0: the rounded reference values
    # resume: needs a longer -t than the info file, restarts from the last snapshot
    p = subprocess.run([binary, "-r", "./input.dat.out.info", "-t", "0.25"], cwd=tmp_path,
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 1 and "greater than the one specifid in the info file" in p.stderr
    p = subprocess.run([binary, "-r", "./input.dat.out.info", "-t", "0.375"], cwd=tmp_path,
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert os.path.exists(tmp_path / "input.dat.out.info.old")
    log = open(tmp_path / "input.dat.out.log").read().splitlines()
    assert log[1].split()[1] == "0.250" and log[-1].split()[1] == "0.375"


CKPT_HEADER = np.dtype([("magic", "S8"), ("n", "<u4"), ("snap_number", "<u4"), ("output_factor", "<u4"),
                        ("iterations", "<u4"), ("nsteps", "<u4"), ("reserved", "<u4"), ("interactions", "<i8"),
                        ("e2", "<f8"), ("eta", "<f4"), ("reserved2", "<f4"), ("ITIME", "<f8"), ("ATIME", "<f8"),
                        ("e_ini", "<f8"), ("e_tmp", "<f8"), ("update_end", "<f8")])


def read_ckpt(path):
    raw = open(path, "rb").read()
    h = np.frombuffer(raw[:CKPT_HEADER.itemsize], dtype=CKPT_HEADER)[0]
    n = int(h["n"])
    body = np.frombuffer(raw[CKPT_HEADER.itemsize:], dtype="<f8")
    assert body.size == n * (4 + 4 + 6 + 1 + 1)
    o = np.cumsum([0, 4 * n, 4 * n, 6 * n, n, n])
    r4, v4, f6, t, dt = (body[o[k]:o[k + 1]] for k in range(5))
    return h, r4.reshape(n, 4), v4.reshape(n, 4), f6.reshape(n, 6), t, dt


@pytest.mark.parametrize("mode", [[], ["--hooks"]])
def test_checkpoint_restart_continues_the_schedule_bit_for_bit(binary, tmp_path, mode):
    """run to t=0.5 straight, and to t=0.25 + --restart to t=0.5: the full-precision state
    (r, v, acc, jerk, t, dt), every counter, the next block time and the text outputs are identical."""
    a, b = tmp_path / "a", tmp_path / "b"
    a.mkdir(); b.mkdir()
    for d in (a, b):
        shutil.copy(os.path.join(REF, "input.dat"), d / "input.dat")
    common = [binary, "-i", "./input.dat", "-o", "./run", "--checkpoint"] + mode
    p = subprocess.run(common + ["-t", "0.5"], cwd=a, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    p = subprocess.run(common + ["-t", "0.25"], cwd=b, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    half = read_ckpt(b / "run.out.ckpt")
    assert half[0]["magic"] == b"GDYCKPT2" and half[0]["ITIME"] == 0.25 and half[0]["n"] == 64
    shutil.copy(b / "run.out.ckpt", b / "half.ckpt")
    p = subprocess.run(common + ["-t", "0.5", "--restart", "./half.ckpt"], cwd=b, capture_output=True, text=True,
                       timeout=300)
    assert p.returncode == 0, p.stderr
    ha, hb = read_ckpt(a / "run.out.ckpt"), read_ckpt(b / "run.out.ckpt")
    for name in CKPT_HEADER.names:
        if name != "update_end":   # wall-clock seconds spent in force evaluations
            assert ha[0][name] == hb[0][name], name
    # the force time carries over the restart, so the GFLOPS column stays interactions / force time (never inf)
    assert hb[0]["update_end"] > half[0]["update_end"] > 0.0
    for ln in open(b / "run.out.log").read().splitlines()[1:]:
        assert np.isfinite(float(ln.split()[-1])), ln
    assert ha[0]["ITIME"] == 0.5 and ha[0]["iterations"] > half[0]["iterations"]
    for x, y in zip(ha[1:], hb[1:]):
        assert np.array_equal(x, y)
    # the snapshots of the second half and the final one are the same files
    for f in ("run.out.snapshot_03", "run.out.snapshot_04", "run.out.snapshot_05"):
        assert open(a / f).read() == open(b / f).read(), f
    # energy log: the lines of t = 0.375 and 0.5 agree in everything but the wall-clock columns
    la = [ln[:92] for ln in open(a / "run.out.log").read().splitlines()]
    lb = [ln[:92] for ln in open(b / "run.out.log").read().splitlines()]
    assert la[-2:] == lb[-2:]
    # a checkpoint of another system, or one that is already past -t, is refused
    p = subprocess.run(common + ["-t", "0.25", "--restart", "./half.ckpt"], cwd=b, capture_output=True, text=True)
    assert p.returncode != 0 and "already at or past" in p.stderr
    p = subprocess.run(common + ["-t", "0.5", "--restart", "./input.dat"], cwd=b, capture_output=True, text=True)
    assert p.returncode != 0 and "not a gravidy-b200 checkpoint" in p.stderr
    # a checkpoint of an earlier layout is named as such (round-2 advisor finding: it used to be "not a checkpoint")
    (b / "v1.ckpt").write_bytes(b"GDYCKPT1" + bytes(4096))
    p = subprocess.run(common + ["-t", "0.5", "--restart", "./v1.ckpt"], cwd=b, capture_output=True, text=True)
    assert p.returncode != 0 and "format version 1" in p.stderr and "reads version 2" in p.stderr


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_gpu_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_two_gpu_host_program_writes_the_same_files(binary, tmp_path):
    """-g 2: one thread per GPU inside the host program, j-sharded, NCCL for big blocks and the fused NVLink
    exchange (same-process peer access) for small ones.  Same schedule and the same text outputs as -g 1
    (forces differ by summation order only, far below the printed digits), and a checkpoint written on two
    GPUs restarts on one."""
    a, b = tmp_path / "a", tmp_path / "b"
    a.mkdir(); b.mkdir()
    run_host(binary, a, ["-g", "1", "--checkpoint"])
    run_host(binary, b, ["-g", "2", "--checkpoint"])
    for f in sorted(os.listdir(a)):
        if not f.startswith("run.out") or f.endswith(".ckpt"):
            continue
        x, y = open(a / f).read(), open(b / f).read()
        if f.endswith(".log"):
            gx, gy = numbers(x), numbers(y)
            assert np.array_equal(gx[:, :4], gy[:, :4])                 # tag, time, Iter, Nsteps: exact
            np.testing.assert_allclose(gx[:, 4], gy[:, 4], rtol=1e-12)  # energy
        elif f.endswith(".info"):
            assert x == y
        else:
            same_layout(x, y)
            np.testing.assert_allclose(numbers(x), numbers(y), rtol=1e-9, atol=1e-14)
    ha, hb = read_ckpt(a / "run.out.ckpt"), read_ckpt(b / "run.out.ckpt")
    for name in ("n", "snap_number", "output_factor", "iterations", "nsteps", "interactions", "ITIME", "ATIME"):
        assert ha[0][name] == hb[0][name], name
    assert np.array_equal(ha[4], hb[4]) and np.array_equal(ha[5], hb[5])   # t, dt: the schedule
    np.testing.assert_allclose(ha[1], hb[1], rtol=1e-10, atol=1e-14)
    # restart the 2-GPU checkpoint on one GPU
    shutil.copy(b / "run.out.ckpt", b / "two.ckpt")
    p = subprocess.run([binary, "-i", "./input.dat", "-o", "./run", "-t", "0.625", "-g", "1", "--restart", "./two.ckpt"],
                       cwd=b, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    log = open(b / "run.out.log").read().splitlines()
    assert log[-1].split()[1] == "0.625"


def test_cli_errors(binary, tmp_path):
    p = subprocess.run([binary], capture_output=True, text=True)
    assert p.returncode == 1 and "option requires an argument -- 'input'" in p.stderr
    p = subprocess.run([binary, "-i", "/no/such/file"], capture_output=True, text=True)
    assert p.returncode == 1 and "cannot access /no/such/file: No such file or directory" in p.stderr
    p = subprocess.run([binary, "--help"], capture_output=True, text=True)
    assert p.returncode == 1 and "Required options" in p.stderr
