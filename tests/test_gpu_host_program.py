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


def test_cli_errors(binary, tmp_path):
    p = subprocess.run([binary], capture_output=True, text=True)
    assert p.returncode == 1 and "option requires an argument -- 'input'" in p.stderr
    p = subprocess.run([binary, "-i", "/no/such/file"], capture_output=True, text=True)
    assert p.returncode == 1 and "cannot access /no/such/file: No such file or directory" in p.stderr
    p = subprocess.run([binary, "--help"], capture_output=True, text=True)
    assert p.returncode == 1 and "Required options" in p.stderr
