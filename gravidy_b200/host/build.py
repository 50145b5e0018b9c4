"""Build the C++ host program ``gravidy_b200/host/gravidy-b200`` (g++; links libgravidy_b200.so).

The host side contains no CUDA: it reaches the GPU only through the C ABI in include/gravidy_b200.h.
"""
from __future__ import annotations

import glob
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "gravidy-b200")
LIBDIR = os.path.dirname(HERE)


def stale() -> bool:
    if not os.path.exists(BIN):
        return True
    t = os.path.getmtime(BIN)
    deps = glob.glob(os.path.join(HERE, "*.cpp")) + glob.glob(os.path.join(HERE, "*.hpp")) + \
        [os.path.join(LIBDIR, "libgravidy_b200.so")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False) -> str:
    if not force and not stale():
        return BIN
    srcs = sorted(glob.glob(os.path.join(HERE, "*.cpp")))
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-pthread", "-o", BIN] + srcs + \
          ["-L" + LIBDIR, "-lgravidy_b200", "-Wl,-rpath,$ORIGIN/..", "-Wl,-rpath," + LIBDIR]
    subprocess.run(cmd, check=True, cwd=HERE)
    return BIN


if __name__ == "__main__":
    print(build(force=True))
