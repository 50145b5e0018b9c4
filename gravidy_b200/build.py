"""Build the CUDA library in-tree: ``gravidy_b200/libgravidy_b200.so`` (sm_100a only).

``python -m gravidy_b200.build`` or ``__graft_entry__.build()``.  nvcc cross-compiles without a GPU.
The built file is git-ignored but travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgravidy_b200.so")
SOURCES = ["gdy_ctx.cu"]
HEADERS = ["gdy_device.cuh", "gdy_force.cuh", "gdy_step.cuh", "gdy_persist.cuh", "gdy_persist2.cuh",
           os.path.join("..", "..", "include", "gravidy_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-O2",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (needed to build gravidy_b200/libgravidy_b200.so)")


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


# The force kernel's speed AND correctness lean on how ptxas 12.9 treats its pinned instruction groups
# (csrc/gdy_force.cuh: never-taken gotos; one nvcc 12.9 code-generation bug is worked around there), so the toolkit
# the numbers were measured with is named here and the generated SASS is checked after every compile.
NVCC_VERIFIED = "12.9"
FORCE_KERNELS = ("k_force_warpILi2ELi1ELi4ELi3ELb0", "k_force_warpILi2ELi4ELi1ELi3ELb0")   # warp_i64, warp_i256


def nvcc_release() -> str:
    out = subprocess.run([_nvcc(), "--version"], capture_output=True, text=True, check=True).stdout
    m = re.search(r"release (\d+\.\d+)", out)
    return m.group(1) if m else "unknown"


def check_force_sass(lib: str = LIB) -> list:
    """Codegen guard, no GPU needed: reads the SASS of the big-block force kernels in `lib` (cuobjdump) and returns the
    list of broken expectations (empty = as measured): 31 FP64-pipe instructions per pair, 11 of them with three
    register sources of which at least 6 are fed by the operand reuse cache, loop addressing in uniform registers,
    and the TMA bulk copy / mbarrier / MUFU.RSQ64H instructions present."""
    sys.path.insert(0, os.path.join(HERE, "..", "tools"))
    try:
        import sass_fp64_cost as S
    finally:
        sys.path.pop(0)
    bad = []
    fns = S.functions(lib)
    for inst in FORCE_KERNELS:
        name = [k for k in fns if inst in k]
        if len(name) != 1:
            bad.append(f"{inst}: {len(name)} instantiations in the library")
            continue
        seg = S.hottest_loop(fns[name[0]])
        total, n64, n3, fed = S.cost(seg)
        if n64 % 31 or n64 < 124:
            bad.append(f"{inst}: {n64} FP64-pipe instructions in the hot loop (31 per pair, >= 4 pairs per trip)")
            continue
        pairs = n64 // 31
        if n3 != 11 * pairs:
            bad.append(f"{inst}: {n3} three-source FP64 instructions for {pairs} pairs (11 per pair)")
        if fed < 6 * pairs:
            bad.append(f"{inst}: only {fed} three-source DFMAs fed by the reuse cache for {pairs} pairs (>= 6 per pair)")
        if total / pairs >= 72.6:
            bad.append(f"{inst}: {total / pairs:.1f} register-read slots per pair (nothing-fed bound 72.6)")
        if S.other_reg_readers(seg) > pairs:
            bad.append(f"{inst}: {S.other_reg_readers(seg)} other vector-register readers in the hot loop (<= {pairs})")
        if not any(re.match(r"LDS\.128 R\d+, \[UR\d+", i) for _, i in seg):
            bad.append(f"{inst}: the hot loop's loads no longer use a uniform address")
        txt = "\n".join(i for _, i in fns[name[0]])
        for need in ("UBLKCP", "MUFU.RSQ64H", "SYNCS"):
            if need not in txt:
                bad.append(f"{inst}: no {need} in the SASS")
    return bad


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"]
    subprocess.run(cmd, check=True, cwd=CSRC)
    # a toolkit that schedules the force kernel differently must not ship silently (GDY_ALLOW_UNVERIFIED_SASS=1 to
    # build anyway, e.g. to measure what the new compiler does)
    rel = nvcc_release()
    if shutil.which("cuobjdump") or os.path.exists("/usr/local/cuda/bin/cuobjdump"):
        bad = check_force_sass(LIB)
        if bad and os.environ.get("GDY_ALLOW_UNVERIFIED_SASS") != "1":
            os.replace(LIB, LIB + ".rejected")
            raise RuntimeError(f"nvcc {rel} (verified: {NVCC_VERIFIED}) generated a force kernel that misses its "
                               "instruction budget; the library was moved to " + LIB + ".rejected:\n  " + "\n  ".join(bad))
    elif rel != NVCC_VERIFIED:
        print(f"gravidy_b200.build: nvcc {rel} is not the verified release {NVCC_VERIFIED} and cuobjdump is missing: "
              "the force kernel's SASS was NOT checked (run tests/test_capi_cpu.py and the every-j-position GPU test)",
              file=sys.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
