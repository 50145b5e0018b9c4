"""TEST INFRASTRUCTURE ONLY — ctypes front end to the two CPU checkers.

* ``RefOracle``  : the UNMODIFIED reference Hermite4CPU path compiled from /root/reference into
  ``oracle/_ref/libgravidy_ref.so`` (``oracle/ref_harness.cpp``).  ``kind = "reference"``.
* ``PortOracle`` : the plain-C restatement ``oracle/hermite_oracle.c`` → ``liboracle_port.so``.
  ``kind = "port"``.

Both expose the hooks of the reference loop (reference:
src/include/cpu/Hermite4CPU_integration.cpp:81-124) under the same method names.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may
import this module; the product package ``gravidy_b200`` never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_port.so")
REF_SO = os.path.join(HERE, "_ref", "libgravidy_ref.so")

_dp = C.POINTER(C.c_double)
_up = C.POINTER(C.c_uint32)


class RunStats(C.Structure):
    _fields_ = [
        ("iterations", C.c_uint64),
        ("nsteps", C.c_uint64),
        ("interactions", C.c_uint64),
        ("itime", C.c_double),
        ("atime", C.c_double),
        ("update_s", C.c_double),
        ("wall_s", C.c_double),
    ]


def build(force: bool = False) -> None:
    """Compile the checkers (``make -C oracle``).  Building the checker is not using it."""
    src_new = os.path.isdir("/root/reference/src") and (
        not os.path.exists(REF_SO)
        or os.path.getmtime(os.path.join(HERE, "ref_harness.cpp")) > os.path.getmtime(REF_SO))
    if force or not os.path.exists(PORT_SO) or src_new:
        subprocess.run(["make", "-C", HERE], check=True, stdout=subprocess.DEVNULL)


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def _ptr(a, ty):
    return None if a is None else a.ctypes.data_as(ty)


def write_input_file(path: str, m, r, v) -> None:
    """8-column text input ``id m x y z vx vy vz`` (reference reader: NbodySystem.cpp:161-180)."""
    n = len(m)
    with open(path, "w") as f:
        for i in range(n):
            f.write(
                "%d %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n"
                % (i, m[i], r[i, 0], r[i, 1], r[i, 2], v[i, 0], v[i, 1], v[i, 2])
            )


class _Base:
    """Shared hook surface; subclasses bind ``self.lib``, ``self.h`` and the symbol prefix."""

    kind = "?"
    _pfx = ""

    def _f(self, name):
        return getattr(self.lib, self._pfx + name)

    def _declare(self):
        L, p = self.lib, self._pfx
        vp = C.c_void_p
        g = lambda nm: getattr(L, p + nm)
        g("n").restype = C.c_uint32; g("n").argtypes = [vp]
        g("e2").restype = C.c_double; g("e2").argtypes = [vp]
        g("eta").restype = C.c_float; g("eta").argtypes = [vp]
        g("total_mass").restype = C.c_float; g("total_mass").argtypes = [vp]
        g("set_threads").argtypes = [C.c_int]
        g("max_threads").restype = C.c_int
        g("init_acc_jrk").argtypes = [vp]
        g("init_dt").restype = C.c_double; g("init_dt").argtypes = [vp, C.c_double]
        g("find_particles_to_move").restype = C.c_uint32
        g("find_particles_to_move").argtypes = [vp, C.c_double]
        g("save_old_acc_jrk").argtypes = [vp, C.c_uint32]
        g("predicted_pos_vel").argtypes = [vp, C.c_double]
        g("update_acc_jrk").argtypes = [vp, C.c_uint32]
        g("correction_pos_vel").argtypes = [vp, C.c_double, C.c_uint32]
        g("next_integration_time").restype = C.c_double
        g("next_integration_time").argtypes = [vp]
        g("get_energy").restype = C.c_double; g("get_energy").argtypes = [vp, _dp, _dp]
        g("get_timestep_normal").restype = C.c_double
        g("get_timestep_normal").argtypes = [vp, C.c_uint32, C.c_float]
        g("get_state").argtypes = [vp] + [_dp] * 9 + [_up]
        g("set_state").argtypes = [vp] + [_dp] * 9 + [_up]
        g("timed_update").restype = C.c_double
        g("timed_update").argtypes = [vp, _up, C.c_uint32]
        g("destroy").argtypes = [vp]

    # ---- hooks -----------------------------------------------------------------------------
    @property
    def n(self): return self._f("n")(self.h)
    @property
    def e2(self): return self._f("e2")(self.h)
    @property
    def eta(self): return self._f("eta")(self.h)
    @property
    def total_mass(self): return self._f("total_mass")(self.h)

    def set_threads(self, k): self._f("set_threads")(int(k))
    def max_threads(self): return self._f("max_threads")()
    def init_acc_jrk(self): self._f("init_acc_jrk")(self.h)
    def init_dt(self, itime=0.0): return self._f("init_dt")(self.h, itime)
    def find_particles_to_move(self, itime): return self._f("find_particles_to_move")(self.h, itime)
    def save_old_acc_jrk(self, nact): self._f("save_old_acc_jrk")(self.h, nact)
    def predicted_pos_vel(self, itime): self._f("predicted_pos_vel")(self.h, itime)
    def update_acc_jrk(self, nact): self._f("update_acc_jrk")(self.h, nact)
    def correction_pos_vel(self, itime, nact): self._f("correction_pos_vel")(self.h, itime, nact)
    def next_integration_time(self): return self._f("next_integration_time")(self.h)

    def get_energy(self):
        ek, ep = C.c_double(), C.c_double()
        e = self._f("get_energy")(self.h, C.byref(ek), C.byref(ep))
        return e, ek.value, ep.value

    def get_timestep_normal(self, i, eta): return self._f("get_timestep_normal")(self.h, i, eta)

    def timed_update(self, move):
        mv = np.ascontiguousarray(move, dtype=np.uint32)
        return self._f("timed_update")(self.h, _ptr(mv, _up), len(mv))

    _FIELDS = (("r4", 4), ("v4", 4), ("f6", 6), ("old6", 6), ("a2", 3), ("a3", 3), ("t", 0),
               ("dt", 0), ("p7", 7))

    def get_state(self, *names):
        """Copy out arrays by name: r4 v4 f6 old6 a2 a3 t dt p7 move (all if none named)."""
        n = self.n
        want = names or [f for f, _ in self._FIELDS] + ["move"]
        out, args = {}, []
        for f, w in self._FIELDS:
            if f in want:
                out[f] = np.zeros((n, w) if w else (n,), dtype=np.float64)
                args.append(_ptr(out[f], _dp))
            else:
                args.append(None)
        if "move" in want:
            out["move"] = np.zeros(n, dtype=np.uint32)
            args.append(_ptr(out["move"], _up))
        else:
            args.append(None)
        self._f("get_state")(self.h, *args)
        return out

    def set_state(self, **arrs):
        args, keep = [], []
        for f, _ in self._FIELDS:
            a = arrs.get(f)
            if a is not None:
                a = np.ascontiguousarray(a, dtype=np.float64); keep.append(a)
            args.append(_ptr(a, _dp))
        mv = arrs.get("move")
        if mv is not None:
            mv = np.ascontiguousarray(mv, dtype=np.uint32); keep.append(mv)
        args.append(_ptr(mv, _up))
        self._f("set_state")(self.h, *args)

    def run(self, atime, t_end, max_steps=0, trace_cap=0, itime0=0.0):
        st = RunStats()
        tn = np.zeros(max(trace_cap, 1), dtype=np.uint32)
        ti = np.zeros(max(trace_cap, 1), dtype=np.float64)
        self._run(atime, itime0, t_end, max_steps, st, tn, ti, trace_cap)
        k = min(trace_cap, st.iterations)
        return st, tn[:k], ti[:k]

    def close(self):
        if getattr(self, "h", None):
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PortOracle(_Base):
    kind = "port"
    _pfx = "go_"

    def __init__(self, m, r, v, softening=1e-4, eta=0.01):
        build()
        self.lib = C.CDLL(PORT_SO)
        self._declare()
        L = self.lib
        L.go_create.restype = C.c_void_p
        L.go_create.argtypes = [C.c_uint32, _dp, _dp, _dp, C.c_float, C.c_float]
        L.go_run.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_uint64,
                             C.POINTER(RunStats), _up, _dp, C.c_uint64]
        L.go_pair.argtypes = [_dp, _dp, C.c_double, _dp]
        L.go_normalize_dt.restype = C.c_double
        L.go_normalize_dt.argtypes = [C.c_double] * 3
        m = np.ascontiguousarray(m, dtype=np.float64)
        r = np.ascontiguousarray(r, dtype=np.float64)
        v = np.ascontiguousarray(v, dtype=np.float64)
        self.h = L.go_create(len(m), _ptr(m, _dp), _ptr(r, _dp), _ptr(v, _dp),
                             C.c_float(softening), C.c_float(eta))

    def normalize_dt(self, new_dt, old_dt, t):
        return self.lib.go_normalize_dt(new_dt, old_dt, t)

    def _run(self, atime, itime0, t_end, max_steps, st, tn, ti, cap):
        self.lib.go_run(self.h, atime, itime0, t_end, max_steps, C.byref(st), _ptr(tn, _up),
                        _ptr(ti, _dp), cap)


class RefOracle(_Base):
    kind = "reference"
    _pfx = "gref_"

    def __init__(self, m, r, v, softening=1e-4, eta=0.01, integration_time=1.0,
                 interval_time=0.125, workdir=None):
        if not have_ref():
            build()
        if not have_ref():
            raise RuntimeError("oracle/_ref/libgravidy_ref.so missing and /root/reference absent")
        self.lib = C.CDLL(REF_SO)
        self._declare()
        L = self.lib
        L.gref_create_from_file.restype = C.c_void_p
        L.gref_create_from_file.argtypes = [C.c_char_p, C.c_char_p, C.c_float, C.c_float,
                                            C.c_float, C.c_float]
        L.gref_run.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_uint64,
                               C.POINTER(RunStats), _up, _dp, C.c_uint64]
        L.gref_normalize_dt.restype = C.c_double
        L.gref_normalize_dt.argtypes = [C.c_void_p] + [C.c_double] * 3
        L.gref_integration.argtypes = [C.c_void_p]
        self._tmp = None
        if workdir is None:
            self._tmp = tempfile.TemporaryDirectory(prefix="gref_")
            workdir = self._tmp.name
        self.workdir = workdir
        self.input_path = os.path.join(workdir, "input.dat")
        self.output_prefix = os.path.join(workdir, "run")
        write_input_file(self.input_path, np.asarray(m), np.asarray(r), np.asarray(v))
        self.h = L.gref_create_from_file(self.input_path.encode(), self.output_prefix.encode(),
                                         C.c_float(softening), C.c_float(eta),
                                         C.c_float(integration_time), C.c_float(interval_time))

    def normalize_dt(self, new_dt, old_dt, t):
        return self.lib.gref_normalize_dt(self.h, new_dt, old_dt, t)

    def center_of_density(self):
        """NbodyUtils::get_center_of_density of the reference, single-threaded (its OpenMP loop races on shared
        temporaries).  Returns (x, y, z); J = self.J neighbours."""
        self.lib.gref_center_of_density.argtypes = [C.c_void_p, _dp]
        prev = self.lib.gref_max_threads()
        self.set_threads(1)
        out = np.zeros(3)
        self.lib.gref_center_of_density(self.h, _ptr(out, _dp))
        self.set_threads(prev)
        return out

    @property
    def J(self):
        return int(self.lib.gref_J())

    def lagrange_radii(self):
        """NbodyUtils::lagrange_radii of the reference on the current positions: the six Lagrange radii
        (1, 5, 10, 20, 50, 75 % of the mass; reference: src/include/utils/NbodyUtils.cpp:262-328)."""
        self.lib.gref_lagrange_radii.argtypes = [C.c_void_p, _dp]
        self.lib.gref_lagrange_radii.restype = None
        out = np.zeros(6)
        self.lib.gref_lagrange_radii(self.h, _ptr(out, _dp))
        return out

    def core_and_relaxation(self):
        """(r_core, t_rlx, r_hm): get_core_radius + get_half_mass_relaxation_time of the reference
        (src/include/utils/NbodyUtils.cpp:113-160, 223-258) around the density centre of the last call."""
        self.lib.gref_core_and_relaxation.argtypes = [C.c_void_p, _dp]
        self.lib.gref_core_and_relaxation.restype = None
        out = np.zeros(3)
        self.lib.gref_core_and_relaxation(self.h, _ptr(out, _dp))
        return tuple(out)

    def pad_layers(self):
        """Room for the reference's -l overrun (see oracle/ref_harness.cpp: gref_pad_layers); call before integration()."""
        self.lib.gref_pad_layers.argtypes = [C.c_void_p]
        self.lib.gref_pad_layers.restype = None
        self.lib.gref_pad_layers(self.h)

    def set_print_options(self, print_all=False, print_lagrange=False):
        self.lib.gref_set_print_options.argtypes = [C.c_void_p, C.c_int, C.c_int]
        self.lib.gref_set_print_options(self.h, int(print_all), int(print_lagrange))

    def integration(self):
        """The reference's own driver: writes <workdir>/run.out.{info,log,snapshot_*}."""
        self.lib.gref_integration(self.h)

    def _run(self, atime, itime0, t_end, max_steps, st, tn, ti, cap):
        self.lib.gref_run(self.h, atime, t_end, max_steps, C.byref(st), _ptr(tn, _up),
                          _ptr(ti, _dp), cap)


def best_oracle(*a, **k):
    """RefOracle when the reference library exists, else the port."""
    return RefOracle(*a, **k) if have_ref() else PortOracle(*a, **k)
