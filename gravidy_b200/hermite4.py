"""Python mirror of the reference's integrator hooks over the C ABI (tests, bench, notebooks).

``Hermite4B200`` exposes the calls the reference loop makes (reference:
src/include/cpu/Hermite4CPU_integration.cpp:81-124; hooks declared in
src/include/gpu/Hermite4GPU.cuh:131-142) under the same names and argument meaning.  Every method is a
thin ctypes call into ``libgravidy_b200.so``: no arithmetic of the hot path happens in Python, and there
is no CPU fallback.  The production host program is the C++ one in ``gravidy_b200/host``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi

_dp, _up = capi._dp, capi._up
ETA_S = np.float32(0.01)   # reference: src/include/common.hpp:114


def softening_to_e2(softening) -> float:
    """e2 as NbodySystem computes it: float product widened (reference: src/include/NbodySystem.cpp:50)."""
    s = np.float32(softening)
    return float(np.float32(s * s))


def _p(a, ty=_dp):
    return None if a is None else a.ctypes.data_as(ty)


class Hermite4B200:
    """One j-shard (= one GPU) of the B200 back end."""

    def __init__(self, m, r, v, softening=1e-4, eta=0.01, device=0, rank=0, world=1, nccl_uid=None,
                 e2=None):
        self.lib = capi.load()
        m = np.ascontiguousarray(m, dtype=np.float64)
        r = np.ascontiguousarray(r, dtype=np.float64)
        v = np.ascontiguousarray(v, dtype=np.float64)
        self.n = int(len(m))
        self.e2 = softening_to_e2(softening) if e2 is None else float(e2)
        self.eta = np.float32(eta)
        self.h = capi._ctx()
        uid = None
        if world > 1 and nccl_uid is not None:
            uid = (C.c_uint8 * capi.NCCL_UID_BYTES).from_buffer_copy(bytes(nccl_uid))
        rc = self.lib.gdy_create_sharded(C.byref(self.h), self.n, self.e2, C.c_float(float(self.eta)),
                                         device, rank, world, uid)
        if rc:
            msg = self.lib.gdy_last_error(None).decode()
            self.h = None
            raise capi.GdyError(rc, msg)
        r4 = np.zeros((self.n, 4)); r4[:, :3] = r; r4[:, 3] = m
        v4 = np.zeros((self.n, 4)); v4[:, :3] = v
        self.upload_particles(r4, v4)

    # ---- plumbing ------------------------------------------------------------------------------
    def _ck(self, rc):
        if rc:
            raise capi.GdyError(rc, self.lib.gdy_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.gdy_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def nccl_unique_id() -> bytes:
        lib = capi.load()
        buf = (C.c_uint8 * capi.NCCL_UID_BYTES)()
        rc = lib.gdy_nccl_unique_id(buf)
        if rc:
            raise capi.GdyError(rc, lib.gdy_last_error(None).decode())
        return bytes(buf)

    @staticmethod
    def shard_range(n, rank, world):
        """[begin, end) of the j-particles shard `rank` of `world` computes (host arithmetic only)."""
        b, e = C.c_uint32(), C.c_uint32()
        rc = capi.load().gdy_shard_range(n, rank, world, C.byref(b), C.byref(e))
        if rc:
            raise capi.GdyError(rc, "bad shard arguments")
        return b.value, e.value

    # ---- transfers -----------------------------------------------------------------------------
    def upload_particles(self, r4, v4):
        r4 = np.ascontiguousarray(r4, dtype=np.float64); v4 = np.ascontiguousarray(v4, dtype=np.float64)
        self._ck(self.lib.gdy_upload_particles(self.h, _p(r4), _p(v4)))

    def upload_times(self, t, dt):
        t = np.ascontiguousarray(t, dtype=np.float64); dt = np.ascontiguousarray(dt, dtype=np.float64)
        self._ck(self.lib.gdy_upload_times(self.h, _p(t), _p(dt)))

    def upload_forces(self, f6):
        f6 = np.ascontiguousarray(f6, dtype=np.float64)
        self._ck(self.lib.gdy_upload_forces(self.h, _p(f6)))

    def upload_derivs(self, old6=None, a2=None, a3=None):
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (old6, a2, a3)]
        self._ck(self.lib.gdy_upload_derivs(self.h, *[_p(a) for a in arrs]))

    def download_forces(self, out=None):
        f6 = np.empty((self.n, 6)) if out is None else out
        self._ck(self.lib.gdy_download_forces(self.h, _p(f6)))
        return f6

    _FIELDS = (("r4", 4), ("v4", 4), ("f6", 6), ("old6", 6), ("a2", 3), ("a3", 3), ("t", 0), ("dt", 0),
               ("p7", 7))

    def get_state(self, *names):
        """Same keys as the oracle's get_state: r4 v4 f6 old6 a2 a3 t dt p7 move."""
        want = names or [f for f, _ in self._FIELDS] + ["move"]
        out, args = {}, []
        for f, w in self._FIELDS:
            if f in want:
                out[f] = np.zeros((self.n, w) if w else (self.n,))
                args.append(_p(out[f]))
            else:
                args.append(None)
        if "move" in want:
            out["move"] = np.zeros(self.n, dtype=np.uint32)
            args.append(_p(out["move"], _up))
        else:
            args.append(None)
        self._ck(self.lib.gdy_download_state(self.h, *args))
        return out

    def set_state(self, r4=None, v4=None, f6=None, old6=None, a2=None, a3=None, t=None, dt=None):
        if r4 is not None or v4 is not None:
            cur = self.get_state("r4", "v4")
            self.upload_particles(cur["r4"] if r4 is None else r4, cur["v4"] if v4 is None else v4)
        if f6 is not None:
            self.upload_forces(f6)
        if old6 is not None or a2 is not None or a3 is not None:
            self.upload_derivs(old6, a2, a3)
        if t is not None or dt is not None:
            cur = self.get_state("t", "dt")
            self.upload_times(cur["t"] if t is None else t, cur["dt"] if dt is None else dt)

    # ---- hooks (names of the reference) --------------------------------------------------------
    def init_acc_jrk(self):
        self._ck(self.lib.gdy_init_acc_jrk(self.h))

    def init_dt(self, itime=0.0, eta_s=ETA_S):
        at = C.c_double()
        self._ck(self.lib.gdy_init_dt(self.h, C.c_float(float(eta_s)), itime, C.byref(at)))
        return at.value

    def find_particles_to_move(self, itime):
        nact = C.c_uint32()
        self._ck(self.lib.gdy_find_particles_to_move(self.h, itime, C.byref(nact)))
        return nact.value

    def save_old_acc_jrk(self, nact):
        self._ck(self.lib.gdy_save_old_acc_jrk(self.h, nact))

    def predicted_pos_vel(self, itime):
        self._ck(self.lib.gdy_predicted_pos_vel(self.h, itime))

    def update_acc_jrk(self, nact):
        self._ck(self.lib.gdy_update_acc_jrk(self.h, nact))

    def correction_pos_vel(self, itime, nact):
        self._ck(self.lib.gdy_correction_pos_vel(self.h, itime, nact))

    def next_integration_time(self):
        at = C.c_double()
        self._ck(self.lib.gdy_next_integration_time(self.h, C.byref(at)))
        return at.value

    def get_energy(self):
        ek, ep = C.c_double(), C.c_double()
        self._ck(self.lib.gdy_energy(self.h, C.byref(ek), C.byref(ep)))
        return ek.value + ep.value, ek.value, ep.value

    def energy_timed(self):
        """get_energy with the pair kernel bracketed by CUDA events; returns its milliseconds."""
        ek, ep, ms = C.c_double(), C.c_double(), C.c_float()
        self._ck(self.lib.gdy_energy_timed(self.h, C.byref(ek), C.byref(ep), C.byref(ms)))
        return ms.value

    def neighbour_timed(self, k):
        ms = C.c_float()
        self._ck(self.lib.gdy_neighbour_timed(self.h, k, C.byref(ms)))
        return ms.value

    def neighbour_distance(self, k):
        out = np.empty(self.n)
        self._ck(self.lib.gdy_neighbour_distance(self.h, k, _p(out)))
        return out

    # ---- fused driver / bench entries ----------------------------------------------------------
    def run(self, atime, t_end, max_steps=0, trace_cap=0, itime0=0.0, trace_us=False):
        st = capi.RunStats()
        tn = np.zeros(max(trace_cap, 1), dtype=np.uint32)
        ti = np.zeros(max(trace_cap, 1), dtype=np.float64)
        tu = np.zeros(max(trace_cap, 1), dtype=np.float32) if trace_us else None
        self._ck(self.lib.gdy_run(self.h, atime, itime0, t_end, max_steps, C.byref(st), _p(tn, _up),
                                  _p(ti), _p(tu, capi._fp), trace_cap))
        k = min(trace_cap, st.iterations)
        if trace_us:
            return st, tn[:k], ti[:k], tu[:k]
        return st, tn[:k], ti[:k]

    def force_sweep(self, move, want_forces=True):
        mv = np.ascontiguousarray(move, dtype=np.uint32)
        f6 = np.empty((len(mv), 6)) if want_forces else None
        ms = C.c_float()
        self._ck(self.lib.gdy_force_sweep(self.h, len(mv), _p(mv, _up), _p(f6), C.byref(ms)))
        return f6, ms.value

    def force_sweep_full(self, reps=1):
        ms = C.c_float()
        self._ck(self.lib.gdy_force_sweep_full(self.h, reps, C.byref(ms)))
        return ms.value

    def predict_timed(self, itime):
        """predicted_pos_vel(itime) for all N, returns the kernel's milliseconds (CUDA events)."""
        ms = C.c_float()
        self._ck(self.lib.gdy_predict_timed(self.h, itime, C.byref(ms)))
        return ms.value

    def set_tuning(self, variant=-1, jsplit=-1):
        self._ck(self.lib.gdy_set_tuning(self.h, variant, jsplit))

    def peer_export(self) -> bytes:
        """Handle of this shard's exchange buffer (for the fused NVLink exchange of the persistent path)."""
        buf = (C.c_uint8 * capi.PEER_HANDLE_BYTES)()
        self._ck(self.lib.gdy_peer_export(self.h, buf))
        return bytes(buf)

    def peer_attach(self, handles):
        """handles: the world's peer_export() results in rank order."""
        blob = b"".join(handles)
        buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        self._ck(self.lib.gdy_peer_attach(self.h, buf))

    def set_persist(self, nact_max=-1):
        """Active sets up to nact_max are stepped by the persistent kernel inside run() (0 = off)."""
        self._ck(self.lib.gdy_set_persist(self.h, nact_max))

    def persist_stats(self):
        """(launches, block steps) done by the persistent small-step kernel so far."""
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.gdy_persist_stats(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def persist_state(self):
        """(state, smem bytes needed per CTA, device limit): state 0 = off, 1 = available, 2 = the slices do not fit."""
        need, lim = C.c_uint32(), C.c_uint32()
        rc = self.lib.gdy_persist_state(self.h, C.byref(need), C.byref(lim))
        if rc < 0:
            self._ck(rc)
        return rc, need.value, lim.value

    def persist_profile(self):
        """ns per phase of the persistent kernel (CTA 0); zeros unless GDY_PERSIST_PROF=1."""
        out = (C.c_uint64 * 144)()
        self._ck(self.lib.gdy_persist_profile(self.h, out))
        return [int(x) for x in out]

    def measure_fp64_peak(self):
        tf, mhz = C.c_double(), C.c_double()
        self._ck(self.lib.gdy_measure_fp64_peak(self.h, C.byref(tf), C.byref(mhz)))
        return tf.value, mhz.value

    def measure_copy_bw(self):
        gb = C.c_double()
        self._ck(self.lib.gdy_measure_copy_bw(self.h, C.byref(gb)))
        return gb.value

    def exchange_stats(self):
        """(force exchanges through peer memory, through ncclAllReduce) so far."""
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.gdy_exchange_stats(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def launch_count(self):
        return int(self.lib.gdy_launch_count(self.h))

    def sync(self):
        self._ck(self.lib.gdy_sync(self.h))
