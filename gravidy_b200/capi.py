"""ctypes binding of the C ABI in ``include/gravidy_b200.h`` — the only way Python reaches the kernels.

There is no CPU fallback: if ``libgravidy_b200.so`` is missing this module raises at ``load()``, and
``gdy_create`` fails with ``GDY_ERR_NO_DEVICE`` when there is no sm_100 GPU.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# GDY_LIB: a lab build of the same library (tools/force_lab.sh compiles inner-loop variants side by side)
LIB_PATH = os.environ.get("GDY_LIB") or os.path.join(HERE, "libgravidy_b200.so")

_dp = C.POINTER(C.c_double)
_up = C.POINTER(C.c_uint32)
_fp = C.POINTER(C.c_float)
_ctx = C.c_void_p

GDY_OK, GDY_ERR_INVALID, GDY_ERR_NO_DEVICE, GDY_ERR_CUDA, GDY_ERR_NCCL, GDY_ERR_STATE = 0, -1, -2, -3, -4, -5
NCCL_UID_BYTES = 128
PEER_HANDLE_BYTES = 128
NOALLOC_MAX_N = 1 << 20   # GDY_NOALLOC_MAX_N


class RunStats(C.Structure):
    _fields_ = [
        ("iterations", C.c_uint64),
        ("nsteps", C.c_uint64),
        ("interactions", C.c_uint64),
        ("itime", C.c_double),
        ("atime", C.c_double),
        ("force_ms", C.c_double),
        ("wall_ms", C.c_double),
        ("last_nact", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


# every symbol include/gravidy_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "gdy_abi_version": (C.c_int, []),
    "gdy_device_count": (C.c_int, []),
    "gdy_neighbour_distance": (C.c_int, [_ctx, C.c_int, _dp]),
    "gdy_create": (C.c_int, [C.POINTER(_ctx), C.c_uint32, C.c_double, C.c_float, C.c_int]),
    "gdy_create_sharded": (C.c_int, [C.POINTER(_ctx), C.c_uint32, C.c_double, C.c_float, C.c_int,
                                     C.c_int, C.c_int, C.POINTER(C.c_uint8)]),
    "gdy_shard_range": (C.c_int, [C.c_uint32, C.c_int, C.c_int, _up, _up]),
    "gdy_nccl_unique_id": (C.c_int, [C.POINTER(C.c_uint8)]),
    "gdy_destroy": (C.c_int, [_ctx]),
    "gdy_last_error": (C.c_char_p, [_ctx]),
    "gdy_upload_particles": (C.c_int, [_ctx, _dp, _dp]),
    "gdy_upload_times": (C.c_int, [_ctx, _dp, _dp]),
    "gdy_upload_forces": (C.c_int, [_ctx, _dp]),
    "gdy_download_forces": (C.c_int, [_ctx, _dp]),
    "gdy_download_state": (C.c_int, [_ctx] + [_dp] * 9 + [_up]),
    "gdy_upload_derivs": (C.c_int, [_ctx, _dp, _dp, _dp]),
    "gdy_init_acc_jrk": (C.c_int, [_ctx]),
    "gdy_init_dt": (C.c_int, [_ctx, C.c_float, C.c_double, _dp]),
    "gdy_find_particles_to_move": (C.c_int, [_ctx, C.c_double, _up]),
    "gdy_save_old_acc_jrk": (C.c_int, [_ctx, C.c_uint32]),
    "gdy_predicted_pos_vel": (C.c_int, [_ctx, C.c_double]),
    "gdy_update_acc_jrk": (C.c_int, [_ctx, C.c_uint32]),
    "gdy_correction_pos_vel": (C.c_int, [_ctx, C.c_double, C.c_uint32]),
    "gdy_next_integration_time": (C.c_int, [_ctx, _dp]),
    "gdy_energy": (C.c_int, [_ctx, _dp, _dp]),
    "gdy_energy_timed": (C.c_int, [_ctx, _dp, _dp, _fp]),
    "gdy_neighbour_timed": (C.c_int, [_ctx, C.c_int, _fp]),
    # device pointers and a cudaStream_t, all as integers (the caller owns the memory: nothing is allocated)
    "gdy_energy_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
    "gdy_run": (C.c_int, [_ctx, C.c_double, C.c_double, C.c_double, C.c_uint64,
                          C.POINTER(RunStats), _up, _dp, _fp, C.c_uint64]),
    "gdy_force_sweep": (C.c_int, [_ctx, C.c_uint32, _up, _dp, _fp]),
    "gdy_force_sweep_full": (C.c_int, [_ctx, C.c_int, _fp]),
    "gdy_predict_timed": (C.c_int, [_ctx, C.c_double, _fp]),
    "gdy_set_tuning": (C.c_int, [_ctx, C.c_int, C.c_int]),
    "gdy_set_persist": (C.c_int, [_ctx, C.c_int]),
    "gdy_persist_stats": (C.c_int, [_ctx, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "gdy_persist_state": (C.c_int, [_ctx, _up, _up]),
    "gdy_peer_export": (C.c_int, [_ctx, C.POINTER(C.c_uint8)]),
    "gdy_peer_attach": (C.c_int, [_ctx, C.POINTER(C.c_uint8)]),
    "gdy_persist_profile": (C.c_int, [_ctx, C.POINTER(C.c_uint64)]),
    "gdy_measure_fp64_peak": (C.c_int, [_ctx, _dp, _dp]),
    "gdy_measure_copy_bw": (C.c_int, [_ctx, _dp]),
    "gdy_plan_dry": (C.c_int, [C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, _up]),
    "gdy_launch_count": (C.c_uint64, [_ctx]),
    "gdy_exchange_stats": (C.c_int, [_ctx, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "gdy_sync": (C.c_int, [_ctx]),
}

_lib = None


class GdyError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"gravidy_b200 error {code}: {msg}")
        self.code = code


def load() -> C.CDLL:
    """dlopen the library and declare every prototype.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GdyError(GDY_ERR_NO_DEVICE,
                       f"{LIB_PATH} is missing: build it with `python -m gravidy_b200.build` "
                       "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError here = header and library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
