"""
ctypes binding of libswarm_b200.so (include/swarm_b200.h).

There is no CPU fallback: if the shared object is missing or a CUDA device is
absent, the calls raise.  `load()` only dlopens the library (works without a GPU,
used by the CPU-side symbol check); every compute entry point needs a B200.
"""
from __future__ import annotations

import ctypes as C
import pathlib

PKG = pathlib.Path(__file__).resolve().parent
import os

LIB_PATH = pathlib.Path(os.environ.get("SWARM_B200_LIB", PKG / "libswarm_b200.so"))

MAX_SPECIES = 16
MAX_TYPES = 8
MAX_NEIGHBORS = 128  # SWARM_MAX_NEIGHBORS: hard ceiling of the Verlet capacity

OK, ERR_INVALID, ERR_STATE, ERR_CUDA, ERR_OVERFLOW, ERR_NOMEM = 0, -1, -2, -3, -4, -5
DECAY_NONE, DECAY_LINEAR, DECAY_EXP, DECAY_INVERSE = 0, 1, 2, 3
INTEGRATOR_BROWNIAN, INTEGRATOR_LANGEVIN = 0, 1
FLAG_NO_PERSISTENT = 1

_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int32)


class Config(C.Structure):
    _fields_ = [
        ("n_dims", C.c_int32),
        ("periodic", C.c_int32 * 3),
        ("box", C.c_double * 3),
        ("dt", C.c_double),
        ("kT", C.c_double),
        ("skin", C.c_double),
        ("seed", C.c_uint64),
        ("device", C.c_int32),
        ("max_neighbors", C.c_int32),
        ("n_envs", C.c_int32),
        ("flags", C.c_int32),
    ]


class StateDev(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in "x y z theta q0 q1 q2 q3 vx vy vz wx wy wz species".split()]


class ActionsDev(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in "f_swim tqx tqy tqz fex fey fez ndx ndy ndz has_nd".split()]


class Stats(C.Structure):
    _fields_ = [
        ("steps_done", C.c_int64),
        ("steps_requested", C.c_int64),
        ("rebuilds", C.c_int64),
        ("launches", C.c_int64),
        ("overflow", C.c_int32),
        ("max_neighbors_seen", C.c_int32),
        ("capacity", C.c_int32),
        ("persistent", C.c_int32),
    ]


DD_MAX_AUX = 8
RUN_CONTINUE, RUN_LEAVE_OPEN = 1, 2


class DDGeom(C.Structure):
    _fields_ = [("slab_w", C.c_double), ("halo", C.c_double), ("n_aux", C.c_int32), ("cap", C.c_int32),
                ("image_shift_left", C.c_int32), ("image_shift_right", C.c_int32)]


class DDArrays(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in "x y theta f_swim tqz gid species image owned".split()] + [("aux", C.c_void_p * DD_MAX_AUX)]


class Policy(C.Structure):
    _fields_ = [("obs_dim", C.c_int32), ("hidden", C.c_int32), ("n_actions", C.c_int32)] + [(n, C.c_void_p) for n in "w1 b1 wa ba wv bv".split()]


# every symbol include/swarm_b200.h declares: (name, restype, argtypes)
_V = C.c_void_p
SYMBOLS = {
    "swarm_last_error": (C.c_char_p, []),
    "swarm_abi_version": (C.c_int, []),
    "swarm_create": (C.c_int, [C.POINTER(Config), C.POINTER(_V)]),
    "swarm_destroy": (C.c_int, [_V]),
    "swarm_set_species": (C.c_int, [_V, C.c_int32, _IP, _DP, _DP, _DP, _DP]),
    "swarm_set_wca": (C.c_int, [_V, C.c_int32, _DP, _DP, _DP]),
    "swarm_set_walls": (C.c_int, [_V, C.c_int32, _DP, _DP, _DP]),
    "swarm_set_thermostat": (C.c_int, [_V, C.c_double, C.c_uint64]),
    "swarm_set_integrator": (C.c_int, [_V, C.c_int32]),
    "swarm_bind_state": (C.c_int, [_V, C.c_int64, C.POINTER(StateDev)]),
    "swarm_bind_actions": (C.c_int, [_V, C.POINTER(ActionsDev)]),
    "swarm_reserve": (C.c_int, [_V, C.c_int64]),
    "swarm_bind_ids": (C.c_int, [_V, _V]),
    "swarm_bind_count": (C.c_int, [_V, _V]),
    "swarm_dd_buffer_doubles": (C.c_int64, [C.POINTER(DDGeom)]),
    "swarm_dd_pack": (C.c_int, [_V, C.POINTER(DDGeom), C.c_int64, _V, C.POINTER(DDArrays), _V, _V, _V]),
    "swarm_dd_merge": (C.c_int, [_V, C.POINTER(DDGeom), C.c_int64, _V, C.POINTER(DDArrays), _V, _V, C.c_int64,
                                 C.POINTER(DDArrays), _V, _V]),
    "swarm_dd_obs_field": (C.c_int, [_V, C.c_int64, _V, C.POINTER(DDArrays), C.c_double, C.c_double, _DP, _DP, C.c_int32,
                                     C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32, _V, _V]),
    "swarm_dd_obs_field_ch": (C.c_int, [_V, C.c_int64, _V, C.POINTER(DDArrays), C.c_double, C.c_double, _DP, _DP, C.c_int32,
                                        C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _V, _V]),
    "swarm_bind_images": (C.c_int, [_V, _V]),
    "swarm_dd_set_lean": (C.c_int, [_V, C.c_int32]),
    "swarm_dd_peer_region_bytes": (C.c_int64, [C.POINTER(DDGeom)]),
    "swarm_dd_peer_alloc": (C.c_int, [C.c_int32, C.POINTER(DDGeom), C.POINTER(_V), C.c_char_p]),
    "swarm_dd_peer_open": (C.c_int, [C.c_int32, C.c_char_p, C.POINTER(_V)]),
    "swarm_dd_peer_close": (C.c_int, [_V]),
    "swarm_dd_peer_free": (C.c_int, [_V]),
    "swarm_dd_peer_recv": (_V, [_V, C.POINTER(DDGeom), C.c_int32, C.c_int32]),
    "swarm_dd_push": (C.c_int, [_V, C.POINTER(DDGeom), _V, _V, _V, _V, C.c_uint64, _V]),
    "swarm_dd_wait": (C.c_int, [_V, _V, C.c_uint64, _V, _V]),
    "swarm_apply_new_direction": (C.c_int, [_V, _V]),
    "swarm_relax": (C.c_int, [_V, C.c_int64, C.c_double, C.c_double, _V]),
    "swarm_run": (C.c_int, [_V, C.c_int64, C.c_uint64, _V]),
    "swarm_run_part": (C.c_int, [_V, C.c_int64, C.c_uint64, C.c_int32, _V]),
    "swarm_sync_stats": (C.c_int, [_V, _V, C.POINTER(Stats)]),
    "swarm_set_max_neighbors": (C.c_int, [_V, C.c_int32]),
    "swarm_debug_timing": (C.c_int, [_V, C.POINTER(C.c_int64)]),
    "swarm_set_tiling": (C.c_int, [_V, C.c_int32, C.c_int32]),
    "swarm_tiling_stats": (C.c_int, [_V, C.POINTER(C.c_int64)]),
    "swarm_tiled_neighbors_dump": (C.c_int, [_V, _IP, C.c_int64, C.POINTER(C.c_int64)]),
    "swarm_directors": (C.c_int, [_V, _V, _V, _V, _V]),
    "swarm_neighbors_dump": (C.c_int, [_V, _IP, C.c_int64, C.POINTER(C.c_int64)]),
    "swarm_cell_index": (C.c_int, [_V, C.POINTER(C.c_int32), _IP]),
    "swarm_forces": (C.c_int, [_V, _V, _V, _V, _V]),
    "swarm_philox_host": (C.c_int, [C.POINTER(C.c_uint64)] * 3),
    "swarm_noise_words_host": (C.c_int, [C.c_uint64] * 4 + [C.POINTER(C.c_uint64)]),
    "swarm_gaussians": (C.c_int, [_V, C.c_int64, C.c_uint64, C.c_uint64, _V, _V]),
    "swarm_obs_field": (C.c_int, [_V, _V, _DP, _DP, _V, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_int32, _V, _V, _V, _V]),
    "swarm_obs_vision": (C.c_int, [_V, _V, C.c_int64, C.c_double, C.c_double, C.c_int32, _V, _V, C.c_int32, C.c_int32, _V, _V, _V]),
    "swarm_obs_pair_field": (C.c_int, [_V, _V, _V, C.c_int64, _DP, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_int32,
                                       C.c_int32, _V, _V, _V, _V]),
    "swarm_obs_position": (C.c_int, [_V, _V, _DP, _V, _V]),
    "swarm_reduce_mean": (C.c_int, [_V, _V, C.c_int64, _V, _V]),
    "swarm_obs_field_mean": (C.c_int, [_V, _V, C.c_int64, _DP, _DP, _V, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_int32, _V, _V, _V]),
    "swarm_obs_field_mean_noalloc": (C.c_int, [_V, _V, C.c_int64, _DP, _DP, _V, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_int32, _V, _V, _V]),
    "swarm_policy_forward": (C.c_int, [_V, C.POINTER(Policy), _V, _V, C.c_uint64, C.c_int32, _V, _V, _V, _V, _V, _V, _V]),
    "swarm_set_actions_host": (C.c_int, [_V, _V, _V, _V]),
    "swarm_set_actions_host_2d": (C.c_int, [_V, _V, _V, _V]),
    "swarm_microbench_dfma": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, _DP]),
    "swarm_snapshot_dev": (C.c_int, [_V, _V, _V, _V, _V]),
    "swarm_get_particle_data_host": (C.c_int, [_V, _V, _V, _V, _V]),
}

_lib = None


class SwarmLibraryError(RuntimeError):
    pass


def load():
    """dlopen the CUDA library; raises (never falls back) when it is absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise SwarmLibraryError(
                f"{LIB_PATH} is missing: build it with `python -m swarmrl_b200.build` "
                "(nvcc, sm_100a). There is no CPU fallback."
            )
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return load().swarm_last_error().decode()


def check(rc: int):
    """Map C status codes onto the exceptions the reference raises."""
    if rc == OK:
        return
    msg = last_error()
    if rc == ERR_INVALID:
        raise ValueError(msg)
    if rc == ERR_STATE:
        raise RuntimeError(msg)
    if rc == ERR_OVERFLOW:
        raise OverflowError(msg)
    if rc == ERR_NOMEM:
        raise MemoryError(msg)
    raise SwarmLibraryError(msg)


def dp(arr):
    """double* view of a 1-D numpy float64 array."""
    return arr.ctypes.data_as(_DP)


def ip(arr):
    return arr.ctypes.data_as(_IP)
