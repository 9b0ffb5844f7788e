"""ctypes view of include/jaybenne_b200.h.

The structures below mirror the C ABI field for field. `bind(lib, prefix)` attaches
argument / result types to every entry point; the same signatures are exported by the CUDA
library with prefix ``jb_`` (the product) and — for tests only — by the CPU oracle with
prefix ``jbo_``. This module never loads the oracle: `load_library()` only knows the
in-tree CUDA build and fails loudly when it is missing.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libjaybenne_b200.so")
HOST_LIB_PATH = os.path.join(HERE, "libjb_host.so")

# ---- enums ---------------------------------------------------------------------------
JB_OK, JB_ERR_INVALID, JB_ERR_CUDA, JB_ERR_CAPACITY, JB_ERR_STATE = 0, 1, 2, 3, 4
JB_ERR_UNSUPPORTED, JB_ERR_ITERATIONS, JB_ERR_NO_DEVICE, JB_ERR_COMM = 5, 6, 7, 8
STATUS_NAMES = {
    0: "JB_OK", 1: "JB_ERR_INVALID", 2: "JB_ERR_CUDA", 3: "JB_ERR_CAPACITY", 4: "JB_ERR_STATE",
    5: "JB_ERR_UNSUPPORTED", 6: "JB_ERR_ITERATIONS", 7: "JB_ERR_NO_DEVICE", 8: "JB_ERR_COMM",
}
JB_OPAC_GRAY, JB_OPAC_EPBREMSS = 0, 1
JB_SCAT_GRAY, JB_SCAT_THOMSON = 0, 1
JB_BC_PERIODIC, JB_BC_REFLECT, JB_BC_OUTFLOW = 0, 1, 2
JB_SOURCE_THERMAL, JB_SOURCE_EMISSION = 0, 1
JB_STRATEGY_UNIFORM, JB_STRATEGY_ENERGY = 0, 1
JB_RNG_PHILOX, JB_RNG_XORSHIFT_POOL, JB_RNG_XORSHIFT_THREADS = 0, 1, 2
JB_GRAY_KIRCHHOFF, JB_GRAY_DIRECT = 0, 1
(JB_FIELD_DENSITY, JB_FIELD_SIE, JB_FIELD_INTERNAL_ENERGY, JB_FIELD_ENERGY_TALLY, JB_FIELD_FLECK_FACTOR,
 JB_FIELD_SOURCE_EW_PER_CELL, JB_FIELD_SOURCE_NUM_PER_CELL, JB_FIELD_ENERGY_DELTA, JB_FIELD_DDMC_FACE_PROB_X1,
 JB_FIELD_DDMC_FACE_PROB_X2, JB_FIELD_DDMC_FACE_PROB_X3) = range(11)
JB_EV_SCATTER, JB_EV_ABSORB, JB_EV_CENSUS, JB_EV_LEFT_BLOCK = 1, 2, 4, 8
JB_EV_DDMC, JB_EV_ALBEDO_REJECT, JB_EV_ESCAPED = 16, 32, 64


class JbError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status


# ---- structures ----------------------------------------------------------------------
class jb_config(C.Structure):
    _fields_ = [
        ("num_particles", C.c_int64), ("dt", C.c_double), ("min_swarm_occupancy", C.c_double),
        ("numin", C.c_double), ("numax", C.c_double), ("unique_rank_seeds", C.c_int32),
        ("seed", C.c_int64), ("max_transport_iterations", C.c_int32), ("use_ddmc", C.c_int32),
        ("tau_ddmc", C.c_double), ("source_strategy", C.c_int32), ("do_emission", C.c_int32),
        ("do_feedback", C.c_int32), ("opacity_model", C.c_int32), ("opacity_kappa", C.c_double),
        ("scattering_model", C.c_int32), ("scattering_sigma", C.c_double), ("apm", C.c_double),
        ("time_scale", C.c_double), ("mass_scale", C.c_double), ("length_scale", C.c_double),
        ("temperature_scale", C.c_double), ("eos_gm1", C.c_double), ("eos_cv", C.c_double),
        ("rng_mode", C.c_int32), ("gray_mode", C.c_int32), ("source_nblocks", C.c_int32),
        ("sort_interval", C.c_int32), ("reserved", C.c_int32 * 8),
    ]


class jb_block(C.Structure):
    _fields_ = [
        ("gid", C.c_int32), ("level", C.c_int32), ("lx", C.c_int64 * 3), ("xmin", C.c_double * 3),
        ("xmax", C.c_double * 3), ("rank", C.c_int32), ("reserved", C.c_int32),
    ]


class jb_mesh(C.Structure):
    _fields_ = [
        ("ndim", C.c_int32), ("nxb", C.c_int32 * 3), ("nrb", C.c_int32 * 3), ("max_level", C.c_int32),
        ("nghost", C.c_int32), ("xmin", C.c_double * 3), ("xmax", C.c_double * 3),
        ("mesh_bc", C.c_int32 * 6), ("swarm_bc", C.c_int32 * 6), ("nblocks", C.c_int32),
        ("blocks", C.POINTER(jb_block)),
    ]


class jb_cycle_stats(C.Structure):
    _fields_ = [
        ("n_start", C.c_int64), ("n_sourced", C.c_int64), ("n_absorbed", C.c_int64), ("n_escaped", C.c_int64),
        ("n_end", C.c_int64), ("n_sent", C.c_int64), ("n_received", C.c_int64), ("n_block_crossings", C.c_int64),
        ("n_segments", C.c_int64), ("n_scatters", C.c_int64), ("n_ddmc_segments", C.c_int64),
        ("n_iterations", C.c_int32), ("n_bface_resampled", C.c_int32), ("e_start", C.c_double), ("e_sourced", C.c_double),
        ("e_absorbed", C.c_double), ("e_escaped", C.c_double), ("e_end", C.c_double),
        ("ms_transport", C.c_double), ("ms_total", C.c_double),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


_pd, _pi, _pu = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint32)


class jb_particles(C.Structure):
    _fields_ = [
        ("x", _pd), ("y", _pd), ("z", _pd), ("time", _pd), ("weight", _pd), ("energy", _pd),
        ("v0", _pd), ("v1", _pd), ("v2", _pd), ("i", _pi), ("j", _pi), ("k", _pi), ("block", _pi), ("rng", _pu),
    ]


class jb_event(C.Structure):
    _fields_ = [
        ("id_idx", C.c_uint32), ("id_gid", C.c_uint32), ("id_call", C.c_uint32), ("draw", C.c_uint32),
        ("kind", C.c_int32), ("block", C.c_int32), ("i", C.c_int32), ("j", C.c_int32), ("k", C.c_int32),
        ("reserved", C.c_int32), ("t", C.c_double), ("x", C.c_double), ("y", C.c_double), ("z", C.c_double),
        ("v0", C.c_double), ("v1", C.c_double), ("v2", C.c_double),
    ]


ctx_p = C.c_void_p
deck_p = C.c_void_p

# name -> (restype, argtypes); names without prefix. Compute entry points exist in both the
# CUDA library (jb_) and the oracle (jbo_); DECK_API only in the CUDA library (host code).
COMPUTE_API = {
    "create": (C.c_int, [C.POINTER(jb_config), C.POINTER(ctx_p)]),
    "destroy": (None, [ctx_p]),
    "last_error": (C.c_char_p, [ctx_p]),
    "get_constants": (C.c_int, [ctx_p, _pd, _pd]),
    "set_mesh": (C.c_int, [ctx_p, C.POINTER(jb_mesh), C.c_int32, C.c_int32]),
    "set_capacity": (C.c_int, [ctx_p, C.c_int64]),
    "set_field": (C.c_int, [ctx_p, C.c_int32, C.c_int32, C.c_void_p]),
    "get_field": (C.c_int, [ctx_p, C.c_int32, C.c_int32, C.c_void_p]),
    "field_size": (C.c_int64, [ctx_p, C.c_int32]),
    "initialize_radiation": (C.c_int, [ctx_p, C.c_int32]),
    "radiation_step": (C.c_int, [ctx_p, C.c_double, C.c_double, C.POINTER(jb_cycle_stats)]),
    "update_derived_transport_fields": (C.c_int, [ctx_p, C.c_double]),
    "source_photons": (C.c_int, [ctx_p, C.c_int32, C.c_double, C.c_double]),
    "transport_photons": (C.c_int, [ctx_p, C.c_double, C.c_double]),
    "swarm_exchange": (C.c_int, [ctx_p]),
    "sample_ddmc_block_face": (C.c_int, [ctx_p]),
    "check_completion": (C.c_int, [ctx_p, C.c_double, C.POINTER(C.c_int64)]),
    "evaluate_radiation_energy": (C.c_int, [ctx_p]),
    "update_fluid": (C.c_int, [ctx_p]),
    "update_derived_material": (C.c_int, [ctx_p]),
    "post_initialization": (C.c_int, [ctx_p]),
    "num_particles": (C.c_int, [ctx_p, C.POINTER(C.c_int64)]),
    "get_particles": (C.c_int, [ctx_p, C.POINTER(jb_particles), C.c_int64, C.POINTER(C.c_int64)]),
    "set_particles": (C.c_int, [ctx_p, C.POINTER(jb_particles), C.c_int64]),
    "reset_streams": (C.c_int, [ctx_p]),
    "get_stats": (C.c_int, [ctx_p, C.POINTER(jb_cycle_stats)]),
    "stats_begin": (C.c_int, [ctx_p]),
    "event_log_enable": (C.c_int, [ctx_p, C.c_int64]),
    "event_log_read": (C.c_int, [ctx_p, C.POINTER(jb_event), C.c_int64, C.POINTER(C.c_int64)]),
    "exchange_record_bytes": (C.c_int64, []),
    "exchange_count": (C.c_int, [ctx_p, C.c_int32, C.POINTER(C.c_int64)]),
    "exchange_take": (C.c_int, [ctx_p, C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "exchange_give": (C.c_int, [ctx_p, C.c_void_p, C.c_int64]),
}
# entry points only the CUDA library has
DEVICE_API = {
    "config_default": (None, [C.POINTER(jb_config)]),
    "abi_version": (C.c_int, []),
    "kernel_launches": (C.c_int64, [ctx_p]),
    "device_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "set_device": (C.c_int, [C.c_int32]),
    "measure_fp64_peak": (C.c_int, [ctx_p, _pd]),
    "selftest_division": (C.c_int, [ctx_p, C.c_int64, C.POINTER(C.c_int64)]),
    "timer_start": (C.c_int, [ctx_p]),
    "timer_stop": (C.c_int, [ctx_p, _pd]),
    "comm_unique_id": (C.c_int, [C.c_void_p]),
    "comm_init": (C.c_int, [ctx_p, C.c_void_p, C.c_int32, C.c_int32]),
    "exchange_take_device": (C.c_int, [ctx_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "exchange_give_device": (C.c_int, [ctx_p, C.c_void_p, C.c_int64]),
    "exchange_clear": (C.c_int, [ctx_p]),
    "set_sort_interval": (C.c_int, [ctx_p, C.c_int32]),
    "exchange_mode": (C.c_int32, [ctx_p]),
    "evaluate_radiation_energy_noalloc": (C.c_int, [ctx_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), _pd]),
    "step_timeline": (C.c_int, [ctx_p, C.POINTER(C.c_int32), C.POINTER(C.c_float), C.c_int32, C.POINTER(C.c_int32)]),
    "set_remote_field": (C.c_int, [ctx_p, C.c_int32, C.c_int32, C.c_void_p]),
    "set_fields": (C.c_int, [ctx_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_void_p)]),
    "get_fields": (C.c_int, [ctx_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_void_p)]),
}
DECK_API = {
    "deck_load": (C.c_int, [C.c_char_p, C.POINTER(deck_p)]),
    "deck_parse": (C.c_int, [C.c_char_p, C.POINTER(deck_p)]),
    "deck_free": (None, [deck_p]),
    "deck_set": (C.c_int, [deck_p, C.c_char_p, C.c_char_p, C.c_char_p]),
    "deck_has": (C.c_int32, [deck_p, C.c_char_p, C.c_char_p]),
    "deck_get": (C.c_char_p, [deck_p, C.c_char_p, C.c_char_p]),
    "deck_config": (C.c_int, [deck_p, C.POINTER(jb_config), C.c_char_p, C.c_int32]),
    "deck_mesh": (C.c_int, [deck_p, C.c_int32, C.POINTER(jb_mesh), C.c_char_p, C.c_int32]),
    "mesh_free": (None, [C.POINTER(jb_mesh)]),
}


class Api:
    """Entry points of one shared library, reachable as attributes without the prefix."""

    def __init__(self, lib, prefix, tables):
        self.lib, self.prefix = lib, prefix
        for table in tables:
            for name, (res, args) in table.items():
                fn = getattr(lib, prefix + name)
                fn.restype, fn.argtypes = res, args
                setattr(self, name, fn)


_LIB = None


def load_library():
    """The in-tree CUDA build. There is no fallback: a missing library is an error."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or make -C jaybenne_b200/csrc). jaybenne_b200 has no CPU fallback.")
        _LIB = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    return _LIB


def device_api():
    return Api(load_library(), "jb_", (COMPUTE_API, DEVICE_API, DECK_API))


_HOST = None


def host_api():
    """The deck reader / mesh-table builder (csrc/host/deck.cpp) as its own host-only library: parsing an input
    deck must not need (or map) the CUDA library."""
    global _HOST
    if _HOST is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise ImportError(f"{HOST_LIB_PATH} is missing: build it with make -C jaybenne_b200/csrc")
        lib = C.CDLL(HOST_LIB_PATH)
        _HOST = Api(lib, "jb_", (DECK_API, {"config_default": DEVICE_API["config_default"]}))
    return _HOST


def declared_symbols():
    """Every symbol include/jaybenne_b200.h declares (tests check the .so exports them all)."""
    names = []
    for table in (COMPUTE_API, DEVICE_API, DECK_API):
        names += ["jb_" + n for n in table]
    return names
