"""Host-side mirror of the jaybenne package interface (src/jaybenne/jaybenne.hpp:50-76).

`Jaybenne` is what mcblock holds after `jaybenne::Initialize(pin, opacity, scattering, eos)`:
the package parameters plus the photons swarm and the package's cell fields, here owned by
the CUDA library behind the C ABI. Method names follow the reference's tasks:

    Initialize            jaybenne::Initialize                 jaybenne.cpp:158
    InitializeRadiation   jaybenne::InitializeRadiation        jaybenne.cpp:570
    RadiationStep         jaybenne::RadiationStep(...).Execute jaybenne.cpp:68
    UpdateDerivedTransportFields / SourcePhotons / TransportPhotons / SampleDDMCBlockFace /
    CheckCompletion / EvaluateRadiationEnergy / UpdateFluid    jaybenne.hpp:59-69

The class is generic over an `Api` (capi.py): the product always passes the CUDA library's
`jb_` entry points; tests may pass the CPU oracle's `jbo_` twins to check one against the
other through the very same host code.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import (JB_FIELD_DENSITY, JB_FIELD_ENERGY_DELTA, JB_FIELD_ENERGY_TALLY, JB_FIELD_FLECK_FACTOR,
                   JB_FIELD_INTERNAL_ENERGY, JB_FIELD_SIE, JB_OK, JB_SOURCE_EMISSION, JB_SOURCE_THERMAL, JbError,
                   jb_config, jb_cycle_stats, jb_event, jb_mesh, jb_particles)

FIELDS = {
    "density": JB_FIELD_DENSITY, "sie": JB_FIELD_SIE, "internal_energy": JB_FIELD_INTERNAL_ENERGY,
    "energy_tally": JB_FIELD_ENERGY_TALLY, "fleck_factor": JB_FIELD_FLECK_FACTOR,
    "source_ew_per_cell": capi.JB_FIELD_SOURCE_EW_PER_CELL, "source_num_per_cell": capi.JB_FIELD_SOURCE_NUM_PER_CELL,
    "energy_delta": JB_FIELD_ENERGY_DELTA, "ddmc_face_prob_x1": capi.JB_FIELD_DDMC_FACE_PROB_X1,
    "ddmc_face_prob_x2": capi.JB_FIELD_DDMC_FACE_PROB_X2, "ddmc_face_prob_x3": capi.JB_FIELD_DDMC_FACE_PROB_X3,
}


def default_config(api=None) -> jb_config:
    """jb_config with the reference defaults of the <jaybenne> block (jaybenne.cpp:163-223)."""
    cfg = jb_config()
    if api is None:
        api = capi.host_api()
    api.config_default(C.byref(cfg))
    return cfg


class Jaybenne:
    def __init__(self, api, cfg: jb_config, mesh: jb_mesh, rank: int = 0, n_ranks: int = 1, capacity: int = 0):
        self.api = api
        self.cfg = cfg
        self.mesh = mesh
        self.rank, self.n_ranks = rank, n_ranks
        self._ctx = capi.ctx_p()
        st = api.create(C.byref(cfg), C.byref(self._ctx))
        if st != JB_OK:
            self._ctx = capi.ctx_p()
            raise JbError(st, "jb_create failed (no CUDA device?)" if st == capi.JB_ERR_NO_DEVICE else "jb_create failed")
        self._check(api.set_mesh(self._ctx, C.byref(mesh), rank, n_ranks))
        if capacity:
            self._check(api.set_capacity(self._ctx, int(capacity)))
        self.local_gids = [mesh.blocks[g].gid for g in range(mesh.nblocks) if mesh.blocks[g].rank == rank]
        self.nxb = tuple(mesh.nxb)

    # reference spelling
    @classmethod
    def Initialize(cls, api, cfg, mesh, **kw):
        return cls(api, cfg, mesh, **kw)

    def close(self):
        if getattr(self, "_ctx", None):
            self.api.destroy(self._ctx)
            self._ctx = capi.ctx_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st):
        if st != JB_OK:
            msg = self.api.last_error(self._ctx)
            raise JbError(st, msg.decode() if msg else "")

    # ---- constants / fields ----------------------------------------------------------
    def constants(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.api.get_constants(self._ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    def field_size(self, name):
        return int(self.api.field_size(self._ctx, FIELDS[name]))

    def field_shape(self, name):
        n1, n2, n3 = self.nxb
        return {"ddmc_face_prob_x1": (n3, n2, n1 + 1), "ddmc_face_prob_x2": (n3, n2 + 1, n1),
                "ddmc_face_prob_x3": (n3 + 1, n2, n1)}.get(name, (n3, n2, n1))

    def set_field(self, gid, name, values):
        a = np.ascontiguousarray(values, dtype=np.float64).ravel()
        assert a.size == self.field_size(name), (a.size, self.field_size(name))
        self._check(self.api.set_field(self._ctx, gid, FIELDS[name], a.ctypes.data_as(C.c_void_p)))

    def set_field_ptr(self, gid, name, ptr):
        """`ptr`: raw host or device address of field_size(name) doubles."""
        self._check(self.api.set_field(self._ctx, gid, FIELDS[name], C.c_void_p(ptr)))

    def set_fields_ptr(self, name, gids, ptrs):
        """One call for many blocks: raw host or device addresses, one per gid (MeshData-style pack)."""
        n = len(gids)
        g = (C.c_int32 * n)(*gids)
        p = (C.c_void_p * n)(*ptrs)
        self._check(self.api.set_fields(self._ctx, FIELDS[name], n, g, p))

    def get_fields_ptr(self, name, gids, ptrs):
        n = len(gids)
        g = (C.c_int32 * n)(*gids)
        p = (C.c_void_p * n)(*ptrs)
        self._check(self.api.get_fields(self._ctx, FIELDS[name], n, g, p))

    def set_remote_field(self, gid, name, values):
        """density / sie of a block another rank owns (multi-rank DDMC, host-driven exchange)."""
        a = np.ascontiguousarray(values, dtype=np.float64).ravel()
        self._check(self.api.set_remote_field(self._ctx, gid, FIELDS[name], a.ctypes.data_as(C.c_void_p)))

    def get_field(self, gid, name, out=None):
        a = out if out is not None else np.empty(self.field_size(name), dtype=np.float64)
        self._check(self.api.get_field(self._ctx, gid, FIELDS[name], a.ctypes.data_as(C.c_void_p)))
        return a.reshape(self.field_shape(name)) if out is None else out

    def get_field_ptr(self, gid, name, ptr):
        self._check(self.api.get_field(self._ctx, gid, FIELDS[name], C.c_void_p(ptr)))

    # ---- tasks -----------------------------------------------------------------------
    def InitializeRadiation(self, is_thermal: bool):
        self._check(self.api.initialize_radiation(self._ctx, 1 if is_thermal else 0))

    def RadiationStep(self, t_start: float, dt: float) -> dict:
        st = jb_cycle_stats()
        self._check(self.api.radiation_step(self._ctx, t_start, dt, C.byref(st)))
        return st.as_dict()

    def UpdateDerivedTransportFields(self, dt):
        self._check(self.api.update_derived_transport_fields(self._ctx, dt))

    def SourcePhotons(self, source_type, t_start, dt):
        self._check(self.api.source_photons(self._ctx, source_type, t_start, dt))

    def TransportPhotons(self, t_start, dt):
        self._check(self.api.transport_photons(self._ctx, t_start, dt))

    def SwarmExchange(self):
        self._check(self.api.swarm_exchange(self._ctx))

    def SampleDDMCBlockFace(self):
        self._check(self.api.sample_ddmc_block_face(self._ctx))

    def CheckCompletion(self, t_end) -> int:
        n = C.c_int64()
        self._check(self.api.check_completion(self._ctx, t_end, C.byref(n)))
        return n.value

    def EvaluateRadiationEnergy(self):
        self._check(self.api.evaluate_radiation_energy(self._ctx))

    def UpdateFluid(self):
        self._check(self.api.update_fluid(self._ctx))

    def UpdateDerivedMaterial(self):
        self._check(self.api.update_derived_material(self._ctx))

    def PostInitialization(self):
        self._check(self.api.post_initialization(self._ctx))

    def stats_begin(self):
        self._check(self.api.stats_begin(self._ctx))

    def stats(self) -> dict:
        st = jb_cycle_stats()
        self._check(self.api.get_stats(self._ctx, C.byref(st)))
        return st.as_dict()

    # ---- swarm -----------------------------------------------------------------------
    def num_particles(self) -> int:
        n = C.c_int64()
        self._check(self.api.num_particles(self._ctx, C.byref(n)))
        return n.value

    def particles(self) -> dict:
        n = self.num_particles()
        out = {k: np.zeros(n, dtype=np.float64) for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2")}
        out.update({k: np.zeros(n, dtype=np.int32) for k in ("i", "j", "k", "block")})
        out["rng"] = np.zeros((n, 4), dtype=np.uint32)
        p = jb_particles()
        for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2"):
            setattr(p, k, out[k].ctypes.data_as(capi._pd))
        for k in ("i", "j", "k", "block"):
            setattr(p, k, out[k].ctypes.data_as(capi._pi))
        p.rng = out["rng"].ctypes.data_as(capi._pu)
        got = C.c_int64()
        self._check(self.api.get_particles(self._ctx, C.byref(p), n, C.byref(got)))
        assert got.value == n
        return out

    def set_particles(self, d: dict):
        n = len(d["x"])
        keep = {}
        p = jb_particles()
        for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2"):
            if k in d:
                keep[k] = np.ascontiguousarray(d[k], dtype=np.float64)
                setattr(p, k, keep[k].ctypes.data_as(capi._pd))
        keep["block"] = np.ascontiguousarray(d["block"], dtype=np.int32)
        p.block = keep["block"].ctypes.data_as(capi._pi)
        if "rng" in d:
            keep["rng"] = np.ascontiguousarray(d["rng"], dtype=np.uint32)
            p.rng = keep["rng"].ctypes.data_as(capi._pu)
        self._check(self.api.set_particles(self._ctx, C.byref(p), n))

    # ---- event log -------------------------------------------------------------------
    def event_log_enable(self, capacity: int):
        self._check(self.api.event_log_enable(self._ctx, capacity))

    def events(self):
        n = C.c_int64()
        self._check(self.api.event_log_read(self._ctx, None, 0, C.byref(n)))
        buf = (jb_event * max(n.value, 1))()
        self._check(self.api.event_log_read(self._ctx, buf, n.value, C.byref(n)))
        return np.frombuffer(buf, dtype=EVENT_DTYPE, count=n.value).copy()

    # ---- host-staged exchange --------------------------------------------------------
    def exchange_take(self, peer) -> bytes:
        n = C.c_int64()
        self._check(self.api.exchange_count(self._ctx, peer, C.byref(n)))
        rb = int(self.api.exchange_record_bytes())
        buf = C.create_string_buffer(max(n.value, 1) * rb)
        got = C.c_int64()
        self._check(self.api.exchange_take(self._ctx, peer, buf, n.value, C.byref(got)))
        return buf.raw[: got.value * rb]

    def exchange_give(self, payload: bytes):
        rb = int(self.api.exchange_record_bytes())
        n = len(payload) // rb
        if n:
            self._check(self.api.exchange_give(self._ctx, payload, n))

    # ---- device-only extras ----------------------------------------------------------
    def kernel_launches(self) -> int:
        return int(self.api.kernel_launches(self._ctx))

    def timer_start(self):
        self._check(self.api.timer_start(self._ctx))

    def timer_stop(self) -> float:
        ms = C.c_double()
        self._check(self.api.timer_stop(self._ctx, C.byref(ms)))
        return ms.value

    def measure_fp64_peak(self) -> float:
        """Sustained FP64 FMA rate of this GPU in TFLOP/s (register-resident DFMA chains)."""
        v = C.c_double()
        self._check(self.api.measure_fp64_peak(self._ctx, C.byref(v)))
        return v.value

    def clear_particles(self):
        self._check(self.api.set_particles(self._ctx, None, 0))

    def exchange_mode(self) -> int:
        """0 none, 1 NCCL send/recv per iteration, 2 direct peer stores over NVLink (jb_exchange_mode)"""
        return int(self.api.exchange_mode(self._ctx))

    TL_NAMES = {1: "fields", 2: "source", 3: "transport", 4: "exchange", 5: "stats", 6: "squeeze", 7: "tally", 8: "end"}

    def step_timeline(self):
        """[(phase name, ms), ...] of the last RadiationStep in stream order (jb_step_timeline)."""
        cap = 96
        labels, ms, n = (C.c_int32 * cap)(), (C.c_float * cap)(), C.c_int32(0)
        self._check(self.api.step_timeline(self._ctx, labels, ms, cap, C.byref(n)))
        return [(self.TL_NAMES.get(labels[q], str(labels[q])), float(ms[q])) for q in range(min(n.value, cap))]

    def evaluate_radiation_energy_noalloc(self, out=None):
        """(tally [local cells], photons in the pool, their energy) through jb_evaluate_radiation_energy_noalloc:
        nothing is allocated on the device; `out` (float64, >= local cells) is the caller's result buffer."""
        ncell = int(np.prod(self.nxb)) * len(self.local_gids)
        a = out if out is not None else np.empty(ncell, dtype=np.float64)
        n, e = C.c_int64(0), C.c_double(0.0)
        self._check(self.api.evaluate_radiation_energy_noalloc(self._ctx, a.ctypes.data_as(C.c_void_p), a.size,
                                                               C.byref(n), C.byref(e)))
        return a[:ncell], n.value, e.value

    def set_sort_interval(self, every_n_cycles: int):
        self._check(self.api.set_sort_interval(self._ctx, int(every_n_cycles)))

    def reset_streams(self):
        """Restart the sourcing-call counter of the per-photon streams: the next run repeats the last one."""
        self._check(self.api.reset_streams(self._ctx))

    def comm_init(self, unique_id: bytes):
        self._check(self.api.comm_init(self._ctx, unique_id, self.rank, self.n_ranks))


EVENT_DTYPE = np.dtype([
    ("id_idx", "<u4"), ("id_gid", "<u4"), ("id_call", "<u4"), ("draw", "<u4"), ("kind", "<i4"), ("block", "<i4"),
    ("i", "<i4"), ("j", "<i4"), ("k", "<i4"), ("reserved", "<i4"), ("t", "<f8"), ("x", "<f8"), ("y", "<f8"),
    ("z", "<f8"), ("v0", "<f8"), ("v1", "<f8"), ("v2", "<f8"),
])
assert EVENT_DTYPE.itemsize == C.sizeof(jb_event)


def radiation_step_hosted(jb: Jaybenne, t_start: float, dt: float, all_to_all, all_reduce_sum,
                          all_gather_fields=None) -> dict:
    """RadiationStep with the swarm exchange driven from the host (jaybenne.cpp:104-138).

    For transports other than the library's own NCCL path (and for the CPU tests, which run
    it over gloo): `all_to_all(list_of_bytes_per_peer) -> list_of_bytes_from_each_peer` and
    `all_reduce_sum(int) -> int` are supplied by the caller.
    """
    jb.stats_begin()
    if jb.cfg.use_ddmc and jb.n_ranks > 1 and all_gather_fields is not None:
        # the ghost exchange of density / sie that precedes the face probabilities
        # (mcblock_driver.cpp:68, jaybenne.cpp:108-110): whole blocks, every rank gets every block
        mine = {g: (jb.get_field(g, "density"), jb.get_field(g, "sie")) for g in jb.local_gids}
        for g, (rho, sie) in all_gather_fields(mine).items():
            if g not in mine:
                jb.set_remote_field(g, "density", rho)
                jb.set_remote_field(g, "sie", sie)
    jb.UpdateDerivedTransportFields(dt)
    jb.SourcePhotons(JB_SOURCE_EMISSION, t_start, dt)
    it = 0
    while True:
        it += 1
        if it > jb.cfg.max_transport_iterations:
            raise JbError(capi.JB_ERR_ITERATIONS, "max_transport_iterations exhausted")
        jb.TransportPhotons(t_start, dt)
        jb.SwarmExchange()
        outgoing = [jb.exchange_take(p) if p != jb.rank else b"" for p in range(jb.n_ranks)]
        for payload in all_to_all(outgoing):
            jb.exchange_give(payload)
        if jb.cfg.use_ddmc:
            jb.SampleDDMCBlockFace()
        if all_reduce_sum(jb.CheckCompletion(t_start + dt)) == 0:
            break
    jb.EvaluateRadiationEnergy()
    jb.UpdateFluid()
    st = jb.stats()
    st["n_iterations"] = it
    return st
