"""mcblock-lite: the smallest host that drives the jaybenne package the way the reference's
driver application does, so the shipped input decks run unchanged.

    ParameterInput          Parthenon's deck reader (C++: csrc/host/deck.cpp)
    Mcblock.__init__        mcblock::Initialize + jaybenne::Initialize     src/mcblock/mcblock.cpp:37-150,288-297
    Mcblock.ProblemGenerator  mcblock::ProblemGenerator + PostInitialization  mcblock.cpp:155-262
    Mcblock.Step            McblockDriver::Step + HostUpdateTasks          src/mcblock/mcblock_driver.cpp:38-74
    Mcblock.Execute         EvolutionDriver::Execute (fixed dt = <jaybenne> dt, clipped at tlim)
    Mcblock.dump            the fields of <parthenon/output0> as numpy arrays + block bounds,
                            laid out like analysis/jhdf.py:36-73 exposes them

Material fields live in the CUDA library; this file only fills initial conditions and
sequences calls. It is generic over the `Api` (see package.py) for the same reason.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import JB_OK, JbError, jb_config, jb_mesh
from .package import Jaybenne


class ParameterInput:
    """Parthenon-style input deck (blocks of key = value) backed by the C++ reader."""

    def __init__(self, path=None, text=None, api=None):
        self.api = api or capi.host_api()
        self._deck = capi.deck_p()
        if path is not None:
            st = self.api.deck_load(str(path).encode(), C.byref(self._deck))
        else:
            st = self.api.deck_parse((text or "").encode(), C.byref(self._deck))
        if st != JB_OK:
            raise JbError(st, f"cannot read input deck {path or '<text>'}")

    def __del__(self):
        try:
            if self._deck:
                self.api.deck_free(self._deck)
                self._deck = capi.deck_p()
        except Exception:
            pass

    def set(self, block, key, value):
        """Override like `mcblock -i deck block/key=value` (tst/regression_test.py modified_inputs)."""
        self.api.deck_set(self._deck, block.encode(), key.encode(), str(value).encode())

    def has(self, block, key):
        return bool(self.api.deck_has(self._deck, block.encode(), key.encode()))

    def get(self, block, key, default=None):
        v = self.api.deck_get(self._deck, block.encode(), key.encode())
        return v.decode() if v is not None else default

    def get_real(self, block, key, default=None):
        v = self.get(block, key)
        return float(v) if v is not None else default

    def config(self) -> jb_config:
        cfg = jb_config()
        err = C.create_string_buffer(512)
        st = self.api.deck_config(self._deck, C.byref(cfg), err, 512)
        if st != JB_OK:
            raise JbError(st, err.value.decode())
        return cfg

    def mesh(self, n_ranks=1) -> "Mesh":
        m = jb_mesh()
        err = C.create_string_buffer(512)
        st = self.api.deck_mesh(self._deck, n_ranks, C.byref(m), err, 512)
        if st != JB_OK:
            raise JbError(st, err.value.decode())
        return Mesh(m, self.api)


class Mesh:
    """Owns a jb_mesh built by jb_deck_mesh (block table in Z-order, contiguous rank chunks)."""

    def __init__(self, m: jb_mesh, api):
        self.c, self.api = m, api

    def __del__(self):
        try:
            self.api.mesh_free(C.byref(self.c))
        except Exception:
            pass

    @property
    def nblocks(self):
        return self.c.nblocks

    def block(self, gid):
        return self.c.blocks[gid]

    def cell_centers(self, gid):
        """Interior cell-centre coordinates of a block per direction (UniformCartesian::Xc)."""
        b, ng = self.c.blocks[gid], self.c.nghost
        out = []
        for d in range(3):
            n = self.c.nxb[d]
            dx = (b.xmax[d] - b.xmin[d]) / n
            g = ng if d < self.c.ndim else 0
            xmin0 = b.xmin[d] - g * dx
            out.append(xmin0 + ((np.arange(n) + g) + 0.5) * dx)
        return out


def uniform_deck_text(nx, nxb, xmin=(0.0, 0.0, 0.0), xmax=(1.0, 1.0, 1.0), bc="periodic") -> str:
    """Deck text of a uniform mesh (used for the synthetic benchmark configuration)."""
    lines = ["<parthenon/mesh>", "refinement = none"]
    for d in range(3):
        lines += [f"nx{d+1} = {nx[d]}", f"x{d+1}min = {xmin[d]!r}", f"x{d+1}max = {xmax[d]!r}",
                  f"ix{d+1}_bc = {bc}", f"ox{d+1}_bc = {bc}"]
    lines += ["<parthenon/meshblock>"] + [f"nx{d+1} = {nxb[d]}" for d in range(3)]
    lines += ["<parthenon/time>", "integrator = rk1"]
    return "\n".join(lines) + "\n"


class Mcblock:
    def __init__(self, pin: ParameterInput, api=None, rank=0, n_ranks=1, capacity=0, cfg_overrides=None):
        self.pin = pin
        self.api = api or capi.device_api()
        self.cfg = pin.config()
        for k, v in (cfg_overrides or {}).items():
            setattr(self.cfg, k, v)
        self.mesh = pin.mesh(n_ranks)
        self.rank, self.n_ranks = rank, n_ranks
        self.problem_id = pin.get("parthenon/job", "problem_id", "")
        self.rho0 = pin.get_real("mcblock", "initial_density")
        self.tt0 = pin.get_real("mcblock", "initial_temperature")
        self.initial_radiation = pin.get("mcblock", "initial_radiation")
        self.tlim = pin.get_real("parthenon/time", "tlim", float("inf"))
        self.nlim = int(pin.get_real("parthenon/time", "nlim", -1))
        self.jb = Jaybenne.Initialize(self.api, self.cfg, self.mesh.c, rank=rank, n_ranks=n_ranks, capacity=capacity)
        self.time, self.ncycle = 0.0, 0
        self.dt = self.cfg.dt
        self.history = []

    # mcblock::ProblemGenerator on every local block, then PostInitialization
    def ProblemGenerator(self, density_fn=None):
        cv = self.cfg.eos_cv
        n1, n2, n3 = self.jb.nxb
        for gid in self.jb.local_gids:
            rho = np.full((n3, n2, n1), self.rho0)
            sie = np.full((n3, n2, n1), cv * self.tt0)
            xc = self.mesh.cell_centers(gid)
            if self.problem_id == "stepdiff":
                ttlow = 1.0e-5 * self.tt0
                sie[:, :, xc[0] >= 0.0] = cv * ttlow
            if density_fn is not None:
                rho = density_fn(xc[0][None, None, :], xc[1][None, :, None], xc[2][:, None, None]) + 0.0 * rho
            self.jb.set_field(gid, "density", rho)
            self.jb.set_field(gid, "sie", sie)
        self.jb.InitializeRadiation(self.initial_radiation == "thermal")
        self.jb.PostInitialization()

    # McblockDriver::Step
    def Step(self, step_fn=None):
        st = (step_fn or self.jb.RadiationStep)(self.time, self.dt)
        # HostUpdateTasks: (ghost exchange) -> FillDerived (sie = u / rho) -> EstimateTimestep
        self.jb.UpdateDerivedMaterial()
        self.history.append(st)
        return st

    # EvolutionDriver::Execute
    def Execute(self, max_cycles=None, step_fn=None):
        while self.time < self.tlim and (self.nlim < 0 or self.ncycle < self.nlim):
            if max_cycles is not None and self.ncycle >= max_cycles:
                break
            self.Step(step_fn)
            self.time += self.dt
            self.ncycle += 1
            self.dt = self.cfg.dt
            if self.time < self.tlim and (self.tlim - self.time) < self.dt:
                self.dt = self.tlim - self.time
        return self.history

    def dump(self, fields=("density", "sie", "internal_energy", "energy_tally")):
        """Per-field arrays [block][k][j][i] over the local blocks + block bounds."""
        out = {"gids": np.array(self.jb.local_gids, dtype=np.int32), "time": self.time, "ncycle": self.ncycle}
        for name in fields:
            out[name] = np.stack([self.jb.get_field(g, name) for g in self.jb.local_gids])
        out["xmin"] = np.array([[self.mesh.block(g).xmin[d] for d in range(3)] for g in self.jb.local_gids])
        out["xmax"] = np.array([[self.mesh.block(g).xmax[d] for d in range(3)] for g in self.jb.local_gids])
        out["level"] = np.array([self.mesh.block(g).level for g in self.jb.local_gids])
        return out
