"""Shared helpers for the GPU-vs-oracle parity tests (TEST INFRASTRUCTURE)."""
import os

import numpy as np

from jaybenne_b200 import Mcblock, ParameterInput, device_api, uniform_deck_text
from oracle_binding import oracle_api

from jaybenne_b200.decks import CI_OVERRIDES, deck_text

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

COUNT_KEYS = ("n_start", "n_sourced", "n_absorbed", "n_escaped", "n_end", "n_segments", "n_scatters",
              "n_ddmc_segments", "n_block_crossings", "n_bface_resampled")
ENERGY_KEYS = ("e_start", "e_sourced", "e_absorbed", "e_escaped", "e_end")


def make_pin(deck=None, text=None, overrides=(), ci=False):
    pin = ParameterInput(text=deck_text(deck) if deck else text)
    for blk, key, val in (CI_OVERRIDES.get(deck, []) if ci else []):
        pin.set(blk, key, val)
    for blk, key, val in overrides:
        pin.set(blk, key, val)
    return pin


def make_pair(deck=None, text=None, overrides=(), cfg_overrides=None, flavour="portable", capacity=0, ci=False):
    """(gpu Mcblock, oracle Mcblock) on the same deck."""
    out = []
    for api in (device_api(), oracle_api(flavour)):
        pin = make_pin(deck, text, overrides, ci)
        out.append(Mcblock(pin, api=api, capacity=capacity if api.prefix == "jb_" else 0,
                           cfg_overrides=cfg_overrides))
    return out


def sorted_particles(jb):
    p = jb.particles()
    rid = p["rng"]
    order = np.lexsort((rid[:, 0], rid[:, 1], rid[:, 2]))
    return {k: v[order] for k, v in p.items()}


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    scale = np.maximum(np.abs(a), np.abs(b))
    scale[scale == 0] = 1.0
    return np.max(np.abs(a - b) / scale) if a.size else 0.0


def assert_swarm_equal(gpu, ora, tol=1e-12, exact=True):
    g, o = sorted_particles(gpu), sorted_particles(ora)
    assert len(g["x"]) == len(o["x"]), (len(g["x"]), len(o["x"]))
    assert np.array_equal(g["rng"], o["rng"]), "stream identities / draw counts differ"
    for k in ("i", "j", "k", "block"):
        assert np.array_equal(g[k], o[k]), k
    for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2"):
        if exact:
            assert np.array_equal(g[k], o[k]), (k, rel_err(g[k], o[k]))
        else:
            assert rel_err(g[k], o[k]) <= tol, (k, rel_err(g[k], o[k]))


def assert_stats_equal(sg, so, tol=1e-12):
    for k in COUNT_KEYS:
        assert sg[k] == so[k], (k, sg[k], so[k])
    for k in ENERGY_KEYS:
        assert rel_err([sg[k]], [so[k]]) <= tol, (k, sg[k], so[k])


def assert_fields_equal(gpu_mb, ora_mb, names=("energy_tally", "energy_delta", "fleck_factor", "internal_energy", "sie"),
                        tol=1e-12, gross=None):
    """Fields are sums whose order differs (atomics): 1e-12 relative to the largest value, and for
    energy_delta — emitted minus absorbed, which cancel in stiff problems — relative to the gross
    energy `gross` that went through a cell (the quantity rounding errors scale with)."""
    for gid in gpu_mb.jb.local_gids:
        for name in names:
            a, b = gpu_mb.jb.get_field(gid, name), ora_mb.jb.get_field(gid, name)
            scale = max(np.max(np.abs(b)), 1e-300)
            if name == "energy_delta" and gross:
                scale = max(scale, gross)
            err = np.max(np.abs(a - b)) / scale
            assert err <= tol, (gid, name, err)


def conservation_residual(st):
    """|E_end - E_start - E_sourced + E_absorbed + E_escaped| / E"""
    e = max(abs(st["e_end"]), abs(st["e_start"]), 1e-300)
    return abs(st["e_end"] - st["e_start"] - st["e_sourced"] + st["e_absorbed"] + st["e_escaped"]) / e


def synthetic_text(n=16, nb=8, num_particles=20000, kappa=None, sigma=None, dt=None, ndim=3, seed=349857,
                   do_emission=True, do_feedback=True, extra=""):
    """Scaled-down twin of the synthetic 3D gray benchmark."""
    nx = [n if d < ndim else 1 for d in range(3)]
    nxb = [nb if d < ndim else 1 for d in range(3)]
    delta = 1.0 / n
    c = 2.99792458e10
    kappa = 0.1 / delta if kappa is None else kappa
    sigma = 0.5 / delta if sigma is None else sigma
    dt = 4 * delta / c if dt is None else dt
    t = uniform_deck_text(nx, nxb)
    t += f"""<parthenon/job>
problem_id = synthetic
<parthenon/time>
integrator = rk1
tlim = 1.0
<jaybenne>
num_particles = {num_particles}
dt = {dt!r}
seed = {seed}
do_emission = {'true' if do_emission else 'false'}
do_feedback = {'true' if do_feedback else 'false'}
<mcblock>
opacity_model = constant
opacity_constant_value = {kappa!r}
scattering_model = constant
scattering_constant_value = {sigma!r}
initial_density = 1.0
initial_temperature = 1.0
initial_radiation = thermal
{extra}
"""
    return t


def synthetic_density(x, y, z):
    return 1.0 + 0.1 * np.sin(2 * np.pi * x) * np.sin(2 * np.pi * y) * np.sin(2 * np.pi * z)
