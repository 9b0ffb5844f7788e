"""TEST INFRASTRUCTURE: the cases behind tests/golden/*.json.

Every function takes an oracle `Api` (tests/oracle_binding.py) and returns a JSON-able
table. tests/golden/make_golden.py evaluates them with the reference-header flavour of the
oracle and commits the result; tests/test_oracle_pin.py evaluates them with the restated
flavours and compares. Floating-point values are stored as IEEE-754 hex strings.
"""
import ctypes as C
import hashlib

import numpy as np

from jaybenne_b200 import Mcblock, capi
from jaybenne_b200.capi import jb_config

C_LIGHT = 2.99792458e10


def fhex(x):
    return float(x).hex()


def unhex(s):
    return float.fromhex(s)


# ------------------------------------------------------------------------------------------
# closures (singularity-opac / singularity-eos through src/mcblock/opacity.hpp, eos.hpp)
# ------------------------------------------------------------------------------------------
CLOSURE_CONFIGS = {
    # name: (opacity_model, kappa, scattering_model, sigma, apm, (time, mass, length, temperature) scales)
    "gray_cgs": (capi.JB_OPAC_GRAY, 1.0, capi.JB_SCAT_GRAY, 1.0e5, 1.0, (1.0, 1.0, 1.0, 1.0)),
    "gray_inf_stiff": (capi.JB_OPAC_GRAY, 1000.0, capi.JB_SCAT_GRAY, 0.0, 1.0, (1.0, 1.0, 1.0, 1.0)),
    # the unit system of singularity-opac/test/test_gray_opacities.cpp:304-334
    "gray_noncgs": (capi.JB_OPAC_GRAY, 1.0, capi.JB_SCAT_GRAY, 2.5, 3.0, (123.0, 456.0, 789.0, 276.0)),
    "epbremss_thomson": (capi.JB_OPAC_EPBREMSS, 0.0, capi.JB_SCAT_THOMSON, 0.0, 1.6726e-24, (1.0, 1.0, 1.0, 1.0)),
    "epbremss_noncgs": (capi.JB_OPAC_EPBREMSS, 0.0, capi.JB_SCAT_THOMSON, 0.0, 2.0e-24, (2.0, 3.0, 5.0, 7.0)),
}
CLOSURE_POINTS = [
    # (rho, temperature, nu)
    (1.0, 1.0, 1.0), (1.0, 1.0, 1.0e10), (1.0, 1.0, 1.0e15), (1.0, 1.0e5, 5.670e-5), (1.1, 0.97, 3.3e-4),
    (0.9, 1.0e-5, 2.0e-9), (1.0e3, 1.0e6, 1.0e17), (1.0e-3, 1.0e4, 1.0e14), (2.5, 3.0e7, 6.0e18), (1.0, 1.0e-5, 1.0),
]


def make_config(api, model):
    om, kappa, sm, sigma, apm, scales = CLOSURE_CONFIGS[model]
    cfg = jb_config()
    api.config_default(C.byref(cfg))
    cfg.num_particles = 1
    cfg.opacity_model, cfg.opacity_kappa = om, kappa
    cfg.scattering_model, cfg.scattering_sigma, cfg.apm = sm, sigma, apm
    cfg.time_scale, cfg.mass_scale, cfg.length_scale, cfg.temperature_scale = scales
    return cfg


def closure_table(api):
    out = {}
    for model in CLOSURE_CONFIGS:
        cfg = make_config(api, model)
        ctx = capi.ctx_p()
        assert api.create(C.byref(cfg), C.byref(ctx)) == 0
        a, b = C.c_double(), C.c_double()
        api.get_constants(ctx, C.byref(a), C.byref(b))
        rows = []
        for rho, temp, nu in CLOSURE_POINTS:
            sie = cfg.eos_cv * temp
            rows.append({
                "rho": fhex(rho), "temp": fhex(temp), "nu": fhex(nu),
                "absorption": fhex(api.probe_absorption(ctx, rho, temp, nu)),
                "scattering": fhex(api.probe_scattering(ctx, rho, temp, nu)),
                "emissivity": fhex(api.probe_emissivity(ctx, rho, temp)),
                "temperature": fhex(api.probe_temperature(ctx, rho, sie)),
                "specific_heat": fhex(api.probe_specific_heat(ctx, rho, sie)),
            })
        out[model] = {"speed_of_light": fhex(a.value), "stefan_boltzmann": fhex(b.value), "cv": fhex(cfg.eos_cv),
                      "points": rows}
        api.destroy(ctx)
    out["eps_imc_offset"] = fhex(api.probe_eps(0))
    out["eps_ddmc_offset"] = fhex(api.probe_eps(1))
    return out


# ------------------------------------------------------------------------------------------
# single calls of the per-particle functions on seeded random inputs
# ------------------------------------------------------------------------------------------
def _dir(rng, ndim):
    mu = 2 * rng.random() - 1
    phi = 2 * np.pi * rng.random()
    s = np.sqrt(1 - mu * mu)
    return C_LIGHT * s * np.cos(phi), C_LIGHT * s * np.sin(phi), C_LIGHT * mu


def step_inputs(seed=20261104, n_per_dim=40):
    """Inputs for probe_imc_step / probe_ddmc_step covering census, face crossings (all six
    faces), effective scatters and absorptions, thick and thin cells."""
    rng = np.random.default_rng(seed)
    cases = []
    for ndim in (1, 2, 3):
        for q in range(n_per_dim):
            d = 10.0 ** rng.uniform(-3, 0)
            w = [d * rng.uniform(0.5, 2.0) if k < ndim else 1.0 for k in range(3)]
            lo = [rng.uniform(-1, 1) if k < ndim else -0.5 for k in range(3)]
            hi = [lo[k] + w[k] for k in range(3)]
            pos = [lo[k] + w[k] * rng.random() for k in range(3)]
            if q % 5 == 4:  # start exactly one nudge inside a face (as after a crossing)
                k = int(rng.integers(0, ndim))
                pos[k] = lo[k] + 2.220446049250313e-09 * w[k] if rng.random() < 0.5 else hi[k] - 2.220446049250313e-09 * w[k]
            vx, vy, vz = _dir(rng, ndim)
            tau = 10.0 ** rng.uniform(-2, 2)     # optical depth of the cell
            dmin = min(w[:ndim])
            ff = rng.random() if q % 3 else 1.0
            aa = tau / dmin * rng.random()
            ss = tau / dmin - aa
            dt = dmin / C_LIGHT * 10.0 ** rng.uniform(-1, 1.5)
            cases.append({"ndim": ndim, "lo": lo, "hi": hi, "pos": pos, "v": [vx, vy, vz], "ff": ff, "aa": aa,
                          "ss": ss, "dt": dt, "dx_push": dmin, "idx": int(rng.integers(0, 2**31)),
                          "P": [float(v) for v in rng.uniform(0.01, 0.6, 6)]})
    return cases


def run_imc_case(api, seed, c):
    lo = (C.c_double * 3)(*c["lo"]); hi = (C.c_double * 3)(*c["hi"])
    io = (C.c_double * 7)(0.0, *c["pos"], *c["v"])
    flags = (C.c_int32 * 2)()
    n = api.probe_imc_step(seed, c["idx"], 0.0, c["dt"], c["ff"], c["aa"], c["ss"], C_LIGHT, c["dx_push"], c["ndim"],
                           lo, hi, io, flags)
    return {"draws": int(n), "flags": [int(flags[0]), int(flags[1])], "io": [fhex(v) for v in io]}


def run_ddmc_case(api, seed, c, on_face):
    lo = (C.c_double * 3)(*c["lo"]); hi = (C.c_double * 3)(*c["hi"])
    pos = list(c["pos"])
    if on_face:  # arrive on the lower x face from an IMC cell: triggers ptcl_ddmc_albedo
        pos[0] = c["lo"][0] + 2.220446049250313e-09 * (c["hi"][0] - c["lo"][0])
    io = (C.c_double * 7)(0.0, *pos, *c["v"])
    P = (C.c_double * 6)(*[p if (k % 3) < c["ndim"] else 0.0 for k, p in enumerate(c["P"])])
    ijk = (C.c_int32 * 3)(5, 5 if c["ndim"] > 1 else 0, 5 if c["ndim"] > 2 else 0)
    flags = (C.c_int32 * 2)()
    n = api.probe_ddmc_step(seed, c["idx"], 0.0, c["dt"], c["ff"], c["aa"], c["ss"], C_LIGHT, c["ndim"], lo, hi, P, io,
                            ijk, flags)
    return {"draws": int(n), "flags": [int(flags[0]), int(flags[1])], "ijk": [int(v) for v in ijk],
            "io": [fhex(v) for v in io]}


def step_table(api, seed=349857):
    cases = step_inputs()
    out = {"seed": seed, "imc": [], "ddmc": [], "ddmc_on_face": [], "planck": []}
    for c in cases:
        out["imc"].append(run_imc_case(api, seed, c))
        out["ddmc"].append(run_ddmc_case(api, seed, c, False))
        out["ddmc_on_face"].append(run_ddmc_case(api, seed, c, True))
    for idx in range(64):
        temp = 10.0 ** (idx % 7 - 3)
        out["planck"].append(fhex(api.probe_planck(seed, idx, 5.6703726225913318e-05, temp)))
    return out


# ------------------------------------------------------------------------------------------
# whole runs of the shipped decks
# ------------------------------------------------------------------------------------------
RUNS = {
    # name: (deck, ci overrides?, extra overrides, cycles)
    "inf": ("inf", False, [], 2),
    "inf_stiff_ddmc": ("inf_stiff", False, [("jaybenne", "num_particles", 400)], 2),
    "stepdiff_ci": ("stepdiff", True, [("jaybenne", "num_particles", 1500)], 2),
    "stepdiff_shipped_2blocks": ("stepdiff", False, [("jaybenne", "num_particles", 1500)], 2),
    "stepdiff_ddmc_ci": ("stepdiff_ddmc", True, [("jaybenne", "num_particles", 3000)], 2),
    "stepdiff_smr_ci": ("stepdiff_smr", True, [("jaybenne", "num_particles", 3000)], 2),
    "stepdiff_smr_ddmc_ci": ("stepdiff_smr_ddmc", True, [("jaybenne", "num_particles", 6000)], 2),
    "stepdiff_smr_hybrid_shipped": ("stepdiff_smr_hybrid", False, [("jaybenne", "num_particles", 6000)], 1),
    # three-dimensional static refinement (jaybenne_b200/decks.py EXTRA_DECKS): level-changing crossings in 3D,
    # SampleDDMCBlockFace::3D / SampleFace3D (n_bface_resampled > 0)
    "smr3d_imc": ("smr3d_imc", False, [], 2),
    "smr3d_ddmc": ("smr3d_ddmc", False, [], 2),
    "smr3d_hybrid": ("smr3d_hybrid", False, [("jaybenne", "num_particles", 12000), ("jaybenne", "dt", "5.0e-12")], 2),
}
COUNT_KEYS = ("n_start", "n_sourced", "n_absorbed", "n_escaped", "n_end", "n_segments", "n_scatters",
              "n_ddmc_segments", "n_block_crossings", "n_bface_resampled")
ENERGY_KEYS = ("e_start", "e_sourced", "e_absorbed", "e_escaped", "e_end")


def swarm_digest(jb):
    """sha256 over the census swarm sorted by stream identity (bit-level pin)."""
    p = jb.particles()
    rid = p["rng"]
    order = np.lexsort((rid[:, 0], rid[:, 1], rid[:, 2]))
    h = hashlib.sha256()
    for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2", "i", "j", "k", "block"):
        h.update(np.ascontiguousarray(p[k][order]).tobytes())
    h.update(np.ascontiguousarray(rid[order]).tobytes())
    return h.hexdigest()


def run_case(api, name):
    import parity_util as pu
    deck, ci, overrides, cycles = RUNS[name]
    pin = pu.make_pin(deck, overrides=overrides, ci=ci)
    mb = Mcblock(pin, api=api)
    mb.ProblemGenerator()
    rec = {"nblocks": mb.mesh.nblocks, "n_init": mb.jb.num_particles(), "cycles": []}
    for _ in range(cycles):
        st = mb.Step()
        mb.time += mb.dt
        row = {k: int(st[k]) for k in COUNT_KEYS}
        row.update({k: fhex(st[k]) for k in ENERGY_KEYS})
        tally = np.concatenate([mb.jb.get_field(g, "energy_tally").ravel() for g in mb.jb.local_gids])
        row["tally_sum"] = fhex(tally.sum())
        row["tally_sha256"] = hashlib.sha256(np.ascontiguousarray(tally).tobytes()).hexdigest()
        row["swarm_sha256"] = swarm_digest(mb.jb)
        rec["cycles"].append(row)
    mb.jb.close()
    return rec


def run_table(api, names=None):
    return {name: run_case(api, name) for name in (names or RUNS)}
