"""The seven problem set-ups the reference ships under inputs/ (inf, inf_stiff, stepdiff,
stepdiff_ddmc, stepdiff_smr, stepdiff_smr_ddmc, stepdiff_smr_hybrid), stated as data and
rendered to Parthenon deck text on demand — `/root/reference` does not exist on the GPU
box, so the tests and the benchmark build their inputs from here. tests/test_decks.py
checks every key/value against the reference's own files when the reference is mounted.
"""
from __future__ import annotations

_OUT = {"file_type": "hdf5",
        "variables": "field.material.density,field.material.sie,field.material.internal_energy,"
                     "field.jaybenne.energy_tally"}
_JB_COMMON = {"numin": "1.e10", "numax": "1.e20", "transport_model": "zone_size", "tracking_algo": "history",
              "source_strategy": "uniform"}


def _mesh(nx, xmin, xmax, bcs, **extra):
    m = dict(extra)
    for d in range(3):
        m[f"nx{d+1}"] = nx[d]
        m[f"x{d+1}min"] = xmin[d]
        m[f"x{d+1}max"] = xmax[d]
        m[f"ix{d+1}_bc"] = bcs[d]
        m[f"ox{d+1}_bc"] = bcs[d]
    return m


def _stepdiff(nx, xmin, xmax, nxb, jb_extra=None, refine=None, swarm_out=True):
    blocks = {
        "parthenon/job": {"problem_id": "stepdiff"},
        "parthenon/mesh": _mesh(nx, xmin, xmax, ("outflow", "periodic", "periodic"),
                                **({"nghost": 2, "refinement": "static"} if refine else {"refinement": "none"})),
        "parthenon/swarm": {f"{s}x{d+1}_bc": ("jaybenne_reflecting" if d == 0 else "periodic")
                            for d in range(3) for s in "io"},
        "parthenon/meshblock": {f"nx{d+1}": nxb[d] for d in range(3)},
    }
    if refine:
        blocks["parthenon/static_refinement1"] = refine
    blocks["parthenon/time"] = {"tlim": "3.335641e-10", "integrator": "rk1"}
    jb = dict(jb_extra or {})
    jb.update({"num_particles": 100000, "dt": "3.335641e-11", "do_emission": "false", "do_feedback": "false",
               "seed": 349857})
    jb.update(_JB_COMMON)
    blocks["jaybenne"] = jb
    blocks["mcblock"] = {"opacity_model": "none", "scattering_model": "constant", "scattering_constant_value": "1.0e3",
                         "specific_heat": "1.0e8", "initial_density": "1.0", "initial_temperature": "1.0e5",
                         "initial_radiation": "thermal"}
    out = dict(_OUT, dt="3.335641e-10")
    if swarm_out:
        out.update({"swarms": "photons", "swarm_variables": "swarm.x, swarm.y"})
    blocks["parthenon/output0"] = out
    return blocks


def _inf(problem_id, nx, jb, mc, tlim, out_dt):
    j = dict(jb)
    j.update(_JB_COMMON)
    m = {"opacity_model": "constant", "scattering_model": "constant", "initial_density": "1.0",
         "initial_temperature": "1.0", "initial_radiation": "thermal"}
    m.update(mc)
    return {
        "parthenon/job": {"problem_id": problem_id},
        "parthenon/mesh": _mesh(nx, ("-0.5",) * 3, ("0.5",) * 3, ("periodic",) * 3, refinement="none"),
        "parthenon/time": {"tlim": tlim, "integrator": "rk1"},
        "jaybenne": j,
        "mcblock": m,
        "parthenon/output0": dict(_OUT, dt=out_dt),
    }


def _smr_box(z):
    return {"level": 1, "x1min": "-0.25", "x1max": "0.25", "x2min": "-0.25", "x2max": "0.25", "x3min": f"-{z}",
            "x3max": z}


DECKS = {
    "inf": _inf("inf", (4, 4, 4),
                {"num_particles": 200, "dt": "1.e-12", "do_emission": "true", "do_feedback": "false", "seed": 349857},
                {"opacity_constant_value": "1.0", "scattering_constant_value": "1.0e5", "specific_heat": "1.0e8"},
                "1.e-10", "0.1"),
    "inf_stiff": _inf("inf_stiff", (4, 1, 1),
                      {"use_ddmc": "true", "num_particles": 10000, "dt": "1.0e-11", "do_emission": "true",
                       "do_feedback": "false", "seed": 349856},
                      {"opacity_constant_value": "1000.0", "scattering_constant_value": "0.0", "specific_heat": "1.0e7"},
                      "1e-10", "1e-11"),
    "stepdiff": _stepdiff((100, 1, 1), ("-0.5",) * 3, ("0.5",) * 3, (50, 1, 1), swarm_out=False),
    "stepdiff_ddmc": _stepdiff((100, 1, 1), ("-0.5",) * 3, ("0.5",) * 3, (50, 1, 1), {"use_ddmc": "true"}),
    "stepdiff_smr": _stepdiff((128, 64, 1), ("-0.5", "-0.25", "-0.5"), ("0.5", "0.25", "0.5"), (32, 32, 1),
                              refine=_smr_box("0.5")),
    "stepdiff_smr_ddmc": _stepdiff((128, 64, 1), ("-0.5", "-0.25", "-0.25"), ("0.5", "0.25", "0.25"), (32, 32, 1),
                                   {"use_ddmc": "true", "tau_ddmc": "2.5"}, refine=_smr_box("0.25")),
    "stepdiff_smr_hybrid": _stepdiff((128, 64, 1), ("-0.5", "-0.25", "-0.25"), ("0.5", "0.25", "0.25"), (32, 32, 1),
                                     {"use_ddmc": "true"}, refine=_smr_box("0.25")),
}

# the overrides the reference's regression scripts apply (tst/stepdiff.py:28-30, tst/stepdiff_smr.py:28-32)
CI_OVERRIDES = {
    "stepdiff": [("parthenon/mesh", "nx1", 128), ("parthenon/meshblock", "nx1", 128)],
    "stepdiff_smr": [("parthenon/mesh", "nx1", 64), ("parthenon/mesh", "nx2", 32), ("parthenon/meshblock", "nx1", 16),
                     ("parthenon/meshblock", "nx2", 16)],
}
CI_OVERRIDES["stepdiff_ddmc"] = CI_OVERRIDES["stepdiff"]
CI_OVERRIDES["stepdiff_smr_ddmc"] = CI_OVERRIDES["stepdiff_smr"]
CI_OVERRIDES["stepdiff_smr_hybrid"] = CI_OVERRIDES["stepdiff_smr"]



def _smr3d(jb_extra, dt, sigma, particles):
    """32^3 root grid in 8^3-cell meshblocks with the central [-0.25, 0.25]^3 refined once: 56 coarse + 64 fine
    blocks, every face / edge / corner of the fine region a level change in three dimensions."""
    d = _stepdiff((32, 32, 32), ("-0.5",) * 3, ("0.5",) * 3, (8, 8, 8), jb_extra, refine=_smr_box("0.25"))
    d["jaybenne"].update({"num_particles": particles, "dt": dt})
    d["parthenon/time"]["tlim"] = dt
    d["mcblock"]["scattering_constant_value"] = sigma
    return d


# NOT reference inputs: the reference ships no three-dimensional refined deck, but its 3D code paths
# (SampleDDMCBlockFace::3D / SampleFace3D, sample_ddmc_bface.cpp:41-78,241-423, and level-changing
# neighbour look-ups in three dimensions, transport.cpp:146-155) are on the hot path. Same physics as
# stepdiff_smr*, on a 3D static-refined mesh: pure IMC, every cell DDMC, and a true hybrid
# (coarse cells tau = 31 > tau_ddmc = 20 > 15.6 = tau of the fine cells).
EXTRA_DECKS = {
    "smr3d_imc": _smr3d(None, "3.335641e-12", "1.0e2", 8000),
    "smr3d_ddmc": _smr3d({"use_ddmc": "true", "tau_ddmc": "2.5"}, "3.335641e-11", "1.0e3", 40000),
    "smr3d_hybrid": _smr3d({"use_ddmc": "true", "tau_ddmc": "20.0"}, "1.0e-11", "1.0e3", 40000),
}


def deck_text(name: str) -> str:
    lines = []
    for block, kv in (DECKS[name] if name in DECKS else EXTRA_DECKS[name]).items():
        lines.append(f"<{block}>")
        lines += [f"{k} = {v}" for k, v in kv.items()]
        lines.append("")
    return "\n".join(lines)
