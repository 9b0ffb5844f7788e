"""GPU parity tests: the CUDA path through the C ABI against the CPU oracle on the same
seeded inputs. Integer results (counts, draw counters, cell indices) must be bit-exact;
photon state is compared bit for bit too (the kernels and the oracle share jb_math.h and
are both built without FMA contraction); cell tallies, which are sums whose order differs
(atomics), must agree to 1e-12 relative — the tolerance BASELINE.json's north_star names.
"""
import numpy as np
import pytest

import parity_util as pu

pytestmark = pytest.mark.gpu


def run_pair(gpu, ora, cycles, density_fn=None, fields=True, exact=True):
    gpu.ProblemGenerator(density_fn)
    ora.ProblemGenerator(density_fn)
    pu.assert_swarm_equal(gpu.jb, ora.jb, exact=exact)
    if fields:
        pu.assert_fields_equal(gpu, ora, names=("energy_tally", "energy_delta", "internal_energy"))
    for c in range(cycles):
        sg, so = gpu.Step(), ora.Step()
        gpu.time += gpu.dt; ora.time += ora.dt
        pu.assert_stats_equal(sg, so)
        assert pu.conservation_residual(sg) <= 1e-12, pu.conservation_residual(sg)
        assert abs(pu.conservation_residual(sg) - pu.conservation_residual(so)) <= 1e-12
        pu.assert_swarm_equal(gpu.jb, ora.jb, exact=exact)
        if fields:
            pu.assert_fields_equal(gpu, ora)
    return sg, so


def test_inf_deck_counts_state_and_fields():
    gpu, ora = pu.make_pair("inf")
    sg, _ = run_pair(gpu, ora, cycles=3)
    assert sg["n_segments"] > 100000


def test_inf_deck_event_by_event():
    gpu, ora = pu.make_pair("inf", overrides=[("jaybenne", "num_particles", 64)])
    gpu.ProblemGenerator(); ora.ProblemGenerator()
    cap = 400000
    gpu.jb.event_log_enable(cap); ora.jb.event_log_enable(cap)
    gpu.Step(); ora.Step()
    eg, eo = gpu.jb.events(), ora.jb.events()
    assert len(eg) == len(eo) and 0 < len(eg) < cap
    key = lambda e: np.lexsort((e["draw"], e["id_idx"], e["id_gid"], e["id_call"]))
    eg, eo = eg[key(eg)], eo[key(eo)]
    for f in ("id_idx", "id_gid", "id_call", "draw", "kind", "block", "i", "j", "k"):
        assert np.array_equal(eg[f], eo[f]), f
    for f in ("t", "x", "y", "z", "v0", "v1", "v2"):
        assert pu.rel_err(eg[f], eo[f]) <= 1e-12, (f, pu.rel_err(eg[f], eo[f]))
        assert np.array_equal(eg[f], eo[f]), f   # in fact bit-exact


def test_stepdiff_ci_reflecting_1d():
    gpu, ora = pu.make_pair("stepdiff", ci=True, overrides=[("jaybenne", "num_particles", 2000)])
    run_pair(gpu, ora, cycles=2)


def test_stepdiff_shipped_two_blocks():
    gpu, ora = pu.make_pair("stepdiff", overrides=[("jaybenne", "num_particles", 2000)])
    sg, _ = run_pair(gpu, ora, cycles=2)
    assert sg["n_block_crossings"] > 0


def test_synthetic_3d_multiblock_emission_feedback():
    text = pu.synthetic_text(n=16, nb=8, num_particles=30000)
    gpu, ora = pu.make_pair(text=text)
    sg, _ = run_pair(gpu, ora, cycles=3, density_fn=pu.synthetic_density)
    assert sg["n_block_crossings"] > 0 and sg["n_absorbed"] > 0 and sg["n_sourced"] > 0


def test_synthetic_2d():
    text = pu.synthetic_text(n=16, nb=8, num_particles=5000, ndim=2)
    gpu, ora = pu.make_pair(text=text)
    run_pair(gpu, ora, cycles=2, density_fn=pu.synthetic_density)


def test_stepdiff_smr_two_levels():
    gpu, ora = pu.make_pair("stepdiff_smr", ci=True, overrides=[("jaybenne", "num_particles", 4000)])
    assert gpu.mesh.c.max_level == 1
    sg, _ = run_pair(gpu, ora, cycles=2)
    assert sg["n_block_crossings"] > 0


def test_outflow_escape():
    text = pu.synthetic_text(n=8, nb=4, num_particles=4000, kappa=0.0, sigma=1.0)
    text = text.replace("periodic", "outflow")
    gpu, ora = pu.make_pair(text=text)
    sg, _ = run_pair(gpu, ora, cycles=1)
    assert sg["n_escaped"] > 0
