"""GPU parity tests: the CUDA path through the C ABI against the CPU oracle on the same
seeded inputs. Integer results (counts, draw counters, cell indices) must be bit-exact;
photon state is compared bit for bit too (the kernels and the oracle share jb_math.h and
are both built without FMA contraction); cell tallies, which are sums whose order differs
(atomics), must agree to 1e-12 relative — the tolerance BASELINE.json's north_star names.
"""
import numpy as np
import pytest

import parity_util as pu
from jaybenne_b200 import capi

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
        # with material feedback the next cycle's source weights inherit the last bits of the
        # energy_delta sums, whose order differs (atomics): bit-exact only while nothing feeds back
        fed_back = c > 0 and gpu.cfg.do_feedback and sg["n_absorbed"] + sg["n_sourced"] > 0
        pu.assert_swarm_equal(gpu.jb, ora.jb, exact=exact and not fed_back)
        if fields:
            ncells = gpu.mesh.nblocks * int(np.prod(gpu.jb.nxb))
            pu.assert_fields_equal(gpu, ora, gross=(so["e_sourced"] + so["e_absorbed"]) / ncells)
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


# ---- DDMC / hybrid ---------------------------------------------------------------------------
def test_stepdiff_ddmc_ci():
    gpu, ora = pu.make_pair("stepdiff_ddmc", ci=True, overrides=[("jaybenne", "num_particles", 4000)])
    sg, _ = run_pair(gpu, ora, cycles=2, fields=False)
    assert sg["n_ddmc_segments"] == sg["n_segments"] > 0


def test_inf_stiff_ddmc_with_emission():
    gpu, ora = pu.make_pair("inf_stiff", overrides=[("jaybenne", "num_particles", 2000)])
    sg, _ = run_pair(gpu, ora, cycles=3)
    assert sg["n_ddmc_segments"] > 0 and sg["n_absorbed"] > 0 and sg["n_sourced"] > 0


def test_stepdiff_smr_ddmc_ci_block_face_resampling():
    gpu, ora = pu.make_pair("stepdiff_smr_ddmc", ci=True, overrides=[("jaybenne", "num_particles", 8000)])
    sg, _ = run_pair(gpu, ora, cycles=2, fields=False)
    assert sg["n_block_crossings"] > 0 and sg["n_ddmc_segments"] > 0
    for gid in gpu.jb.local_gids[:6]:
        for d in ("x1", "x2"):
            a, b = gpu.jb.get_field(gid, "ddmc_face_prob_" + d), ora.jb.get_field(gid, "ddmc_face_prob_" + d)
            assert np.array_equal(a, b), (gid, d)


def test_stepdiff_smr_hybrid_shipped_resolution_true_hybrid():
    gpu, ora = pu.make_pair("stepdiff_smr_hybrid", overrides=[("jaybenne", "num_particles", 8000)])
    sg, _ = run_pair(gpu, ora, cycles=1, fields=False)
    assert 0 < sg["n_ddmc_segments"] < sg["n_segments"]      # IMC in the fine region, DDMC outside


def test_synthetic_3d_ddmc_thick():
    text = pu.synthetic_text(n=8, nb=4, num_particles=6000, kappa=8.0, sigma=80.0,
                             extra="").replace("<jaybenne>", "<jaybenne>\nuse_ddmc = true\ntau_ddmc = 5.0")
    gpu, ora = pu.make_pair(text=text)
    sg, _ = run_pair(gpu, ora, cycles=2, density_fn=pu.synthetic_density)
    assert sg["n_ddmc_segments"] > 0


# ---- the committed golden fixtures (reference headers) against the CUDA path ------------------
def test_cuda_path_against_reference_golden_runs():
    import json, os
    import golden_cases as gc
    from jaybenne_b200 import device_api
    want = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "runs.json")))
    got = gc.run_table(device_api())
    for name, w in want.items():
        if name.startswith("_"):
            continue
        g = got[name]
        assert g["nblocks"] == w["nblocks"] and g["n_init"] == w["n_init"], name
        for cyc, (gr, wr) in enumerate(zip(g["cycles"], w["cycles"])):
            for k in gc.COUNT_KEYS:
                assert gr[k] == wr[k], (name, cyc, k, gr[k], wr[k])
            for k in gc.ENERGY_KEYS + ("tally_sum",):
                a, b = gc.unhex(gr[k]), gc.unhex(wr[k])
                assert abs(a - b) <= 1e-12 * max(abs(a), abs(b), 1e-300), (name, cyc, k)


# ---- other closures -----------------------------------------------------------------------------
def test_epbremss_thomson_closures():
    text = pu.synthetic_text(n=8, nb=4, num_particles=3000)
    text = text.replace("opacity_model = constant", "opacity_model = ep_bremss").replace(
        "initial_temperature = 1.0", "initial_temperature = 1.0e8").replace("initial_density = 1.0", "initial_density = 1.0e-12")
    # free-free absorption ~ 26 / cm at the Planck-sampled "frequencies", Thomson scattering ~ 10 / cm
    gpu, ora = pu.make_pair(text=text, cfg_overrides={"scattering_model": 1, "apm": 3.3e-38})
    run_pair(gpu, ora, cycles=2)


# ---- reference acceptance + size-independent properties at scale --------------------------------
def test_stepdiff_analytic_acceptance_full():
    """tst/stepdiff.py as the CI runs it: 128 cells, 1e5 photons, 10 cycles, tolerance 0.05."""
    import regression_util as ru
    from jaybenne_b200 import Mcblock, device_api
    mb = Mcblock(pu.make_pin("stepdiff", ci=True), api=device_api())
    mb.ProblemGenerator()
    mb.Execute()
    assert mb.ncycle == 10
    err = ru.weighted_mean_frac_error(mb.dump())
    assert err <= ru.TOL_STEPDIFF, err


def test_stepdiff_smr_analytic_acceptance_full():
    """tst/stepdiff_smr.py: 64x32 base, 16x16 blocks, tolerance 0.3 — IMC, DDMC and hybrid decks."""
    import regression_util as ru
    from jaybenne_b200 import Mcblock, device_api
    for deck in ("stepdiff_smr", "stepdiff_smr_ddmc", "stepdiff_smr_hybrid"):
        mb = Mcblock(pu.make_pin(deck, ci=True), api=device_api())
        mb.ProblemGenerator()
        mb.Execute()
        err = ru.weighted_mean_frac_error(mb.dump())
        assert err <= ru.TOL_STEPDIFF_SMR, (deck, err)


def test_large_run_conservation_and_count_identities():
    """128^3 cells / 1.25e8 photons (the single-GPU twin of the BASELINE workload): properties
    that do not need the oracle — energy conservation to 1e-12, photon bookkeeping, tally sum."""
    from jaybenne_b200 import Mcblock, device_api
    n = 128
    text = pu.synthetic_text(n=n, nb=32, num_particles=int(1.25e8))
    mb = Mcblock(pu.make_pin(text=text), api=device_api(), capacity=int(1.3e8), cfg_overrides={"source_nblocks": 64})
    mb.ProblemGenerator(pu.synthetic_density)
    for _ in range(2):
        st = mb.Step()
        mb.time += mb.dt
        assert pu.conservation_residual(st) <= 1e-12, pu.conservation_residual(st)
        assert st["n_end"] == st["n_start"] + st["n_sourced"] - st["n_absorbed"] - st["n_escaped"]
        assert st["n_iterations"] == 1 and st["n_segments"] > 5 * st["n_start"]
        dv = (1.0 / n) ** 3
        tally = sum(mb.jb.get_field(g, "energy_tally").sum() for g in mb.jb.local_gids) * dv
        assert abs(tally - st["e_end"]) <= 1e-11 * st["e_end"]
        edelta = sum(mb.jb.get_field(g, "energy_delta").sum() for g in mb.jb.local_gids)
        assert abs(edelta - (st["e_absorbed"] - st["e_sourced"])) <= 1e-9 * st["e_absorbed"]


def test_mid_size_run_is_bit_identical_to_the_oracle():
    """32^3 cells in 16^3 blocks, 1e6 photons, two cycles with emission and feedback: the largest case the serial
    oracle finishes in seconds (8e6 segments per cycle). Thousands of chunks per warp-private pool, every phase of
    k_transport_pool at full width, block crossings in all 26 directions — counts, census swarm (positions,
    directions, times, draw counters, cells) bit for bit, tallies to 1e-12."""
    text = pu.synthetic_text(n=32, nb=16, num_particles=1000000)
    gpu, ora = pu.make_pair(text=text)
    gpu.ProblemGenerator(pu.synthetic_density); ora.ProblemGenerator(pu.synthetic_density)
    pu.assert_swarm_equal(gpu.jb, ora.jb)
    for c in range(2):
        sg, so = gpu.Step(), ora.Step()
        gpu.time += gpu.dt; ora.time += ora.dt
        for k in pu.COUNT_KEYS:
            assert sg[k] == so[k], (k, sg[k], so[k])
        # energy sums over 1e6 photons: the oracle adds them serially (rounding ~ sqrt(n) ulp), the device pairwise
        for k in pu.ENERGY_KEYS:
            assert pu.rel_err([sg[k]], [so[k]]) <= 1e-10, (k, sg[k], so[k])
        assert pu.conservation_residual(sg) <= 1e-12
        # cycle 2's source weights inherit the last bits of cycle 1's energy_delta sums (atomic order): bit-exact
        # swarm after the first cycle, 1e-12 after the second
        pu.assert_swarm_equal(gpu.jb, ora.jb, exact=(c == 0))
        ncells = gpu.mesh.nblocks * int(np.prod(gpu.jb.nxb))
        pu.assert_fields_equal(gpu, ora, gross=(so["e_sourced"] + so["e_absorbed"]) / ncells)
    assert sg["n_segments"] > 4000000 and sg["n_block_crossings"] > 100000


# ---- north_star check 3: seed ensembles against the reference-faithful generator ------------------
def _ensemble(api, rng_mode, deck, overrides, cycles, seeds, ci=False):
    from jaybenne_b200 import Mcblock
    tallies, sies = [], []
    for seed in seeds:
        pin = pu.make_pin(deck, overrides=list(overrides) + [("jaybenne", "seed", seed)], ci=ci)
        mb = Mcblock(pin, api=api, cfg_overrides={"rng_mode": rng_mode})
        mb.ProblemGenerator()
        for _ in range(cycles):
            mb.Step()
            mb.time += mb.dt
        d = mb.dump()
        tallies.append(np.concatenate([t.ravel() for t in d["energy_tally"]]))
        sies.append(np.concatenate([t.ravel() for t in d["sie"]]))
    return np.array(tallies), np.array(sies)


def _assert_within_3_sigma(a, b, label):
    """a, b: [seed][bin]. Means agree within 3 standard errors; with tens of bins one excursion
    up to 4.5 sigma is allowed (p(3 sigma) = 0.27 % per bin)."""
    se = np.sqrt(a.var(axis=0, ddof=1) / a.shape[0] + b.var(axis=0, ddof=1) / b.shape[0])
    se = np.maximum(se, 1e-12 * np.maximum(np.abs(a.mean(axis=0)), np.abs(b.mean(axis=0))))
    live = se > 0
    z = np.abs(a.mean(axis=0) - b.mean(axis=0))[live] / se[live]
    assert (z > 3.0).sum() <= 1 and z.max() <= 4.5, (label, np.sort(z)[-3:])


def test_seed_ensemble_profiles_match_reference_generator_within_3_sigma():
    """stepdiff (radiation profile of the Marshak wave) and inf (equilibration with feedback):
    CUDA path with per-photon Philox streams vs the CPU oracle running the reference's
    XorShift64 pool in serial order, 16 seeds each; T_rad ~ tally^(1/4) and T_mat ~ sie profiles
    must agree within 3 sigma of the Monte Carlo noise."""
    from jaybenne_b200 import capi, device_api
    from oracle_binding import oracle_api
    seeds = [349857 + k for k in range(16)]
    ov = [("jaybenne", "num_particles", 4000)]
    g_t, _ = _ensemble(device_api(), capi.JB_RNG_PHILOX, "stepdiff", ov, 3, seeds, ci=True)
    o_t, _ = _ensemble(oracle_api("std"), capi.JB_RNG_XORSHIFT_POOL, "stepdiff", ov, 3, seeds, ci=True)
    nb = 16                                     # 128 cells -> 16 bins of 8 cells
    g_b = g_t.reshape(len(seeds), nb, -1).sum(axis=2)
    o_b = o_t.reshape(len(seeds), nb, -1).sum(axis=2)
    _assert_within_3_sigma(g_b, o_b, "stepdiff energy_tally")
    ov = [("jaybenne", "num_particles", 200), ("jaybenne", "do_feedback", "true")]
    g_t, g_s = _ensemble(device_api(), capi.JB_RNG_PHILOX, "inf", ov, 3, seeds)
    o_t, o_s = _ensemble(oracle_api("std"), capi.JB_RNG_XORSHIFT_POOL, "inf", ov, 3, seeds)
    _assert_within_3_sigma(g_t.sum(axis=1, keepdims=True), o_t.sum(axis=1, keepdims=True), "inf radiation energy")
    _assert_within_3_sigma(g_s.mean(axis=1, keepdims=True), o_s.mean(axis=1, keepdims=True), "inf material sie")


def test_gray_direct_mode_matches_oracle_and_stays_within_event_tolerance():
    """JB_GRAY_DIRECT (sigma_a = rho * kappa instead of the reference's j_nu / B_nu round trip,
    gray_opacity_photons.hpp:50-54): bit-exact against the oracle in the same mode, and against the
    reference arithmetic the counts are identical and the swarm agrees to the 1e-12 event tolerance."""
    text = pu.synthetic_text(n=16, nb=8, num_particles=20000)
    direct = {"gray_mode": capi.JB_GRAY_DIRECT}
    gpu, ora = pu.make_pair(text=text, cfg_overrides=direct)
    _, ora_ref = pu.make_pair(text=text)
    for mb in (gpu, ora, ora_ref):
        mb.ProblemGenerator(pu.synthetic_density)
    sg, so, sr = gpu.Step(), ora.Step(), ora_ref.Step()
    pu.assert_stats_equal(sg, so)
    pu.assert_swarm_equal(gpu.jb, ora.jb)
    pu.assert_stats_equal(sg, sr)
    pu.assert_swarm_equal(gpu.jb, ora_ref.jb, tol=1e-12, exact=False)


def test_pool_holes_are_squeezed_by_occupancy_not_every_cycle():
    """Absorbed photons leave holes; the pool is compacted only when occupancy drops below 1/2 (or
    <jaybenne> min_swarm_occupancy), on swarm access, or for room. Every cycle must still match the
    oracle, whichever state the pool is in."""
    text = pu.synthetic_text(n=16, nb=8, num_particles=40000)
    gpu, ora = pu.make_pair(text=text)
    gpu.ProblemGenerator(pu.synthetic_density); ora.ProblemGenerator(pu.synthetic_density)
    launches = []
    for _ in range(4):                       # no swarm access in between: the pool keeps its holes
        l0 = gpu.jb.kernel_launches()
        sg, so = gpu.Step(), ora.Step()
        launches.append(gpu.jb.kernel_launches() - l0)
        pu.assert_stats_equal(sg, so)
        assert pu.conservation_residual(sg) <= 1e-12
    # ~1/3 of the photons are absorbed per cycle: cycle 1 stays above 1/2 occupancy (no compaction), a later one
    # crosses it (two more launches: find + move)
    assert max(launches) >= min(launches) + 2, launches
    pu.assert_swarm_equal(gpu.jb, ora.jb, exact=False)
    pu.assert_fields_equal(gpu, ora, gross=None, names=("energy_tally", "fleck_factor", "internal_energy", "sie"))
    # a pool capped at exactly what a dense pool needs still sources: the holes are squeezed out for room
    n_now = gpu.jb.num_particles()
    assert n_now == sg["n_end"]


def test_batched_field_access_equals_per_block_access():
    """jb_set_fields / jb_get_fields (every block of a pack in one call) against jb_set_field / jb_get_field."""
    text = pu.synthetic_text(n=16, nb=8, num_particles=4000)
    gpu, _ = pu.make_pair(text=text)
    gpu.ProblemGenerator(pu.synthetic_density)
    gids = gpu.jb.local_gids
    n = gpu.jb.field_size("density")
    rng = np.random.default_rng(3)
    rho = [1.0 + rng.random(n) for _ in gids]
    gpu.jb.set_fields_ptr("density", gids, [a.ctypes.data for a in rho])
    for g, a in zip(gids, rho):
        assert np.array_equal(gpu.jb.get_field(g, "density").ravel(), a)
    out = [np.empty(n) for _ in gids]
    gpu.jb.get_fields_ptr("energy_tally", gids, [a.ctypes.data for a in out])
    for g, a in zip(gids, out):
        assert np.array_equal(gpu.jb.get_field(g, "energy_tally").ravel(), a)
    with pytest.raises(Exception):
        gpu.jb.set_fields_ptr("density", [10 ** 6], [rho[0].ctypes.data])


def test_cell_sorted_pool_gives_identical_results():
    """sort_interval = 1: after every cycle the pool is compacted and sorted by (meshblock, cell) (counting sort +
    gather). Streams travel with the photons, so nothing may change except the slot order."""
    text = pu.synthetic_text(n=16, nb=8, num_particles=30000)
    gpu, ora = pu.make_pair(text=text, cfg_overrides={"sort_interval": 1})
    gpu.ProblemGenerator(pu.synthetic_density); ora.ProblemGenerator(pu.synthetic_density)
    for _ in range(3):
        sg, so = gpu.Step(), ora.Step()
        pu.assert_stats_equal(sg, so)
    pu.assert_swarm_equal(gpu.jb, ora.jb, exact=False)
    p = gpu.jb.particles()
    ng = gpu.mesh.c.nghost
    n1, n2, n3 = gpu.jb.nxb
    cell = ((p["k"] - ng) * n2 + (p["j"] - ng)) * n1 + (p["i"] - ng)
    key = p["block"].astype(np.int64) * (n1 * n2 * n3) + cell
    assert len(key) == sg["n_end"] and np.all(np.diff(key) >= 0), "pool is not sorted by (block, cell)"


def test_pool_stays_ordered_through_transport_and_squeeze():
    """k_transport_pool packs a chunk's survivors at the front of that chunk and the squeeze (k_squeeze_*) keeps the
    relative order of the live photons: a photon never overtakes one that was sourced >= 2 chunks (2048 slots)
    ahead of it, cycle after cycle — the (meshblock, cell) order of the sourcing pass only blurs locally. With
    do_emission off nothing is appended, so the thermal photons' birth slot is their rank in (block, birth index)."""
    text = pu.synthetic_text(n=16, nb=8, num_particles=60000, do_emission=False)
    gpu, ora = pu.make_pair(text=text)
    gpu.ProblemGenerator(pu.synthetic_density); ora.ProblemGenerator(pu.synthetic_density)
    p0 = gpu.jb.particles()
    rid0 = p0["rng"].astype(np.int64)
    # birth slot of (birth index, birth gid): the position straight after sourcing
    birth = {(int(a), int(b)): q for q, (a, b) in enumerate(zip(rid0[:, 0], rid0[:, 1]))}
    n0 = len(rid0)
    for cyc in range(3):
        sg, so = gpu.Step(), ora.Step()
        for k in pu.COUNT_KEYS:   # (energy sums of 60000 equal weights: the oracle's serial sum is the noisier one)
            assert sg[k] == so[k], (k, sg[k], so[k])
        p = gpu.jb.particles()   # squeezes (order-preserving) and reads the pool in slot order
        rid = p["rng"].astype(np.int64)
        assert len(rid) == sg["n_end"] and sg["n_end"] < n0
        slot = np.array([birth[(int(a), int(b))] for a, b in zip(rid[:, 0], rid[:, 1])])
        chunk = slot // 1024
        # non-decreasing birth chunk along the pool: a chunk's photons stay together and chunks stay in order
        assert np.all(np.diff(chunk) >= 0), f"cycle {cyc}: the pool lost its order"
    pu.assert_swarm_equal(gpu.jb, ora.jb)


def test_step_timeline_reports_the_phases_of_the_last_step():
    """jb_step_timeline: device-timed phases of the last RadiationStep in stream order; one transport phase on a
    single rank, every duration >= 0, and together no longer than the step's own event bracket."""
    gpu, _ = pu.make_pair(text=pu.synthetic_text(n=16, nb=8, num_particles=20000))
    gpu.ProblemGenerator(pu.synthetic_density)
    st = gpu.Step()
    tl = gpu.jb.step_timeline()
    names = [n for n, _ in tl]
    assert names.count("transport") == 1 and "exchange" not in names
    assert names[0] == "fields" and names[-1] == "end" and "tally" in names and "source" in names
    assert all(ms >= 0.0 for _, ms in tl)
    assert sum(ms for _, ms in tl) <= st["ms_total"] * 1.001 + 0.01
    assert abs(dict(tl)["transport"] - st["ms_transport"]) <= 0.05 + 0.05 * st["ms_transport"]
    assert gpu.jb.exchange_mode() == 0


def test_range_test_free_division_is_the_ieee_quotient():
    """jbm::div_normal (nvcc's fast-path sequence without the operand test and slow-path call) must return the
    correctly rounded quotient, i.e. the bits of `/`, over the operand ranges the kernels use it on."""
    import ctypes as C
    gpu, _ = pu.make_pair("inf")
    bad = C.c_int64(-1)
    assert gpu.api.selftest_division(gpu.jb._ctx, 200_000_000, C.byref(bad)) == 0
    assert bad.value == 0, bad.value


def test_hybrid_deck_event_by_event():
    """The IMC / DDMC hybrid kernel's event log against the oracle's on the shipped stepdiff_smr_hybrid deck (fine
    cells IMC, coarse cells DDMC, two refinement levels): every segment of every photon — kind bits (DDMC, albedo
    reject, left block, absorb, census), cell, block and state — in stream order."""
    gpu, ora = pu.make_pair("stepdiff_smr_hybrid", overrides=[("jaybenne", "num_particles", 600)])
    gpu.ProblemGenerator(); ora.ProblemGenerator()
    cap = 2000000
    gpu.jb.event_log_enable(cap); ora.jb.event_log_enable(cap)
    sg, so = gpu.Step(), ora.Step()
    pu.assert_stats_equal(sg, so)
    eg, eo = gpu.jb.events(), ora.jb.events()
    assert len(eg) == len(eo) and 0 < len(eg) < cap
    assert sg["n_ddmc_segments"] > 0 and sg["n_segments"] > sg["n_ddmc_segments"]  # a true hybrid run
    key = lambda e: np.lexsort((e["draw"], e["id_idx"], e["id_gid"], e["id_call"]))
    eg, eo = eg[key(eg)], eo[key(eo)]
    for f in ("id_idx", "id_gid", "id_call", "draw", "kind", "block", "i", "j", "k"):
        assert np.array_equal(eg[f], eo[f]), f
    for f in ("t", "x", "y", "z", "v0", "v1", "v2"):
        assert np.array_equal(eg[f], eo[f]), f


# ---- kernel variants and the VALIDATE cross-checks -------------------------------------------------
VARIANT_DECKS = {
    "synthetic3d": dict(text=pu.synthetic_text(n=16, nb=8, num_particles=30000), density=True, cycles=3),
    "synthetic2d": dict(text=pu.synthetic_text(n=16, nb=8, num_particles=6000, ndim=2), density=True, cycles=2),
    "stepdiff_reflecting": dict(deck="stepdiff", overrides=[("jaybenne", "num_particles", 3000)], cycles=2),
    "stepdiff_smr_levels": dict(deck="stepdiff_smr", ci=True, overrides=[("jaybenne", "num_particles", 4000)], cycles=2),
    "outflow": dict(text=pu.synthetic_text(n=8, nb=4, num_particles=4000, kappa=0.0, sigma=1.0).replace("periodic", "outflow"),
                    cycles=1),
}


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("name", sorted(VARIANT_DECKS))
def test_imc_kernel_variants_are_bit_identical_to_the_oracle(monkeypatch, variant, name):
    """JB_TRANSPORT_VARIANT: 0 = thread per photon, 1 = persistent warps with lane refill (v9), 2 = warp-private
    photon pools in shared memory (v10, the default). Streams are per photon, so every variant must give the
    oracle's swarm bit for bit."""
    monkeypatch.setenv("JB_TRANSPORT_VARIANT", str(variant))
    d = VARIANT_DECKS[name]
    gpu, ora = pu.make_pair(d.get("deck"), text=d.get("text"), overrides=d.get("overrides", ()), ci=d.get("ci", False))
    sg, _ = run_pair(gpu, ora, cycles=d["cycles"], density_fn=pu.synthetic_density if d.get("density") else None)
    assert sg["n_segments"] > 0


@pytest.mark.parametrize("slots,refill,low", [(56, 1, 1), (56, 32, 32), (64, 16, 24), (64, 8, 32)])
def test_pool_kernel_tuning_does_not_change_results(monkeypatch, slots, refill, low):
    """The pool kernel's phase thresholds only change the order in which a warp visits its photons."""
    monkeypatch.setenv("JB_TRANSPORT_VARIANT", "2")
    monkeypatch.setenv("JB_POOL_SLOTS", str(slots))
    monkeypatch.setenv("JB_POOL_REFILL_MIN", str(refill))
    monkeypatch.setenv("JB_POOL_READY_LOW", str(low))
    gpu, ora = pu.make_pair(text=pu.synthetic_text(n=16, nb=8, num_particles=30000))
    run_pair(gpu, ora, cycles=3, density_fn=pu.synthetic_density)


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("name", sorted(VARIANT_DECKS))
def test_validate_build_finds_no_disagreement(monkeypatch, variant, name):
    """JB_IMC_VALIDATE=1 runs the instantiation that re-derives every cell index / block with Xtoijk / find_block
    beside the index-increment shortcut and evaluates the IEEE `/` beside every div_normal; a disagreement makes
    the step fail (DStats::n_validate)."""
    monkeypatch.setenv("JB_TRANSPORT_VARIANT", str(variant))
    monkeypatch.setenv("JB_IMC_VALIDATE", "1")
    d = VARIANT_DECKS[name]
    gpu, ora = pu.make_pair(d.get("deck"), text=d.get("text"), overrides=d.get("overrides", ()), ci=d.get("ci", False))
    run_pair(gpu, ora, cycles=d["cycles"], density_fn=pu.synthetic_density if d.get("density") else None)


def test_validate_build_at_128_cubed(monkeypatch):
    """The same cross-checks at 128^3 cells / 1.25e8 photons (1e9 segments, 6e9 divisions): the step must not
    report a single disagreement, and must give the counts of the production instantiation."""
    from jaybenne_b200 import Mcblock, device_api
    text = pu.synthetic_text(n=128, nb=32, num_particles=int(1.25e8))
    out = []
    for validate in ("1", "0"):
        monkeypatch.setenv("JB_IMC_VALIDATE", validate)
        mb = Mcblock(pu.make_pin(text=text), api=device_api(), capacity=int(1.3e8), cfg_overrides={"source_nblocks": 64})
        mb.ProblemGenerator(pu.synthetic_density)
        st = mb.Step()           # raises on n_validate != 0
        out.append({k: st[k] for k in pu.COUNT_KEYS})
        del mb
    assert out[0] == out[1], out
    assert out[0]["n_segments"] > 9e8


# ---- three-dimensional static refinement (decks.EXTRA_DECKS; goldens: runs.json smr3d_*) ---------------
def test_smr3d_imc_level_changes_in_three_dimensions():
    """Pure IMC on 56 coarse + 64 fine blocks: photons cross faces, edges and corners between levels in 3D
    (find_block + Xtoijk on the finer / coarser side, transport.cpp:146-155)."""
    gpu, ora = pu.make_pair("smr3d_imc")
    assert gpu.mesh.c.max_level == 1 and gpu.mesh.nblocks == 120 and gpu.mesh.c.ndim == 3
    sg, _ = run_pair(gpu, ora, cycles=2)
    assert sg["n_block_crossings"] > 1000


def test_smr3d_ddmc_block_face_resampling_in_three_dimensions():
    """Every cell DDMC (tau_ddmc = 2.5): leaks out of coarse blocks into fine ones arrive on a fine-cell VERTEX
    and go through SampleDDMCBlockFace::3D / SampleFace3D (sample_ddmc_bface.cpp:41-78,241-423)."""
    gpu, ora = pu.make_pair("smr3d_ddmc")
    sg, so = run_pair(gpu, ora, cycles=2, fields=False)
    assert sg["n_ddmc_segments"] == sg["n_segments"] > 0
    assert sg["n_bface_resampled"] == so["n_bface_resampled"] > 100      # SampleFace3D draws happened
    for gid in gpu.jb.local_gids[::7]:
        for d in ("x1", "x2", "x3"):
            a, b = gpu.jb.get_field(gid, "ddmc_face_prob_" + d), ora.jb.get_field(gid, "ddmc_face_prob_" + d)
            assert np.array_equal(a, b), (gid, d)


def test_smr3d_hybrid_true_hybrid_in_three_dimensions():
    """tau_ddmc = 20 between the optical depths of the fine (15.6) and the coarse (31) cells: IMC inside the
    refined region, DDMC outside, albedo + block-face resampling at the 3D level boundary."""
    gpu, ora = pu.make_pair("smr3d_hybrid", overrides=[("jaybenne", "num_particles", 12000), ("jaybenne", "dt", "5.0e-12")])
    sg, _ = run_pair(gpu, ora, cycles=2, fields=False)
    assert 0 < sg["n_ddmc_segments"] < sg["n_segments"] and sg["n_bface_resampled"] > 0


def test_smr3d_hybrid_event_by_event():
    gpu, ora = pu.make_pair("smr3d_hybrid", overrides=[("jaybenne", "num_particles", 12000), ("jaybenne", "dt", "1.0e-12")])
    gpu.ProblemGenerator(); ora.ProblemGenerator()
    cap = 3000000
    gpu.jb.event_log_enable(cap); ora.jb.event_log_enable(cap)
    sg, so = gpu.Step(), ora.Step()
    pu.assert_stats_equal(sg, so)
    eg, eo = gpu.jb.events(), ora.jb.events()
    assert len(eg) == len(eo) and 0 < len(eg) < cap
    assert sg["n_ddmc_segments"] > 0 and sg["n_segments"] > sg["n_ddmc_segments"] and sg["n_block_crossings"] > 0
    key = lambda e: np.lexsort((e["draw"], e["id_idx"], e["id_gid"], e["id_call"]))
    eg, eo = eg[key(eg)], eo[key(eo)]
    for f in ("id_idx", "id_gid", "id_call", "draw", "kind", "block", "i", "j", "k"):
        assert np.array_equal(eg[f], eo[f]), f
    for f in ("t", "x", "y", "z", "v0", "v1", "v2"):
        assert np.array_equal(eg[f], eo[f]), f
