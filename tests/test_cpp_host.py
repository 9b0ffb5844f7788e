"""The C++ host side above the C ABI: include/jaybenne_b200.hpp + the `mcblock_b200` driver
(jaybenne_b200/csrc/host/mcblock_main.cpp), i.e. the reference's mcblock application
(src/mcblock/main.cpp, mcblock.cpp, mcblock_driver.cpp) re-hosted on libjaybenne_b200.so.

CPU: the binary exists, parses every shipped deck like the Python mirror, reproduces the
reference's input errors and refuses to compute without a GPU. GPU: a run of the binary is
bit-identical (swarm) to the CPU oracle driven through the Python mirror, the host-sequenced
task list equals the fused step, and the re-hosted tst/stepdiff.py acceptance passes on its dump.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import parity_util as pu
import regression_util as ru
from jaybenne_b200 import Mcblock, ParameterInput, capi
from jaybenne_b200.decks import CI_OVERRIDES, DECKS, deck_text
from oracle_binding import oracle_api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "jaybenne_b200", "mcblock_b200")


def run_bin(args, env=None, check=True, timeout=600):
    e = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        e.pop(k, None)
    e.update(env or {})
    r = subprocess.run([BIN] + [str(a) for a in args], capture_output=True, text=True, env=e, timeout=timeout)
    if check:
        assert r.returncode == 0, r.stderr + r.stdout
    return r


def write_deck(tmp_path, name):
    path = tmp_path / f"{name}.in"
    path.write_text(deck_text(name))
    return path


def ci_args(name):
    return [f"{blk}/{key}={val}" for blk, key, val in CI_OVERRIDES.get(name, [])]


def load_dump(outdir, problem_id, tag="final", particles=False):
    """The .npy dump directory of every rank as one dict shaped like Mcblock.dump()."""
    d = os.path.join(outdir, f"{problem_id}.out0.{tag}")
    ranks = sorted({f.split(".")[0] for f in os.listdir(d)})
    out = {}
    names = ["density", "sie", "internal_energy", "energy_tally", "xmin", "xmax", "level", "gids"]
    if particles:
        names += ["photons." + k for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2", "i", "j", "k",
                                            "block", "rng")]
    for n in names:
        out[n] = np.concatenate([np.load(os.path.join(d, f"{r}.{n}.npy")) for r in ranks])
    infos = [json.load(open(os.path.join(d, f"{r}.info.json"))) for r in ranks]
    out["time"], out["ncycle"] = infos[0]["time"], infos[0]["ncycle"]
    out["info"] = infos
    return out


def read_history(outdir, problem_id, rank=0):
    import csv
    with open(os.path.join(outdir, f"{problem_id}.hst.r{rank}.csv")) as f:
        return list(csv.DictReader(f))


# ------------------------------------------------------------------------------------------
# CPU
# ------------------------------------------------------------------------------------------
def test_binary_is_built_and_links_the_cuda_library():
    assert os.access(BIN, os.X_OK), "make -C jaybenne_b200/csrc builds mcblock_b200"
    out = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libjaybenne_b200.so" in out and "not found" not in out, out


@pytest.mark.parametrize("name", sorted(DECKS))
def test_parse_only_agrees_with_the_python_mirror(tmp_path, name):
    path = write_deck(tmp_path, name)
    r = run_bin(["-i", path, "--parse-only"] + ci_args(name))
    kv = {}
    for tok in r.stdout.split():
        k, _, v = tok.partition("=")
        kv[k] = v
    pin = pu.make_pin(name, ci=True)
    cfg, mesh = pin.config(), pin.mesh()
    assert int(kv["nbtotal"]) == mesh.nblocks and int(kv["ndim"]) == mesh.c.ndim
    assert kv["block"] == "x".join(str(v) for v in mesh.c.nxb)
    assert int(kv["num_particles"]) == cfg.num_particles and int(kv["seed"]) == cfg.seed
    assert float(kv["dt"]) == cfg.dt and float(kv["tau_ddmc"]) == cfg.tau_ddmc
    assert int(kv["use_ddmc"]) == cfg.use_ddmc and int(kv["do_emission"]) == cfg.do_emission
    assert float(kv["kappa"]) == cfg.opacity_kappa and float(kv["sigma"]) == cfg.scattering_sigma
    assert float(kv["cv"]) == cfg.eos_cv
    assert kv["swarm_bc"] == ",".join(str(v) for v in mesh.c.swarm_bc)


def test_input_errors_read_like_the_reference(tmp_path):
    path = write_deck(tmp_path, "inf")
    # McblockDriver requires all six mesh BC keys (mcblock_driver.cpp:23-28) — checked before any GPU work
    text = "\n".join(l for l in deck_text("inf").splitlines() if not l.startswith("ox3_bc"))
    bad = tmp_path / "bad.in"
    bad.write_text(text)
    r = run_bin(["-i", bad, "--parse-only"], check=False)
    assert r.returncode == 1 and "Parameter name 'ox3_bc' not found in block 'parthenon/mesh'" in r.stderr
    r = run_bin(["-i", tmp_path / "missing.in", "--parse-only"], check=False)
    assert r.returncode == 1 and "cannot read input deck" in r.stderr
    r = run_bin(["-i", path, "--parse-only", "jaybenne/min_swarm_occupancy=1.5"], check=False)
    assert r.returncode == 1 and "FATAL" in r.stderr
    r = run_bin(["--parse-only"], check=False)
    assert r.returncode == 2 and "usage" in r.stderr


def test_refuses_to_compute_without_a_device(tmp_path):
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    r = run_bin(["-i", write_deck(tmp_path, "inf"), "--outdir", tmp_path / "out"], check=False)
    assert r.returncode == 1 and "no CUDA device" in r.stderr


# ------------------------------------------------------------------------------------------
# GPU
# ------------------------------------------------------------------------------------------
def oracle_run(name, cycles, ci=False):
    ora = Mcblock(pu.make_pin(name, ci=ci), api=oracle_api("portable"))
    ora.ProblemGenerator()
    ora.Execute(max_cycles=cycles)
    return ora


def assert_photons_match(dump, ora_jb):
    o = pu.sorted_particles(ora_jb)
    rid = dump["photons.rng"]
    order = np.lexsort((rid[:, 0], rid[:, 1], rid[:, 2]))
    assert len(rid) == len(o["x"])
    assert np.array_equal(rid[order], o["rng"])
    for k in ("x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2", "i", "j", "k", "block"):
        assert np.array_equal(dump["photons." + k][order], o[k]), k


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fused", "tasklist"])
@pytest.mark.parametrize("name,cycles,ci", [("inf", 3, False), ("stepdiff_smr_hybrid", 2, True)])
def test_cpp_driver_is_bit_identical_to_the_oracle(tmp_path, name, cycles, ci, mode):
    out = tmp_path / "out"
    args = ["-i", write_deck(tmp_path, name), "--outdir", out, "--max-cycles", cycles, "--dump-particles", "--quiet"]
    args += ci_args(name) if ci else []
    if mode == "tasklist":
        args.append("--tasklist")
    run_bin(args)
    pid = DECKS[name]["parthenon/job"]["problem_id"]
    dump = load_dump(out, pid, particles=True)
    assert dump["ncycle"] == cycles and dump["info"][0]["kernel_launches"] > 0
    ora = oracle_run(name, cycles, ci)
    assert_photons_match(dump, ora.jb)
    od = ora.dump()
    assert np.array_equal(dump["gids"], od["gids"]) and np.array_equal(dump["xmin"], od["xmin"])
    for f in ("density", "sie", "internal_energy"):
        assert pu.rel_err(dump[f], od[f]) <= 1e-12, f
    assert pu.rel_err(dump["energy_tally"], od["energy_tally"]) <= 1e-12
    hist = read_history(out, pid)
    assert len(hist) == cycles
    for row, st in zip(hist, ora.history):
        for k in pu.COUNT_KEYS:
            assert int(row[k]) == st[k], k
        for k in pu.ENERGY_KEYS:
            assert abs(float(row[k]) - st[k]) <= 1e-12 * max(abs(st[k]), 1e-300), k


@pytest.mark.gpu
def test_cpp_driver_passes_the_stepdiff_acceptance(tmp_path):
    """tst/stepdiff.py through the C++ host: CI overrides, run to tlim, weighted-mean error <= 0.05."""
    out = tmp_path / "out"
    run_bin(["-i", write_deck(tmp_path, "stepdiff"), "--outdir", out, "--quiet"] + ci_args("stepdiff"))
    dump = load_dump(out, "stepdiff")
    assert dump["ncycle"] == 10
    assert abs(dump["time"] - 3.335641e-10) <= 1e-20
    assert ru.weighted_mean_frac_error(dump) <= ru.TOL_STEPDIFF
    # Parthenon-style numbered outputs: t = 0 and t = <parthenon/output0> dt
    assert os.path.isdir(out / "stepdiff.out0.00000") and os.path.isdir(out / "stepdiff.out0.00001")


@pytest.mark.gpu
def test_cpp_driver_two_ranks_over_nccl(tmp_path):
    """One process per GPU (RANK / WORLD_SIZE / LOCAL_RANK), NCCL id through a file: the union of
    the ranks' swarms equals the one-rank swarm bit for bit."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    name, cycles = "stepdiff_smr", 2
    deck = write_deck(tmp_path, name)
    common = ["-i", deck, "--max-cycles", cycles, "--dump-particles", "--quiet"] + ci_args(name)
    run_bin(common + ["--outdir", tmp_path / "one"])
    procs = []
    for r in range(2):
        e = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_PORT="29517")
        procs.append(subprocess.Popen([BIN] + [str(a) for a in common + ["--outdir", tmp_path / "two"]], env=e,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    for p in procs:
        so, se = p.communicate(timeout=600)
        assert p.returncode == 0, se + so
    one = load_dump(tmp_path / "one", "stepdiff", particles=True)
    two = load_dump(tmp_path / "two", "stepdiff", particles=True)
    assert len(two["info"]) == 2 and len(one["photons.x"]) == len(two["photons.x"])

    def key(d):
        rid = d["photons.rng"]
        return np.lexsort((rid[:, 0], rid[:, 1], rid[:, 2]))
    a, b = key(one), key(two)
    for k in ("rng", "x", "y", "z", "time", "weight", "energy", "v0", "v1", "v2", "i", "j", "k", "block"):
        assert np.array_equal(one["photons." + k][a], two["photons." + k][b]), k
    g1, g2 = np.argsort(one["gids"]), np.argsort(two["gids"])
    assert pu.rel_err(one["energy_tally"][g1], two["energy_tally"][g2]) <= 1e-12
