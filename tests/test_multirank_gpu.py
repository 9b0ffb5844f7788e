"""The N > 1 path of the CUDA library.

1. two rank contexts on ONE GPU, swarm exchange staged through the host (jb_exchange_take /
   jb_exchange_give): exercises the device outbox written by the transport kernel, the unpack
   kernel and the iteration-until-complete loop, and runs wherever one GPU is available;
2. two processes on TWO GPUs, exchange over NCCL inside jb_radiation_step (skipped with < 2 GPUs).

Random streams are keyed per photon, so in both cases the union of the ranks' swarms must equal
the single-rank swarm bit for bit and every block's fields must agree to the atomic-order
tolerance (1e-12) — a property the reference does not have.
"""
import os
import pickle
import socket
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import parity_util as pu

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
FIELDS = ("energy_tally", "energy_delta", "internal_energy", "fleck_factor")


def _build(text, rank, world, nblocks_total, capacity=0):
    from jaybenne_b200 import Mcblock, device_api
    return Mcblock(pu.make_pin(text=text), api=device_api(), rank=rank, n_ranks=world, capacity=capacity,
                   cfg_overrides={"source_nblocks": nblocks_total})


def _collect(mb, hist):
    return {"particles": mb.jb.particles(), "history": hist,
            "fields": {g: {n: mb.jb.get_field(g, n) for n in FIELDS} for g in mb.jb.local_gids}}


def _single_rank(text, nblocks_total, cycles):
    mb = _build(text, 0, 1, nblocks_total)
    mb.ProblemGenerator(pu.synthetic_density)
    hist = []
    for _ in range(cycles):
        hist.append(mb.Step())
        mb.time += mb.dt
    return _collect(mb, hist)


def _compare(one, parts, cycles, extra_counts=()):
    for c in range(cycles):
        for k in ("n_sourced", "n_absorbed", "n_end", "n_segments", "n_scatters", "n_block_crossings") + tuple(extra_counts):
            assert sum(p["history"][c][k] for p in parts) == one["history"][c][k], (c, k)
        sent = sum(p["history"][c]["n_sent"] for p in parts)
        assert sent == sum(p["history"][c]["n_received"] for p in parts) and sent > 0
        e2 = sum(p["history"][c]["e_end"] for p in parts)
        assert abs(e2 - one["history"][c]["e_end"]) <= 1e-12 * abs(e2)
    merged = {k: np.concatenate([p["particles"][k] for p in parts]) for k in one["particles"]}

    def order(p):
        return np.lexsort((p["rng"][:, 0], p["rng"][:, 1], p["rng"][:, 2]))
    o1, o2 = order(one["particles"]), order(merged)
    assert len(o1) == len(o2)
    for k in one["particles"]:
        assert np.array_equal(one["particles"][k][o1], merged[k][o2]), k
    seen = set()
    for p in parts:
        for g, f in p["fields"].items():
            seen.add(g)
            for name, arr in f.items():
                ref = one["fields"][g][name]
                scale = max(np.max(np.abs(ref)), 1e-300)
                assert np.max(np.abs(arr - ref)) / scale <= 1e-12, (g, name)
    assert seen == set(one["fields"])


def _step_ranks_hosted(mbs, t, dt):
    """radiation_step_hosted (jaybenne_b200/package.py) for several rank contexts in one process."""
    from jaybenne_b200 import capi
    jbs = [mb.jb for mb in mbs]
    n = len(jbs)
    if jbs[0].cfg.use_ddmc:   # ghost exchange of density / sie ahead of the face probabilities
        owned = {g: (jb.get_field(g, "density"), jb.get_field(g, "sie")) for jb in jbs for g in jb.local_gids}
        for jb in jbs:
            for g, (rho, sie) in owned.items():
                if g not in jb.local_gids:
                    jb.set_remote_field(g, "density", rho)
                    jb.set_remote_field(g, "sie", sie)
    for jb in jbs:
        jb.stats_begin()
        jb.UpdateDerivedTransportFields(dt)
        jb.SourcePhotons(capi.JB_SOURCE_EMISSION, t, dt)
    it = 0
    while True:
        it += 1
        assert it < 100
        out = []
        for jb in jbs:
            jb.TransportPhotons(t, dt)
            jb.SwarmExchange()
            out.append([jb.exchange_take(p) if p != jb.rank else b"" for p in range(n)])
        for r, jb in enumerate(jbs):
            for src in range(n):
                if src != r:
                    jb.exchange_give(out[src][r])
            if jb.cfg.use_ddmc:
                jb.SampleDDMCBlockFace()
        if sum(jb.CheckCompletion(t + dt) for jb in jbs) == 0:
            break
    stats = []
    for mb in mbs:
        mb.jb.EvaluateRadiationEnergy()
        mb.jb.UpdateFluid()
        st = mb.jb.stats()
        st["n_iterations"] = it
        mb.jb.UpdateDerivedMaterial()   # McblockDriver::Step's FillDerived
        stats.append(st)
    return stats


@pytest.mark.parametrize("ndim,n,nb,world", [(3, 16, 8, 2), (3, 16, 4, 4), (2, 32, 8, 2)])
def test_rank_contexts_on_one_gpu_equal_one_rank(ndim, n, nb, world):
    # no material feedback: the second cycle's source weights would inherit the last bits of
    # energy_delta, an atomic sum whose order differs between runs
    text = pu.synthetic_text(n=n, nb=nb, num_particles=40000, ndim=ndim, do_feedback=False)
    nblocks_total = (n // nb) ** ndim
    cycles = 2
    one = _single_rank(text, nblocks_total, cycles)
    mbs = [_build(text, r, world, nblocks_total, capacity=60000) for r in range(world)]
    for mb in mbs:
        mb.ProblemGenerator(pu.synthetic_density)
    hists = [[] for _ in mbs]
    for _ in range(cycles):
        sts = _step_ranks_hosted(mbs, mbs[0].time, mbs[0].dt)
        for mb, h, st in zip(mbs, hists, sts):
            h.append(st)
            mb.time += mb.dt
    _compare(one, [_collect(mb, h) for mb, h in zip(mbs, hists)], cycles)


@pytest.mark.parametrize("deck,world", [("stepdiff_smr_ddmc", 2), ("stepdiff_smr_hybrid", 4), ("stepdiff_ddmc", 2)])
def test_ddmc_rank_contexts_on_one_gpu_equal_one_rank(deck, world):
    """Multi-rank DDMC / hybrid on the static-refinement decks: face probabilities at rank
    boundaries come from the neighbour rank's density / sie (jb_set_remote_field), photons that
    change rank are re-sampled on the block face after they arrive."""
    from jaybenne_b200 import Mcblock, device_api
    ci = deck != "stepdiff_smr_hybrid"          # the hybrid deck at shipped resolution is the true hybrid
    ov = [("jaybenne", "num_particles", 20000)]
    if deck == "stepdiff_ddmc":
        ov.append(("parthenon/meshblock", "nx1", 32))   # CI runs one 128-cell block: split it in four
    cycles = 2

    def build(rank, n_ranks):
        return Mcblock(pu.make_pin(deck, overrides=ov, ci=ci), api=device_api(), rank=rank, n_ranks=n_ranks,
                       capacity=40000)
    one_mb = build(0, 1)
    one_mb.ProblemGenerator()
    h1 = []
    for _ in range(cycles):
        h1.append(one_mb.Step())
        one_mb.time += one_mb.dt
    one = _collect(one_mb, h1)
    assert h1[-1]["n_ddmc_segments"] > 0
    mbs = [build(r, world) for r in range(world)]
    for mb in mbs:
        mb.ProblemGenerator()
    hists = [[] for _ in mbs]
    for _ in range(cycles):
        sts = _step_ranks_hosted(mbs, mbs[0].time, mbs[0].dt)
        for mb, h, st in zip(mbs, hists, sts):
            h.append(st)
            mb.time += mb.dt
    _compare(one, [_collect(mb, h) for mb, h in zip(mbs, hists)], cycles, extra_counts=("n_ddmc_segments",))
    ndim = one_mb.mesh.c.ndim
    for mb in mbs:   # face probabilities are deterministic: bit-identical to the one-rank values
        for g in mb.jb.local_gids:
            for d in range(ndim):
                name = "ddmc_face_prob_x%d" % (d + 1)
                assert np.array_equal(mb.jb.get_field(g, name), one_mb.jb.get_field(g, name)), (g, name)


_WORKER = r"""
import os, pickle, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {here!r})
import torch, torch.distributed as dist
import parity_util as pu
import test_multirank_gpu as T
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
deck = {deck!r}
if deck:
    from jaybenne_b200 import Mcblock, device_api
    mb = Mcblock(pu.make_pin(deck, overrides={ov!r}, ci=True), api=device_api(), rank=rank, n_ranks=world,
                 capacity=200000)
else:
    mb = T._build(open({textfile!r}).read(), rank, world, {nblocks_total}, capacity=200000)
idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    import ctypes as C
    raw = C.create_string_buffer(128)
    assert mb.jb.api.comm_unique_id(raw) == 0
    idbuf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
dist.broadcast(idbuf, 0)
mb.jb.comm_init(bytes(idbuf.cpu().numpy().tobytes()))
mb.ProblemGenerator(None if deck else pu.synthetic_density)
hist = []
for _ in range({cycles}):
    hist.append(mb.Step())          # jb_radiation_step: NCCL exchange inside the library
    mb.time += mb.dt
res = T._collect(mb, hist)
res["exchange_mode"] = mb.jb.exchange_mode()
pickle.dump(res, open(os.path.join({outdir!r}, "rank%d.pkl" % rank), "wb"))
dist.barrier()
dist.destroy_process_group()
"""


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("p2p", ["1", "0"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_equal_one_rank(world, p2p, monkeypatch):
    """JB_P2P=1 (default): photons travel by direct stores into the peers' IPC-mapped inboxes and the loop state
    lives on the device; JB_P2P=0: count all-gather + grouped ncclSend / ncclRecv per iteration."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("JB_P2P", p2p)
    n, nb, ndim, cycles = 32, 8, 3, 2
    text = pu.synthetic_text(n=n, nb=nb, num_particles=200000, ndim=ndim, do_feedback=False)
    nblocks_total = (n // nb) ** ndim
    one = _single_rank(text, nblocks_total, cycles)
    with tempfile.TemporaryDirectory() as d:
        textfile = os.path.join(d, "deck.in")
        open(textfile, "w").write(text)
        script = os.path.join(d, "worker.py")
        open(script, "w").write(_WORKER.format(root=ROOT, here=HERE, textfile=textfile, nblocks_total=nblocks_total,
                                               cycles=cycles, outdir=d, deck=None, ov=None))
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), script]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
        parts = [pickle.load(open(os.path.join(d, f"rank{q}.pkl"), "rb")) for q in range(world)]
    assert all(p.pop("exchange_mode") == (2 if p2p == "1" else 1) for p in parts), "the requested exchange path did not run"
    _compare(one, parts, cycles)


@pytest.mark.parametrize("p2p", ["1", "0"])
@pytest.mark.parametrize("world", [2, 4])
def test_nccl_ddmc_ranks_equal_one_rank(world, p2p, monkeypatch):
    """stepdiff_smr_ddmc over NCCL: the density / sie ghost exchange and the photon exchange both
    run inside jb_radiation_step (photon exchange by direct peer stores when JB_P2P=1)."""
    import torch
    from jaybenne_b200 import Mcblock, device_api
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("JB_P2P", p2p)
    deck, ov, cycles = "stepdiff_smr_ddmc", [("jaybenne", "num_particles", 50000)], 2
    one_mb = Mcblock(pu.make_pin(deck, overrides=ov, ci=True), api=device_api())
    one_mb.ProblemGenerator()
    h1 = []
    for _ in range(cycles):
        h1.append(one_mb.Step())
        one_mb.time += one_mb.dt
    one = _collect(one_mb, h1)
    with tempfile.TemporaryDirectory() as d:
        script = os.path.join(d, "worker.py")
        open(script, "w").write(_WORKER.format(root=ROOT, here=HERE, textfile="", nblocks_total=0, cycles=cycles,
                                               outdir=d, deck=deck, ov=ov))
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), script]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
        parts = [pickle.load(open(os.path.join(d, f"rank{q}.pkl"), "rb")) for q in range(world)]
    assert all(p.pop("exchange_mode") == (2 if p2p == "1" else 1) for p in parts), "the requested exchange path did not run"
    _compare(one, parts, cycles, extra_counts=("n_ddmc_segments",))
