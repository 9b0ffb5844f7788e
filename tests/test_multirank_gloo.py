"""The N > 1 path on CPU: two processes over torch.distributed `gloo`, meshblocks split in
Z-order chunks, photons that leave a rank's blocks moved by the host-staged swarm exchange
(jb_exchange_take / jb_exchange_give — the transport-agnostic flavour of the exchange the
CUDA library does over NCCL), completion decided by an all-reduce of CheckCompletion.

Because random streams are keyed per photon, the union of the two ranks' swarms must equal
the single-rank swarm bit for bit, and every block's fields must be identical.
"""
import os
import pickle
import socket
import sys
import tempfile

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _build(text, rank, world, nblocks_total):
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import oracle_binding as ob
    import parity_util as pu
    from jaybenne_b200 import Mcblock
    mb = Mcblock(pu.make_pin(text=text), api=ob.oracle_api("std"), rank=rank, n_ranks=world,
                 cfg_overrides={"source_nblocks": nblocks_total})
    return mb, pu


def _collect(mb, pu, hist):
    fields = {g: {n: mb.jb.get_field(g, n) for n in ("energy_tally", "energy_delta", "internal_energy", "fleck_factor")}
              for g in mb.jb.local_gids}
    return {"particles": mb.jb.particles(), "fields": fields, "history": hist}


def _worker(rank, world, port, text, nblocks_total, cycles, outdir):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    from jaybenne_b200 import radiation_step_hosted
    mb, pu = _build(text, rank, world, nblocks_total)
    mb.ProblemGenerator(pu.synthetic_density)

    def all_to_all(outgoing):
        gathered = [None] * world
        dist.all_gather_object(gathered, outgoing)
        return [gathered[src][rank] for src in range(world) if src != rank]

    def all_reduce_sum(x):
        import torch
        t = torch.tensor([int(x)], dtype=torch.int64)
        dist.all_reduce(t)
        return int(t.item())

    hist = []
    for _ in range(cycles):
        st = mb.Step(step_fn=lambda t, dt: radiation_step_hosted(mb.jb, t, dt, all_to_all, all_reduce_sum))
        mb.time += mb.dt
        hist.append(st)
    with open(os.path.join(outdir, f"rank{rank}.pkl"), "wb") as f:
        pickle.dump(_collect(mb, pu, hist), f)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("ndim,n,nb", [(3, 8, 4), (2, 16, 8)])
def test_two_ranks_over_gloo_equal_one_rank(ndim, n, nb):
    import torch.multiprocessing as mp
    sys.path.insert(0, HERE)
    import parity_util as pu
    text = pu.synthetic_text(n=n, nb=nb, num_particles=4000, ndim=ndim)
    nblocks_total = (n // nb) ** ndim
    cycles = 2
    # single rank
    mb, _ = _build(text, 0, 1, nblocks_total)
    mb.ProblemGenerator(pu.synthetic_density)
    hist1 = []
    for _ in range(cycles):
        hist1.append(mb.Step())
        mb.time += mb.dt
    one = _collect(mb, pu, hist1)
    # two ranks
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(2, _free_port(), text, nblocks_total, cycles, d), nprocs=2, join=True)
        parts = [pickle.load(open(os.path.join(d, f"rank{r}.pkl"), "rb")) for r in range(2)]
    # counts add up; photons were exchanged
    for c in range(cycles):
        for k in ("n_sourced", "n_absorbed", "n_end", "n_segments", "n_scatters", "n_block_crossings"):
            assert sum(p["history"][c][k] for p in parts) == hist1[c][k], (c, k)
        assert sum(p["history"][c]["n_sent"] for p in parts) == sum(p["history"][c]["n_received"] for p in parts) > 0
        e2 = sum(p["history"][c]["e_end"] for p in parts)
        assert abs(e2 - hist1[c]["e_end"]) <= 1e-12 * abs(e2)
    # the union of the swarms is the single-rank swarm, bit for bit
    merged = {k: np.concatenate([p["particles"][k] for p in parts]) for k in one["particles"]}
    def order(p):
        return np.lexsort((p["rng"][:, 0], p["rng"][:, 1], p["rng"][:, 2]))
    o1, o2 = order(one["particles"]), order(merged)
    for k in one["particles"]:
        assert np.array_equal(one["particles"][k][o1], merged[k][o2]), k
    # owner-computes fields: every block identical to the single-rank run
    seen = set()
    for p in parts:
        for g, f in p["fields"].items():
            seen.add(g)
            for name, arr in f.items():
                ref = one["fields"][g][name]
                assert np.allclose(arr, ref, rtol=1e-12, atol=0.0), (g, name)
    assert seen == set(one["fields"])
