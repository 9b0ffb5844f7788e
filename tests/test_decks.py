"""Host logic without a GPU: the Parthenon-style deck reader (csrc/host/deck.cpp), the
<jaybenne>/<mcblock> -> jb_config mapping (jaybenne.cpp:163-223, mcblock.cpp:40-143) and the
block-table builder (uniform + static refinement, Z-order ids, contiguous rank chunks).
"""
import os
import re

import numpy as np
import pytest

from jaybenne_b200 import JbError, ParameterInput, capi
from jaybenne_b200.decks import CI_OVERRIDES, DECKS, deck_text

REF_INPUTS = "/root/reference/inputs"


def parse_plain(text):
    out, block = {}, None
    joined, pending = [], ""
    for line in text.splitlines():
        line = line.split("#")[0].strip()
        if line.endswith("&"):          # Parthenon line continuation
            pending += line[:-1].strip()
            continue
        joined.append(pending + line)
        pending = ""
    for line in joined:
        if not line:
            continue
        m = re.match(r"<(.+)>", line)
        if m:
            block = m.group(1).strip()
            out.setdefault(block, {})
            continue
        k, v = line.split("=", 1)
        out[block][k.strip()] = v.strip()
    return out


@pytest.mark.skipif(not os.path.isdir(REF_INPUTS), reason="reference not mounted")
@pytest.mark.parametrize("name", sorted(DECKS))
def test_deck_data_equals_the_reference_input_file(name):
    ref = parse_plain(open(os.path.join(REF_INPUTS, name + ".in")).read())
    ours = parse_plain(deck_text(name))
    for block, kv in ref.items():
        for k, v in kv.items():
            assert block in ours and k in ours[block], (block, k)
            a, b = ours[block][k], v
            try:
                assert float(a) == float(b), (block, k, a, b)
            except ValueError:
                assert a.replace(" ", "") == b.replace(" ", ""), (block, k, a, b)
    for block, kv in ours.items():
        for k in kv:
            assert k in ref.get(block, {}), ("not in the reference deck", block, k)


@pytest.mark.skipif(not os.path.isdir(REF_INPUTS), reason="reference not mounted")
def test_reader_loads_reference_files_directly():
    pin = ParameterInput(path=os.path.join(REF_INPUTS, "stepdiff_smr_hybrid.in"))
    assert pin.get("jaybenne", "use_ddmc") == "true"
    assert pin.get("parthenon/static_refinement1", "x1max") == "0.25"
    assert pin.get("parthenon/output0", "swarm_variables").replace(" ", "") == "swarm.x,swarm.y"
    cfg = pin.config()
    assert cfg.use_ddmc == 1 and cfg.num_particles == 100000


def test_reader_comments_overrides_and_missing_keys():
    pin = ParameterInput(text="# c\n<a/b>\nx = 1.5 # trailing\n  y=hello world\n\n<jaybenne>\nnum_particles = 7\n")
    assert pin.get("a/b", "x") == "1.5" and pin.get("a/b", "y") == "hello world"
    assert pin.get("a/b", "z") is None and not pin.has("a/b", "z")
    pin.set("a/b", "x", 3)
    pin.set("new", "k", "v")
    assert pin.get_real("a/b", "x") == 3.0 and pin.get("new", "k") == "v"


def test_config_defaults_and_errors():
    from jaybenne_b200 import uniform_deck_text
    base = uniform_deck_text([4, 4, 4], [4, 4, 4]) + "<mcblock>\ninitial_density = 1\ninitial_temperature = 1\ninitial_radiation = none\nopacity_model = none\nscattering_model = none\n"
    cfg = ParameterInput(text=base + "<jaybenne>\nnum_particles = 10\n").config()
    assert cfg.seed == 123 and cfg.tau_ddmc == 5.0 and cfg.do_emission == 1 and cfg.do_feedback == 1
    assert cfg.eos_gm1 == pytest.approx(0.66666666667) and cfg.eos_cv == pytest.approx(1.0 / 0.66666666667, rel=1e-15)
    assert cfg.opacity_model == capi.JB_OPAC_GRAY and cfg.opacity_kappa == 0.0   # "none" == Gray(0)
    with pytest.raises(JbError):   # num_particles is required (jaybenne.cpp:163)
        ParameterInput(text=base + "<jaybenne>\ndt = 1\n").config()
    with pytest.raises(JbError):   # min_swarm_occupancy in [0, 1) (jaybenne.cpp:171)
        ParameterInput(text=base + "<jaybenne>\nnum_particles = 10\nmin_swarm_occupancy = 1.5\n").config()
    with pytest.raises(JbError):   # integrator must be rk1 (mcblock.cpp:40-41)
        ParameterInput(text=base.replace("rk1", "rk2") + "<jaybenne>\nnum_particles = 10\n").config()
    with pytest.raises(JbError):   # all six mesh BC keys are required (mcblock_driver.cpp:23-28)
        ParameterInput(text=base.replace("ox3_bc = periodic\n", "") + "<jaybenne>\nnum_particles = 10\n").config()
    with pytest.raises(JbError):   # unknown opacity model (mcblock.cpp:95-120)
        ParameterInput(text=base.replace("opacity_model = none", "opacity_model = tabular") + "<jaybenne>\nnum_particles = 10\n").config()


@pytest.mark.parametrize("name", sorted(DECKS))
def test_every_shipped_deck_configures_and_meshes(name):
    pin = ParameterInput(text=deck_text(name))
    cfg = pin.config()
    mesh = pin.mesh(1)
    m = mesh.c
    assert cfg.num_particles > 0
    vol = 0.0
    for g in range(m.nblocks):
        b = m.blocks[g]
        assert b.gid == g and b.rank == 0
        vol += np.prod([b.xmax[d] - b.xmin[d] for d in range(3)])
    assert vol == pytest.approx(np.prod([m.xmax[d] - m.xmin[d] for d in range(3)]), rel=1e-14)
    if "smr" in name:
        assert m.max_level == 1 and m.ndim == 2
    expect_bc = capi.JB_BC_REFLECT if name.startswith("stepdiff") else capi.JB_BC_PERIODIC
    assert m.swarm_bc[0] == expect_bc and m.swarm_bc[1] == expect_bc


def test_smr_ci_mesh_has_twenty_blocks_in_z_order():
    pin = ParameterInput(text=deck_text("stepdiff_smr"))
    for blk, key, val in CI_OVERRIDES["stepdiff_smr"]:
        pin.set(blk, key, val)
    mesh = pin.mesh(1)
    m = mesh.c
    levels = [m.blocks[g].level for g in range(m.nblocks)]
    assert m.nblocks == 20 and levels.count(0) == 4 and levels.count(1) == 16
    for g in range(m.nblocks):
        b = m.blocks[g]
        if b.level == 1:   # the refined region is x, y in [-0.25, 0.25]
            assert -0.25 <= b.xmin[0] and b.xmax[0] <= 0.25
            assert (b.xmax[0] - b.xmin[0]) == pytest.approx(0.125)
    # fine siblings are contiguous in gid (Z-order within the parent)
    first = levels.index(1)
    assert all(lv == 1 for lv in levels[first:first + 16]) or levels.count(1) == 16


@pytest.mark.parametrize("n_ranks", [2, 3, 8])
def test_rank_assignment_is_contiguous_and_balanced(n_ranks):
    from jaybenne_b200 import uniform_deck_text
    text = uniform_deck_text([32, 32, 32], [8, 8, 8]) + "<jaybenne>\nnum_particles = 1\n"
    mesh = ParameterInput(text=text).mesh(n_ranks)
    ranks = [mesh.c.blocks[g].rank for g in range(mesh.c.nblocks)]
    assert ranks == sorted(ranks) and set(ranks) == set(range(n_ranks))
    counts = np.bincount(ranks)
    assert counts.max() - counts.min() <= 1
    if n_ranks == 8:   # Z-order chunks of a 4x4x4 block grid are 2x2x2 octants
        for r in range(8):
            lx = np.array([[mesh.c.blocks[g].lx[d] for d in range(3)] for g in range(64) if ranks[g] == r])
            assert (lx.max(axis=0) - lx.min(axis=0) == 1).all()
