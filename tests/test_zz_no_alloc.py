"""jb_evaluate_radiation_energy_noalloc (include/jaybenne_b200.h): the census tally for callers that cannot call
cudaMalloc — scratch in a file-scope device array, nothing allocated by the call. No counterpart in the reference.
Collected last on purpose.

Checked through the C ABI against the existing entry point jb_evaluate_radiation_energy (whose result is the
context's energy_tally field) on two calls in a row — the second call finds the first one's contents in the shared
scratch array — to the 1e-12 the other tally comparisons use (atomic order differs between launches), and that a
mesh beyond the fixed scratch size is refused with a clear error."""
import numpy as np
import pytest

import parity_util as pu
from jaybenne_b200 import Mcblock, capi, device_api

pytestmark = pytest.mark.gpu


def _tally_of_existing_entry_point(mb):
    mb.jb._check(mb.jb.api.evaluate_radiation_energy(mb.jb._ctx))
    return np.concatenate([mb.jb.get_field(g, "energy_tally").ravel() for g in mb.jb.local_gids])


def test_noalloc_tally_matches_the_existing_entry_point_twice_in_a_row():
    text = pu.synthetic_text(n=16, nb=8, num_particles=40000)
    mb = Mcblock(pu.make_pin(text=text), api=device_api())
    mb.ProblemGenerator(pu.synthetic_density)
    out = np.full(16 ** 3, -1.0)
    for call in range(2):
        ref = _tally_of_existing_entry_point(mb)
        n_ref = mb.jb.num_particles()
        got, n, e = mb.jb.evaluate_radiation_energy_noalloc(out)
        assert n == n_ref and n > 0
        assert got.shape == ref.shape and pu.rel_err(got, ref) <= 1e-12, pu.rel_err(got, ref)
        dv = (1.0 / 16) ** 3
        assert abs(e - ref.sum() * dv) <= 1e-11 * e
        # the context's own field is untouched by the new entry point
        again = np.concatenate([mb.jb.get_field(g, "energy_tally").ravel() for g in mb.jb.local_gids])
        assert np.array_equal(again, ref)
        # a different population for the second call: one cycle later (absorptions, emission, moved photons)
        mb.Step()
    # and once more on the final state, straight after the previous call's contents
    ref = _tally_of_existing_entry_point(mb)
    got, n, e = mb.jb.evaluate_radiation_energy_noalloc(out)
    assert pu.rel_err(got, ref) <= 1e-12 and n == mb.jb.num_particles()


def test_noalloc_tally_refuses_a_mesh_beyond_its_fixed_scratch():
    text = pu.synthetic_text(n=128, nb=32, num_particles=1000)   # 2 097 152 cells > JB_NOALLOC_MAX_CELLS = 262 144
    mb = Mcblock(pu.make_pin(text=text), api=device_api())
    out = np.zeros(128 ** 3)
    with pytest.raises(capi.JbError) as ei:
        mb.jb.evaluate_radiation_energy_noalloc(out)
    assert "JB_ERR_CAPACITY" in str(ei.value) and "262144" in str(ei.value)
    small = Mcblock(pu.make_pin(text=pu.synthetic_text(n=16, nb=8, num_particles=1000)), api=device_api())
    with pytest.raises(capi.JbError) as ei:
        small.jb.evaluate_radiation_energy_noalloc(np.zeros(10))
    assert "JB_ERR_INVALID" in str(ei.value)
