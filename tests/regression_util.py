"""The reference's regression acceptance logic, re-hosted on Mcblock.dump() (there is no
HDF5 / Parthenon phdf here): tst/stepdiff.py:33-46 and tst/stepdiff_smr.py:35-48 (analytic
diffusion solution), tst/regression_test.py:376-419 (the `weighted_mean` comparison, which
is the default the CI uses). TEST INFRASTRUCTURE.
"""
import numpy as np
from scipy.special import erf

TAU = 1.000692e-7   # tst/stepdiff.py:33
UR0 = 7.5646e5      # tst/stepdiff.py:34
SHIFT = 0.5         # tst/stepdiff.py:35
TOL_STEPDIFF = 0.05      # tst/stepdiff.py:54
TOL_STEPDIFF_SMR = 0.3   # tst/stepdiff_smr.py:56


def ur_solution(t, x):
    return UR0 / 2.0 * (erf(((x + SHIFT) + 0.5) / (2.0 * np.sqrt(t / TAU))) - erf(((x + SHIFT) - 0.5) / (2.0 * np.sqrt(t / TAU))))


def cell_centers_x(dump):
    """X1 cell centres per block, the analysis/jhdf.py:52-73 convention (uniform within a block)."""
    nb, _, _, n1 = dump["energy_tally"].shape
    xs = np.empty((nb, n1))
    for b in range(nb):
        lo, hi = dump["xmin"][b, 0], dump["xmax"][b, 0]
        xs[b] = lo + (np.arange(n1) + 0.5) * (hi - lo) / n1
    return xs


def weighted_mean_frac_error(dump, variable="energy_tally"):
    """sum(frac_err * sol) / sum(sol) over every cell of every block, frac_err = |sol - val| / |(sol + val) / 2|."""
    val = dump[variable]
    xs = cell_centers_x(dump)
    sol = ur_solution(dump["time"], xs)[:, None, None, :] + 0.0 * val
    err = np.abs(sol - val)
    frac = err / np.abs((sol + val) / 2)
    return float(np.sum(frac * sol) / np.sum(sol))
