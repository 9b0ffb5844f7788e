"""jb_math.h (the portable FP64 elementary functions the CUDA kernels and the "portable"
oracle share) pinned against libm: <= 1 ulp on the ranges the hot path uses, exact special
values. numpy's float64 functions call the platform libm.
"""
import ctypes as C

import numpy as np

import oracle_binding as ob

LOG, SIN, COS, ACOS, EXPM1, EXP, POW4, LOG_UNIT = range(8)


def evaluate(fn, x):
    api = ob.oracle_api("portable")
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    api.math_array(fn, x.ctypes.data_as(C.POINTER(C.c_double)), y.ctypes.data_as(C.POINTER(C.c_double)), x.size)
    return y


def ulp_err(got, want):
    want = np.asarray(want, dtype=np.float64)
    spacing = np.spacing(np.abs(want))
    spacing[spacing == 0] = np.finfo(np.float64).tiny
    return np.max(np.abs(got - want) / spacing)


def test_log_uniform_draws():
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.random(200000), 10.0 ** rng.uniform(-300, 300, 50000), [1.0, 0.5, 2.0, 1e-320]])
    assert ulp_err(evaluate(LOG, x), np.log(x)) <= 1.0
    assert evaluate(LOG, [0.0])[0] == -np.inf
    assert np.isnan(evaluate(LOG, [-1.0])[0])
    assert evaluate(LOG, [1.0])[0] == 0.0


def test_log_unit_equals_log_on_the_draw_lattice():
    """log_unit (the transport kernel's logarithm) drops log's special-case branches; on uniform
    deviates n * 2^-53 it must return log's result bit for bit, including log(0) = -inf."""
    rng = np.random.default_rng(7)
    n = np.concatenate([rng.integers(0, 2 ** 53, 400000, dtype=np.int64), np.arange(0, 4096, dtype=np.int64),
                        2 ** 53 - 1 - np.arange(0, 4096, dtype=np.int64), 2 ** np.arange(0, 53, dtype=np.int64)])
    x = n.astype(np.float64) * 1.1102230246251565e-16
    a, b = evaluate(LOG, x), evaluate(LOG_UNIT, x)
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    assert b[np.where(n == 0)[0][0]] == -np.inf


def test_sincos_on_zero_two_pi():
    rng = np.random.default_rng(2)
    x = np.concatenate([2 * np.pi * rng.random(200000), np.pi * rng.random(50000), [0.0, np.pi / 2, np.pi, 2 * np.pi]])
    # relative to the function value the error near the zeros of sin/cos is bounded by the
    # argument reduction; the hot path needs absolute accuracy (direction cosines)
    assert np.max(np.abs(evaluate(SIN, x) - np.sin(x))) <= 2.3e-16
    assert np.max(np.abs(evaluate(COS, x) - np.cos(x))) <= 2.3e-16
    assert evaluate(SIN, [0.0])[0] == 0.0 and evaluate(COS, [0.0])[0] == 1.0


def test_acos():
    rng = np.random.default_rng(3)
    x = np.concatenate([2 * rng.random(200000) - 1, [-1.0, 1.0, 0.0, 0.5, -0.5]])
    assert ulp_err(evaluate(ACOS, x), np.arccos(x)) <= 1.0
    assert evaluate(ACOS, [1.0])[0] == 0.0
    assert np.isnan(evaluate(ACOS, [1.5])[0])


def test_expm1_and_exp():
    rng = np.random.default_rng(4)
    x = np.concatenate([10.0 ** rng.uniform(-20, 2.8, 100000), -(10.0 ** rng.uniform(-20, 2.8, 100000)), [0.0]])
    with np.errstate(over="ignore"):
        assert ulp_err(evaluate(EXPM1, x), np.expm1(x)) <= 1.0
        xe = np.concatenate([rng.uniform(-700, 700, 100000), [0.0, 1.0, -1.0]])
        assert ulp_err(evaluate(EXP, xe), np.exp(xe)) <= 1.0
    assert evaluate(EXPM1, [1000.0])[0] == np.inf       # B_nu overflow branch of the gray closure
    assert evaluate(EXPM1, [-1000.0])[0] == -1.0


def test_pow4_is_correctly_rounded():
    from fractions import Fraction
    rng = np.random.default_rng(5)
    x = np.concatenate([10.0 ** rng.uniform(-5, 7, 2000), [1.0, 1e5, 1e-5, 0.97]])
    got = evaluate(POW4, x)
    for xi, gi in zip(x, got):
        exact = Fraction(float(xi)) ** 4
        lo, hi = np.nextafter(gi, -np.inf), np.nextafter(gi, np.inf)
        assert abs(Fraction(float(gi)) - exact) <= min(abs(Fraction(float(lo)) - exact), abs(Fraction(float(hi)) - exact))
