"""Draw stream: CPU checks of the oracle's implementation; GPU check that the device produces the same bits."""
import numpy as np
import pytest


def test_oracle_draw_moments(oracle):
    z = oracle.draws(7, 3, 11, 2, 0, 0, 0.0, 200000)
    u = oracle.draws(7, 3, 11, 2, 0, 1, 0.0, 200000)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1) < 0.01 and abs((z ** 4).mean() - 3) < 0.1
    assert 0 < u.min() and u.max() < 1 and abs(u.mean() - 0.5) < 0.005
    for a in (0.3, 1.0, 2.5, 40.0):
        g = oracle.draws(7, 3, 11, 2, 0, 2, a, 100000)
        assert abs(g.mean() - a) < 0.03 * max(1, a) and abs(g.var() - a) < 0.08 * max(1, a)
    for mu in (0.7, 42.0):  # kind 4: Poisson by inversion
        x = oracle.draws(7, 3, 11, 2, 0, 4, mu, 100000)
        assert np.all(x == np.floor(x)) and abs(x.mean() - mu) < 0.03 * max(1, mu ** 0.5) and abs(x.var() - mu) < 0.05 * max(1, mu)
    x = oracle.draws(7, 3, 11, 2, 0, 5, 0.31, 100000)  # kind 5: binomial(25, 0.31)
    assert x.min() >= 0 and x.max() <= 25 and abs(x.mean() - 7.75) < 0.03 and abs(x.var() - 25 * 0.31 * 0.69) < 0.1
    for a in (0.3, 6.5):  # kind 3: beta(a, 1 + a/2)
        b = 1.0 + 0.5 * a
        x = oracle.draws(7, 3, 11, 2, 0, 3, a, 100000)
        assert 0 < x.min() and x.max() < 1
        assert abs(x.mean() - a / (a + b)) < 0.005 and abs(x.var() - a * b / ((a + b) ** 2 * (a + b + 1))) < 0.003


def test_oracle_streams_are_keyed(oracle):
    base = oracle.draws(1, 0, 1, 0, 0, 0, 0.0, 8)
    for args in ((2, 0, 1, 0, 0), (1, 1, 1, 0, 0), (1, 0, 2, 0, 0), (1, 0, 1, 1, 0), (1, 0, 1, 0, 1)):
        assert not np.array_equal(base, oracle.draws(*args, 0, 0.0, 8))
    assert np.array_equal(base, oracle.draws(1, 0, 1, 0, 0, 0, 0.0, 8))


def test_oracle_det_math(oracle):
    import ctypes as C
    import math
    L = oracle.lib
    L.ref_det_log.restype = L.ref_det_exp.restype = C.c_double
    L.ref_det_log.argtypes = L.ref_det_exp.argtypes = [C.c_double]
    for x in (1e-300, 1e-9, 0.3, 0.70710678, 0.999999, 1.0, 1.41421357, 7.7, 1e9):
        assert abs(L.ref_det_log(x) - math.log(x)) <= 4e-16 * max(1.0, abs(math.log(x)))
    for x in (-690.0, -30.5, -1.0, -1e-9, 0.0):
        assert abs(L.ref_det_exp(x) - math.exp(x)) <= 4e-16 * math.exp(x)


@pytest.mark.gpu
def test_device_draws_bit_identical(oracle, b200):
    for kind, a in ((0, 0.0), (1, 0.0), (2, 0.4), (2, 1.0), (2, 17.3), (3, 0.3), (3, 6.5), (4, 0.7), (4, 42.0), (5, 0.31)):
        for key in ((5, 0, 1, 0, 0), (2 ** 40 + 9, 1023, 4999, 3, 77)):
            d = b200.draws(*key, kind, a, 4096)
            r = oracle.draws(*key, kind, a, 4096)
            assert np.array_equal(d, r), (kind, a, key, np.abs(d - r).max())
