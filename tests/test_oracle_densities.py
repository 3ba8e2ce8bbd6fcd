"""CPU: pin the oracle's nmath restatement (oracle/nmath_ref.c) against the 50-digit golden vectors."""
import ctypes as C
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
VEC = json.load(open(os.path.join(HERE, "golden", "density_vectors.json")))


def _call(lib, v):
    d, a = v["dist"], v["args"]
    if d == "dcat":
        pr = (C.c_double * (len(a) - 1))(*a[1:])
        return lib.ref_dcat(a[0], pr, len(a) - 1)
    name = {"dexp": "ref_dexp_nimble"}.get(d, "ref_" + d)
    return getattr(lib, name)(*a)


def test_all_golden_vectors(oracle):
    worst = {}
    for v in VEC:
        got, want = _call(oracle.lib, v), float(v["logp"])
        err = abs(got - want) / max(1.0, abs(want))
        worst[v["dist"]] = max(worst.get(v["dist"], 0.0), err)
        assert err < 5e-14, (v, got)
    assert set(worst) == {"dnorm", "dgamma", "dpois", "dbinom", "dbeta", "dlogis", "dexp", "dcat", "dlnorm"}


def test_edge_cases(oracle):
    L = oracle.lib
    inf = float("inf")
    assert L.ref_dnorm(0.0, 0.0, -1.0) != L.ref_dnorm(0.0, 0.0, -1.0)          # sd < 0 -> NaN
    assert L.ref_dnorm(0.0, 0.0, inf) == -inf                                   # as in binary out-of-bounds test
    assert L.ref_dnorm(1.0, 1.0, 0.0) == inf and L.ref_dnorm(1.0, 0.0, 0.0) == -inf
    assert L.ref_dpois(2.5, 1.0) == -inf and L.ref_dpois(-1.0, 1.0) == -inf     # non-integer / negative x
    assert L.ref_dpois(0.0, 0.0) == 0.0 and L.ref_dpois(1.0, 0.0) == -inf
    assert L.ref_dbinom(0.0, 1.0, 0.0) == 0.0 and L.ref_dbinom(1.0, 1.0, 0.0) == -inf
    assert L.ref_dbinom(1.0, 1.0, 1.0) == 0.0 and L.ref_dbinom(0.0, 1.0, 1.0) == -inf
    assert L.ref_dbinom(2.0, 1.0, 0.5) == -inf
    assert np.isnan(L.ref_dbinom(0.0, 1.0, 1.5))
    assert L.ref_dgamma(-1.0, 2.0, 1.0) == -inf and L.ref_dgamma(0.0, 0.5, 1.0) == inf
    assert np.isnan(L.ref_dgamma(1.0, 2.0, -1.0))
    assert L.ref_dbeta(1.5, 2.0, 2.0) == -inf
    assert L.ref_dexp_nimble(-0.1, 1.0) == -inf


def test_dcat_invariant_to_normalisation(oracle):
    # P/tests/testthat/test-distributions.R:259-271
    p = np.array([0.2, 0.5, 0.3])
    a = (C.c_double * 3)(*p)
    b = (C.c_double * 3)(*(7.0 * p))
    for x in (1.0, 2.0, 3.0):
        assert abs(oracle.lib.ref_dcat(x, a, 3) - oracle.lib.ref_dcat(x, b, 3)) < 1e-15
    assert oracle.lib.ref_dcat(4.0, a, 3) == -float("inf")
    neg = (C.c_double * 3)(0.2, -0.5, 0.3)
    assert np.isnan(oracle.lib.ref_dcat(1.0, neg, 3))


def test_stirlerr_bd0_identities(oracle):
    import mpmath as mp
    mp.mp.dps = 40
    L = oracle.lib
    for n in (0.5, 1.0, 7.5, 15.0, 15.5, 36.0, 81.0, 501.0, 3.3):
        x = mp.mpf(n)
        want = mp.loggamma(x + 1) - (x + 0.5) * mp.log(x) + x - mp.log(2 * mp.pi) / 2
        assert abs(L.ref_stirlerr(n) - want) < 1e-15
    for x, np_ in ((1.0, 1.05), (10.0, 10.5), (3.0, 9.0), (100.0, 99.0), (1.0, 0.93)):
        want = mp.mpf(x) * mp.log(mp.mpf(x) / mp.mpf(np_)) + mp.mpf(np_) - mp.mpf(x)
        assert abs(L.ref_bd0(x, np_) - want) < 1e-15 * max(1.0, abs(want))
