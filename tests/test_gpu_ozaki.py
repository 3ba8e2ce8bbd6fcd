"""GPU: the tcgen05 experiment (csrc/ozaki_logit.cu) -- the logistic block sum with its contraction as exact int8 slice products
on the 5th-generation tensor cores.  Not the default path (measured slower than the FP64 mma.sync kernel), kept as a
checked diagnostic: linear predictors against numpy, block sums against the oracle's dbinom(y, 1, ilogit(eta)) sums at the
parity tolerance, the awkward shapes, and the refusal / NaN behaviour for values the slicing cannot represent."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
dp = C.POINTER(C.c_double)


def _call(lib, X, y, beta, want_eta=True):
    f = lib.nimb200_debug_ozaki
    f.argtypes = [dp, dp, C.c_int32, C.c_int32, dp, C.c_int32, dp, dp, dp]
    f.restype = C.c_int
    n, p = X.shape
    Cn = beta.shape[1]
    eta = np.empty((n, Cn)) if want_eta else None
    sums = np.empty(Cn)
    rc = f(np.ascontiguousarray(X).ctypes.data_as(dp), y.ctypes.data_as(dp), n, p, np.ascontiguousarray(beta).ctypes.data_as(dp), Cn,
           eta.ctypes.data_as(dp) if want_eta else None, sums.ctypes.data_as(dp), None)
    return rc, eta, sums


def _oracle_sums(oracle, X, y, beta):
    want = np.empty(beta.shape[1])
    for c in range(beta.shape[1]):
        eta = X @ beta[:, c]
        pr = 1.0 / (1.0 + np.exp(-eta))
        want[c] = sum(oracle.lib.ref_dbinom(float(yy), 1.0, float(pp)) for yy, pp in zip(y, pr))
    return want


@pytest.mark.parametrize("n,p,Cn,scale", [(256, 50, 64, 1.0), (1000, 50, 70, 0.3), (129, 1, 3, 2.0), (5000, 64, 130, 0.2), (777, 7, 64, 1e-3)])
def test_slice_products_reproduce_the_fp64_contraction(b200, oracle, n, p, Cn, scale):
    rng = np.random.default_rng(n + p)
    X = rng.standard_normal((n, p)) / np.sqrt(p)
    X[:, 0] = 1.0
    X[3 % n] *= 1e-6                      # rows of very different magnitude get their own exponent
    X[5 % n] *= 1e5
    beta = scale * rng.standard_normal((p, Cn))
    beta[:, 1 % Cn] *= 1e-8
    y = (rng.random(n) < 0.5).astype(np.float64)
    rc, eta, sums = _call(b200.lib, X, y, beta)
    assert rc == 0, b200._err()
    want = X @ beta
    # absolute error below 2^-50 of (row maximum x chain maximum), the bound stated in the kernel's header
    bound = 2.0 ** -48 * np.abs(X).max(1)[:, None] * np.abs(beta).max(0)[None, :] * p
    assert np.all(np.abs(eta - want) <= bound + 1e-300), np.max(np.abs(eta - want) / (bound + 1e-300))
    if n <= 1000 and Cn <= 70:
        ref = _oracle_sums(oracle, X, y, beta)
        ok = np.isfinite(ref)
        assert np.max(np.abs(sums[ok] - ref[ok]) / np.maximum(1.0, np.abs(ref[ok]))) < 1e-12


def test_values_the_slicing_cannot_represent(b200):
    rng = np.random.default_rng(1)
    X = rng.standard_normal((300, 5))
    y = np.ones(300)
    beta = rng.standard_normal((5, 4))
    beta[2, 1] = np.nan                   # a poisoned chain: its block sum is NaN, the others are untouched
    rc, _, sums = _call(b200.lib, X, y, beta, want_eta=False)
    assert rc == 0 and np.isnan(sums[1]) and np.all(np.isfinite(sums[[0, 2, 3]]))
    X[7, 2] = np.inf                      # a non-finite design matrix is refused (the caller keeps the FP64 kernel)
    rc, _, _ = _call(b200.lib, X, y, rng.standard_normal((5, 4)), want_eta=False)
    assert rc != 0 and b"not representable" in b200._err()
