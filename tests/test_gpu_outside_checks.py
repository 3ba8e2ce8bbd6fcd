"""GPU: checks of the benchmarked configurations that do not pass through the oracle (tests/outside_checks.py): truth recovery
at medium sizes, where the family kernels and the vectorised sweeps are the path taken, and the 592 golden density vectors
through the device's generic evaluation."""
import pytest

from tests import outside_checks as oc

pytestmark = pytest.mark.gpu


def test_golden_vectors_on_device(b200):
    oc.check_golden_vectors(b200, 1e-12)


def test_logistic_recovers_mle(b200):
    oc.check_logistic(b200, N=40000, p=10, chains=64, niter=6000)


def test_lmm_recovers_truth(b200):
    oc.check_lmm(b200, N=200000, G=100, chains=16, niter=3000)
    oc.check_lmm(b200, N=3007, G=1000, chains=16, niter=3000)      # small unequal groups: SEL_SEG, sensitive to one misplaced node


def test_lmm_permuted_groups_recover_truth(b200):
    oc.check_lmm(b200, N=100000, G=50, chains=16, niter=3000, permute=True)


def test_occupancy_recovers_truth(b200):
    oc.check_occupancy(b200, S=20000, J=6, chains=32, niter=3000)


def test_glmm_recovers_truth(b200):
    oc.check_glmm(b200, N=100000, G=100, chains=64, niter=4000)
    oc.check_glmm(b200, N=3007, G=1000, chains=64, niter=4000)


# ------------------------------------------------------------------ full-size declarations (BASELINE.json configs[2..4])
# The family kernels are timed at these sizes; here their block sums are compared (1) with a closed form computed in numpy
# for a parameter setting where one exists, (2) with the sum over four node ranges, (3) with the oracle on a 10 000-node
# range that straddles group boundaries -- the way test_full_size_logistic_properties does it for configs[1].
import numpy as np

from nimble_b200 import configs
from tests.helpers import compile_on, rel_err

RTOL_LP = 1e-12


def _decl(cm, var):
    return [cm.lowering.decl_ids[d.index] for d in cm.model.decls if d.var == var][0]


def _ranges_and_oracle(b200, oracle, conf, inits, di, n, lo, hi):
    cg = compile_on(b200, conf)
    dg = cg._ensure(inits.shape[0])
    dg.set_values(inits)
    total = dg.calculate(di)
    q = n // 4
    parts = sum(dg.calculate(di, a, n if k == 3 else a + q) for k, a in enumerate(range(0, 4 * q, q)))
    assert rel_err(parts, total).max() < RTOL_LP, rel_err(parts, total).max()
    co = compile_on(oracle, conf)
    do = co._ensure(2)
    do.set_values(inits[:2])
    d2 = compile_on(b200, conf)._ensure(2)
    d2.set_values(inits[:2])
    assert rel_err(d2.calculate(di, lo, hi), do.calculate(di, lo, hi)).max() < RTOL_LP
    pn = rel_err(d2.calculate_nodes(di, lo, lo + 300), do.calculate_nodes(di, lo, lo + 300)).max()
    assert pn < RTOL_LP, pn
    return total


def test_full_size_lmm_properties(oracle, b200):
    m, conf = configs.lmm()                                   # N = 1e7, G = 1e4
    N, C = 10_000_000, 8
    rng = np.random.default_rng(5)
    inits = np.tile(m.init_vector(), (C, 1))
    s = m.var_slot
    inits[:, s["mu"]] = rng.normal(2, 0.5, C)
    inits[:, s["sig_e"]] = rng.uniform(0.8, 2.0, C)
    inits[:, s["u"]:s["u"] + 10_000] = rng.normal(0, 1, (C, 10_000))
    inits[0, s["mu"]], inits[0, s["sig_e"]] = 0.0, 1.0        # chain 0: every node is dnorm(y; 0, 1)
    inits[0, s["u"]:s["u"] + 10_000] = 0.0
    cm = compile_on(b200, conf)
    total = _ranges_and_oracle(b200, oracle, conf, inits, _decl(cm, "y"), N, 4_309_500, 4_319_500)
    y = np.asarray(m.data["y"])
    want = -0.5 * N * np.log(2 * np.pi) - 0.5 * float(np.sum(y * y))
    assert abs(total[0] - want) < 1e-12 * abs(want), (total[0], want)


def test_full_size_glmm_properties(oracle, b200):
    from scipy.special import gammaln
    m, conf = configs.glmm()                                  # N = 1e6, G = 1e4
    N, C = 1_000_000, 64
    rng = np.random.default_rng(6)
    inits = np.tile(m.init_vector(), (C, 1))
    s = m.var_slot
    inits[:, s["b0"]], inits[:, s["b1"]] = rng.normal(0.5, 0.2, C), rng.normal(0.3, 0.1, C)
    inits[:, s["lam"]:s["lam"] + 10_000] = rng.gamma(2.0, 0.5, (C, 10_000))
    inits[0, s["b0"]], inits[0, s["b1"]] = 0.0, 0.0           # chain 0: every node is dpois(y; 1)
    inits[0, s["lam"]:s["lam"] + 10_000] = 1.0
    cm = compile_on(b200, conf)
    total = _ranges_and_oracle(b200, oracle, conf, inits, _decl(cm, "y"), N, 430_950, 440_950)
    y = np.asarray(m.data["y"])
    want = float(np.sum(-1.0 - gammaln(y + 1.0)))
    assert abs(total[0] - want) < 1e-12 * abs(want), (total[0], want)


def test_full_size_occupancy_properties(oracle, b200):
    m, conf = configs.occupancy()                             # S = 1e5, J = 10
    S, J, C = 100_000, 10, 16
    rng = np.random.default_rng(7)
    inits = np.tile(m.init_vector(), (C, 1))
    s = m.var_slot
    for k, (mu_, sd_) in {"b0": (0.3, 0.2), "b1": (1.0, 0.2), "a0": (-0.5, 0.2), "a1": (0.7, 0.2)}.items():
        inits[:, s[k]] = rng.normal(mu_, sd_, C)
        inits[0, s[k]] = 0.0                                  # chain 0: psi = p = 1/2 everywhere, z = 1
    cm = compile_on(b200, conf)
    ty = _ranges_and_oracle(b200, oracle, conf, inits, _decl(cm, "y"), S * J, 431_005, 441_005)
    tz = _ranges_and_oracle(b200, oracle, conf, inits, _decl(cm, "z"), S, 43_100, 53_100)
    assert abs(ty[0] - S * J * np.log(0.5)) < 1e-12 * S * J and abs(tz[0] - S * np.log(0.5)) < 1e-12 * S
