"""GPU: the conjugacy system against closed forms (not against the oracle, which shares the host lowering): Bayesian linear
regression through inprod / sum / written-out terms (P/tests/testthat/test-mcmc.R:1684-1803), two coefficients of one
variable, and the rmnorm_chol relation behind sampler_RW_block's proposal (test-dsl_dists.R:20-45)."""
import numpy as np
import pytest

from tests import conjugacy_cases as cc
from tests.helpers import compile_on, rel_err, run

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("form,intercept", [("inprod", False), ("inprod", True), ("sum", True), ("terms", False)])
def test_regression_posterior_closed_form(b200, form, intercept):
    m, conf, mean, cov = cc.regression(form, intercept)
    s = run(compile_on(b200, conf), np.tile(m.init_vector(), (64, 1)), 1500, 300, seed=13)
    cc.check_posterior(s, mean, cov)


def test_two_coefficients_of_one_variable(b200):
    m, conf, mean, cov = cc.two_coefficients()
    s = run(compile_on(b200, conf), np.tile(m.init_vector(), (64, 1)), 1500, 300, seed=5)
    cc.check_posterior(s, mean, cov)


def test_formulations_give_equal_samples(b200, oracle):
    out = []
    for form in ("inprod", "sum", "matmul", "terms"):
        m, conf, _, _ = cc.regression(form, True)
        out.append(run(compile_on(b200, conf), np.tile(m.init_vector(), (5, 1)), 50, 0, seed=1))
    for o in out[1:]:
        np.testing.assert_allclose(o, out[0], rtol=1e-10, atol=1e-12)
    m, conf, _, _ = cc.regression("inprod", True)
    ref = run(compile_on(oracle, conf), np.tile(m.init_vector(), (5, 1)), 50, 0, seed=1)
    assert rel_err(out[0], ref).max() < 1e-9


def test_rmnorm_chol_relation(b200):
    cc.rmnorm_chol_relation(b200)


def test_lognormal_conjugacies(b200, oracle):
    cc.check_lognormal(b200, 64, 1500)
    m, conf, _ = cc.lognormal_cases()
    inits = np.tile(m.init_vector(), (7, 1))
    a, b = run(compile_on(b200, conf), inits, 200, 0, seed=2), run(compile_on(oracle, conf), inits, 200, 0, seed=2)
    assert rel_err(a, b).max() < 1e-9
