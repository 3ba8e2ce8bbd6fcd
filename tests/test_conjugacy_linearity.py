"""CPU: the conjugacy classifier seen from OUTSIDE the oracle.  The GPU-vs-oracle parity tests share the host lowering, so an
error in link classes or dependency lists is invisible to them (round-1 review); every expectation here is either the
reference's own unit case (P/tests/testthat/test-mcmc.R:1684-1803, 1948-1964, 2033-2099, 2172-2283) or a closed-form posterior."""
import numpy as np
import pytest

import nimble_b200 as nb
from nimble_b200 import table as T
from nimble_b200.mcmc import _linear_form, _link_code
from tests import conjugacy_cases as cc
from tests.helpers import compile_on, run


def _classify(expr, extra_latent=""):
    """cc_checkLinearity(expr, 'b') on the lowered mean of y ~ dnorm(expr, 1); a, phi, d are constants."""
    code = "{ b ~ dnorm(0, 1)\n %s y ~ dnorm(%s, 1) }" % (extra_latent, expr)
    m = nb.nimbleModel(code, constants={"a": 0.37, "phi": 1.7, "d": -2.5}, data={"y": 0.3}, inits={"b": 0.1})
    f = _linear_form(m.decls[-1].bugs_params["mean"], "b")
    return f if f in (None, "free") else (f[0], f[1])


# test-mcmc.R:2172-2236 ('cc_checkLinearity and cc_replace01 unit tests'): (offset present, coefficient present) of each case;
# the reference tags literal 0 / 1 with +1i so that `0+b` is additive and `1*b` multiplicative, not identity
@pytest.mark.parametrize("expr,want", [
    ("b", (False, False)),            # offset 0, scale 1
    ("0+b", (True, False)),           # offset 0+1i, scale 1
    ("1*b", (False, True)),           # offset 0, scale 1+1i
    ("b+a", (True, False)),
    ("b*3", (False, True)),
    ("a+phi*b", (True, True)),
    ("a+phi*(3+b)", (True, True)),
    ("b/phi", (False, True)),
    ("(b+a)/phi", (True, True)),
    ("(phi+a)/b", None),
    ("b+b", (False, True)),           # offset 0, scale 2
    ("a+d", "free"),                  # scale 0
    ("b*(2*b)", None),
    ("-b", (False, True)),
    ("b-(a-b)", (True, True)),        # offset -a, scale 2
    ("exp(b)", None),
    ("a+exp(phi*b)", None),
])
def test_check_linearity_unit_cases(expr, want):
    assert _classify(expr) == want


def test_scale_identically_zero_corner_case():
    # test-mcmc.R:1948-1962: beta[4] * 0 * alpha stays a coefficient (not folded into "no dependence")
    code = ("{ for(j in 1:4) { beta[j] ~ dnorm(0, 1) }\n"
            "y ~ dnorm(beta[1] + beta[2]*0 + beta[3]*al + beta[4]*0*al, 1) }")
    m = nb.nimbleModel(code, constants={"al": 0.8}, data={"y": 0.3}, inits={"beta": np.zeros(4)})
    f = _linear_form(m.decls[-1].bugs_params["mean"], "beta")
    assert f[:2] == (True, True) and _link_code(f) == T.LINK_LINEAR


def _links(conf, var):
    sc = [s for s in conf.samplerConfs if s.var == var][0]
    return sc.name, {conf.model.decls[k].var: v[0] for k, v in (sc.conj or {}).items()}


@pytest.mark.parametrize("form", sorted(cc.FORMS))
@pytest.mark.parametrize("intercept", [True, False])
def test_inprod_formulations_are_linear_conjugate(form, intercept):
    # test-mcmc.R:1684-1716, 1793-1803: conjugate_dnorm_dnorm_linear for beta[j] via inprod(), sum(a*b), %*% and written-out terms
    m, conf, _, _ = cc.regression(form, intercept)
    name, links = _links(conf, "beta")
    assert name == "conjugate_dnorm" and links == {"y": T.LINK_LINEAR}
    if intercept:  # b0 + ...: the additive link
        assert _links(conf, "b0") == ("conjugate_dnorm", {"y": T.LINK_ADDITIVE})


def test_nonlinear_use_is_not_conjugate():
    # test-mcmc.R:1770-1780: b0 + exp(inprod(beta[1:p], X[i,1:p])) -> RW on beta
    rng = np.random.default_rng(1)
    X, y = rng.normal(size=(5, 3)), rng.normal(size=5)
    code = ("{ for(i in 1:n) { y[i] ~ dnorm(b0 + exp(inprod(beta[1:p], X[i, 1:p])), 1) }\n"
            "for(j in 1:p) { beta[j] ~ dnorm(0, 1) }\n b0 ~ dnorm(0, 1) }")
    m = nb.nimbleModel(code, constants={"n": 5, "p": 3}, data={"y": y, "X": X}, inits={"beta": np.zeros(3), "b0": 0.0})
    conf = nb.configureMCMC(m)
    assert _links(conf, "beta")[0] == "RW" and _links(conf, "b0")[0] == "conjugate_dnorm"


def test_gamma_prior_needs_a_zero_offset():
    # dgamma prior: identity / multiplicative links only (MCMC_conjugacy.R:80-93; cc_linkCheck :1271-1291).  An inprod over
    # several elements, or th[1]*t + th[2]*u, has a non-zero offset for each element -> not conjugate
    rng = np.random.default_rng(2)
    X = rng.uniform(0.5, 1.5, size=(6, 3))
    y = rng.poisson(2.0, size=6).astype(float)
    code = "{ for(i in 1:n) { y[i] ~ dpois(inprod(X[i, 1:3], lam[1:3])) }\n for(j in 1:3) { lam[j] ~ dgamma(1, 1) } }"
    m = nb.nimbleModel(code, constants={"n": 6}, data={"y": y, "X": X}, inits={"lam": np.ones(3)})
    assert _links(nb.configureMCMC(m), "lam")[0] == "RW"
    code = "{ for(i in 1:n) { y[i] ~ dpois(th[1]*t[i] + th[2]*u[i]) }\n for(j in 1:2) { th[j] ~ dgamma(1, 1) } }"
    m = nb.nimbleModel(code, constants={"n": 6, "t": X[:, 0], "u": X[:, 1]}, data={"y": y}, inits={"th": np.ones(2)})
    assert _links(nb.configureMCMC(m), "th")[0] == "RW"
    # one element per node keeps the multiplicative link (pump: theta[i] * t[i])
    code = "{ for(i in 1:n) { y[i] ~ dpois(th[i]*t[i]) }\n for(j in 1:n) { th[j] ~ dgamma(1, 1) } }"
    m = nb.nimbleModel(code, constants={"n": 6, "t": X[:, 0]}, data={"y": y}, inits={"th": np.ones(6)})
    assert _links(nb.configureMCMC(m), "th") == ("conjugate_dgamma", {"y": T.LINK_MULTIPLICATIVE})


def test_realized_links():
    # test-mcmc.R:2033-2099 ("realized conjugacy links are working"), dnorm and dgamma cases
    rng = np.random.default_rng(3)
    data = {"y%d" % k: rng.normal(size=2) for k in range(1, 5)}
    code = ("{ for(i in 1:2) { y1[i] ~ dnorm(mu, 1) }\n for(i in 1:2) { y2[i] ~ dnorm(a + mu, 1) }\n"
            "for(i in 1:2) { y3[i] ~ dnorm(b*mu, 1) }\n for(i in 1:2) { y4[i] ~ dnorm(a + b*mu, 1) }\n mu ~ dnorm(0, 1) }")
    m = nb.nimbleModel(code, constants={"a": 0.4, "b": 1.9}, data=data, inits={"mu": 0.0})
    assert _links(nb.configureMCMC(m), "mu") == ("conjugate_dnorm", {
        "y1": T.LINK_IDENTITY, "y2": T.LINK_ADDITIVE, "y3": T.LINK_MULTIPLICATIVE, "y4": T.LINK_LINEAR})
    data = {"y1": rng.normal(size=2), "y2": rng.normal(size=2), "y3": np.array([1.0, 0.0]), "y4": np.array([2.0, 1.0])}
    code = ("{ for(i in 1:2) { y1[i] ~ dnorm(0.3, tau) }\n for(i in 1:2) { y2[i] ~ dnorm(0.3, tau/d) }\n"
            "for(i in 1:2) { y3[i] ~ dpois(tau) }\n for(i in 1:2) { y4[i] ~ dpois(tau/d) }\n tau ~ dgamma(1, 1) }")
    m = nb.nimbleModel(code, constants={"d": 2.0}, data=data, inits={"tau": 1.0})
    assert _links(nb.configureMCMC(m), "tau") == ("conjugate_dgamma", {
        "y1": T.LINK_IDENTITY, "y2": T.LINK_MULTIPLICATIVE, "y3": T.LINK_IDENTITY, "y4": T.LINK_MULTIPLICATIVE})


def test_links_with_constants():
    # test-mcmc.R:2245-2283 ('cc_checkLinearity behaves correctly with constants')
    y = np.array([0.1, -0.4, 0.9, 1.3])
    cst = {"n": 4, "x": np.arange(1.0, 5.0), "alpha": np.arange(0.0, 4.0), "al": 0.2, "tau": 1.0}
    for mean, link in (("al+beta*x[i]", T.LINK_LINEAR), ("alpha[i]+beta", T.LINK_ADDITIVE), ("alpha[i]+beta*x[i]", T.LINK_LINEAR)):
        m = nb.nimbleModel("{ beta ~ dnorm(0, 0.001)\n for(i in 1:n) { y[i] ~ dnorm(%s, tau) } }" % mean,
                           constants=cst, data={"y": y}, inits={"beta": 0.0})
        assert _links(nb.configureMCMC(m), "beta") == ("conjugate_dnorm", {"y": link}), mean
    m = nb.nimbleModel("{ beta ~ dgamma(1, 1)\n for(i in 1:n) { y[i] ~ dpois(beta*x[i]) } }",
                       constants=cst, data={"y": np.array([1.0, 3.0, 2.0, 5.0])}, inits={"beta": 1.0})
    assert _links(nb.configureMCMC(m), "beta") == ("conjugate_dgamma", {"y": T.LINK_MULTIPLICATIVE})


@pytest.mark.parametrize("form,intercept", [("inprod", False), ("inprod", True), ("sum", True), ("terms", False)])
def test_regression_posterior_closed_form_on_oracle(oracle, form, intercept):
    # Bayesian linear regression: Gibbs over scalar conjugate updates must reach N((Z'Z + I)^-1 Z'y, (Z'Z + I)^-1)
    m, conf, mean, cov = cc.regression(form, intercept)
    cm = compile_on(oracle, conf)
    s = run(cm, np.tile(m.init_vector(), (4, 1)), 6000, 500, seed=13)
    cc.check_posterior(s, mean, cov)


def test_formulations_give_equal_samples_on_oracle(oracle):
    # test-mcmc.R:1716-1760: inprod, sum and written-out terms produce the same chain (same draws, same conditional posteriors)
    out = []
    for form in ("inprod", "sum", "matmul", "terms"):
        m, conf, _, _ = cc.regression(form, True)
        out.append(run(compile_on(oracle, conf), np.tile(m.init_vector(), (2, 1)), 50, 0, seed=1))
    for o in out[1:]:
        np.testing.assert_allclose(o, out[0], rtol=1e-10, atol=1e-12)


def test_two_coefficients_of_one_variable_on_oracle(oracle):
    m, conf, mean, cov = cc.two_coefficients()
    assert _links(conf, "b") == ("conjugate_dnorm", {"y": T.LINK_LINEAR})
    s = run(compile_on(oracle, conf), np.tile(m.init_vector(), (4, 1)), 6000, 500, seed=5)
    cc.check_posterior(s, mean, cov)


def test_rmnorm_chol_relation_on_oracle(oracle):
    cc.rmnorm_chol_relation(oracle)


def test_lognormal_conjugacies_on_oracle(oracle):
    cc.check_lognormal(oracle, 4, 6000)
