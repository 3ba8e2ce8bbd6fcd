"""Closed-form fixtures for the conjugacy system, independent of the host lowering: every expected value below comes from
linear algebra on the data (numpy), never from the other back end.  Shared by the CPU (oracle) and GPU test files.

Reference fixtures restated: P/tests/testthat/test-mcmc.R:1684-1803 (dnorm-dnorm conjugacy through inprod / sum / %*% and
its refusal through exp), :2245-2283 (links with constants)."""
import numpy as np

import nimble_b200 as nb

FORMS = {
    "inprod": "inprod(beta[1:p], X[i, 1:p])",
    "inprod_swapped": "inprod(X[i, 1:p], beta[1:p])",
    "sum": "sum(beta[1:p] * X[i, 1:p])",
    "matmul": "(X[i, 1:p] %*% beta[1:p])[1, 1]",
    "terms": "beta[1]*X[i,1] + beta[2]*X[i,2] + beta[3]*X[i,3]",
}


def regression(form="inprod", intercept=True, n=40, p=3, seed=7, prior="dnorm(0, 1)"):
    """y[i] ~ dnorm([b0 +] <linear predictor>, 1), beta[j] ~ dnorm(0, 1), b0 ~ dnorm(0, 1).
    Returns (model, conf, posterior mean, posterior covariance) of (beta, [b0]) in closed form."""
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, p))
    bt = np.array([1.3, -2.2, 0.6][:p])
    y = X @ bt + (0.7 if intercept else 0.0) + rng.normal(size=n)
    mean = ("b0 + " if intercept else "") + FORMS[form]
    code = "{ for(i in 1:n) { y[i] ~ dnorm(%s, 1) }\n for(j in 1:p) { beta[j] ~ %s }\n" % (mean, prior)
    if intercept:
        code += " b0 ~ dnorm(0, 1)\n"
    code += "}"
    inits = {"beta": np.zeros(p) if prior.startswith("dnorm") else np.ones(p)}
    if intercept:
        inits["b0"] = 0.0
    m = nb.nimbleModel(code, constants={"n": n, "p": p}, data={"y": y, "X": X}, inits=inits)
    mons = ["beta", "b0"] if intercept else ["beta"]
    conf = nb.configureMCMC(m, monitors=mons)
    Z = np.column_stack([X, np.ones(n)]) if intercept else X
    cov = np.linalg.inv(Z.T @ Z + np.eye(Z.shape[1]))
    return m, conf, cov @ (Z.T @ y), cov


def two_coefficients(n=200, seed=3):
    """y[i] ~ dnorm(b[1] + b[2]*x[i], sd = 1.5): for b[1] the other element is an offset, for b[2] a coefficient
    applies as well.  Flat-ish priors, so the posterior mean is (numerically) the ridge solution."""
    rng = np.random.default_rng(seed)
    x = rng.normal(size=n)
    y = 1.9 + 2.9 * x + 1.5 * rng.normal(size=n)
    code = "{ for(i in 1:n) { y[i] ~ dnorm(b[1] + b[2]*x[i], sd = 1.5) }\n for(j in 1:2) { b[j] ~ dnorm(0, sd = 10) } }"
    m = nb.nimbleModel(code, constants={"n": n, "x": x}, data={"y": y}, inits={"b": [0.0, 0.0]})
    conf = nb.configureMCMC(m, monitors=["b"])
    Z = np.column_stack([np.ones(n), x])
    cov = np.linalg.inv(Z.T @ Z / 1.5 ** 2 + np.eye(2) / 100.0)
    return m, conf, cov @ (Z.T @ y) / 1.5 ** 2, cov


def check_posterior(samples, mean, cov, nsig=5.0):
    """Pooled chains against a Gaussian posterior: mean within nsig Monte Carlo standard errors (batch means), sd within 6 %."""
    s = samples.reshape(-1, samples.shape[-1])
    nb_ = 40
    L = s.shape[0] // nb_
    bm = s[:L * nb_].reshape(nb_, L, -1).mean(1)
    mcse = bm.std(0, ddof=1) / np.sqrt(nb_)
    got = s.mean(0)
    assert np.all(np.abs(got - mean) < nsig * mcse + 1e-3), (got, mean, mcse)
    assert np.all(np.abs(s.std(0) / np.sqrt(np.diag(cov)) - 1.0) < 0.06), (s.std(0), np.sqrt(np.diag(cov)))


def rmnorm_chol_relation(backend, chains=3, niter=6, seed=29):
    """test-dsl_dists.R:20-45: rmnorm_chol(mn, chol(cov), prec_param = FALSE) == mn + t(chol(cov)) %*% z for the z the
    generator drew.  Here through sampler_RW_block's proposal (MCMC_samplers.R:690, dists.cpp:1132-1164) with a non-diagonal
    propCov on an essentially flat target (every proposal is accepted), z replayed from the draw stream: every accepted move
    must equal scale * t(chol(propCov)) %*% z, computed here with numpy's Cholesky factor."""
    d, scale = 4, 0.7
    rng = np.random.default_rng(5)
    A = rng.normal(size=(d, d))
    cov = A @ A.T + 0.5 * np.eye(d)
    code = "{ for(j in 1:4) { x[j] ~ dnorm(0, sd = 1.0e8) } }"
    m = nb.nimbleModel(code, inits={"x": np.arange(1.0, d + 1)})
    conf = nb.configureMCMC(m, nodes=[], monitors=["x"])
    conf.addSampler("x[1:4]", type="RW_block", control={"propCov": cov, "scale": scale, "adaptive": False})
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    s = cm.run_chains(np.tile(m.init_vector(), (chains, 1)), niter, 0, 1, True, seed)
    U = np.linalg.cholesky(cov).T   # upper factor, as chol() in R
    prev = np.tile(m.init_vector(), (chains, 1))
    moved = 0
    for it in range(niter):
        for c in range(chains):
            z = backend.draws(seed, c, it + 1, 0, 0, 0, 0.0, d)   # iterations count from 1 in the counter
            want = prev[c] + scale * (U.T @ z)
            if not np.array_equal(s[c, it], prev[c]):
                moved += 1
                np.testing.assert_allclose(s[c, it], want, rtol=1e-12, atol=1e-12)
        prev = s[:, it].copy()
    assert moved >= chains * niter - 1


def lognormal_cases():
    """Normal-lognormal and gamma-lognormal conjugacy (P/R/MCMC_conjugacy.R:86, 113): a dlnorm dependent contributes through
    log(value).  Returns (model, conf, {column: (posterior mean, posterior sd)}) with the closed forms computed here."""
    rng = np.random.default_rng(11)
    n = 25
    y = np.exp(rng.normal(0.8, 0.5, n))
    code = ("{ mu ~ dnorm(0, sd = 10)\n tau ~ dgamma(2, 1)\n for(i in 1:n) { y[i] ~ dlnorm(mu, sdlog = 0.5) }\n"
            "for(i in 1:n) { w[i] ~ dlnorm(0.3, taulog = tau) }\n ypred ~ dlnorm(mu, sdlog = 0.5) }")
    w = np.exp(rng.normal(0.3, 0.7, n))
    m = nb.nimbleModel(code, constants={"n": n}, data={"y": y, "w": w}, inits={"mu": 0.0, "tau": 1.0, "ypred": 1.0})
    conf = nb.configureMCMC(m, monitors=["mu", "tau", "ypred"])
    prec = 1 / 100.0 + n / 0.25
    mu_mean, mu_sd = (np.log(y).sum() / 0.25) / prec, prec ** -0.5
    shape, rate = 2 + n / 2.0, 1 + 0.5 * np.sum((np.log(w) - 0.3) ** 2)
    return m, conf, {"mu": (mu_mean, mu_sd), "tau": (shape / rate, np.sqrt(shape) / rate)}


def check_lognormal(backend, chains, niter):
    m, conf, want = lognormal_cases()
    kinds = {s.var: s.name for s in conf.samplerConfs}
    assert kinds == {"mu": "conjugate_dnorm", "tau": "conjugate_dgamma", "ypred": "posterior_predictive"}, kinds
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    s = cm.run_chains(np.tile(m.init_vector(), (chains, 1)), niter, niter // 10, 1, True, 3)
    s = s.reshape(-1, s.shape[-1])
    for name, (mean, sd) in want.items():
        col = s[:, cm.colnames.index(name)]
        assert abs(col.mean() - mean) < 5 * sd / np.sqrt(col.size) + 1e-3 * sd, (name, col.mean(), mean)
        assert abs(col.std() / sd - 1) < 0.05, (name, col.std(), sd)
    lp = np.log(s[:, cm.colnames.index("ypred")])      # predictive: log ypred ~ N(mu, 0.5^2) mixed over the posterior of mu
    assert abs(lp.mean() - want["mu"][0]) < 0.02 and abs(lp.var() - (0.25 + want["mu"][1] ** 2)) < 0.02
    return cm, s
