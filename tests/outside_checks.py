"""Checks of the benchmarked configurations (BASELINE.json configs[1..4]) that do not pass through the other back end: a
back end is run on a scaled-down model and its posterior is compared with what the data generator put in (or with a fit
computed here in numpy), so an error in the host lowering -- dependency lists, node selections, family matching -- that the
oracle would share shows up.  Used on the CPU with the oracle (small sizes) and on the GPU (medium sizes)."""
import json
import os

import numpy as np

import nimble_b200 as nb
from nimble_b200 import configs

HERE = os.path.dirname(os.path.abspath(__file__))


def _run(backend, m, conf, inits, niter, nburnin, seed):
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    s = cm.run_chains(np.asarray(inits, dtype=np.float64), niter, nburnin, 1, True, seed)
    return cm, s.reshape(-1, s.shape[-1])


def _cols(cm, s, name):
    idx = [i for i, c in enumerate(cm.colnames) if c == name or c.startswith(name + "[")]
    return s[:, idx]


def logistic_mle(X, y, iters=40):
    """Newton iterations for the logistic MLE and its asymptotic covariance (numpy only)."""
    b = np.zeros(X.shape[1])
    for _ in range(iters):
        pr = 1.0 / (1.0 + np.exp(-(X @ b)))
        H = X.T @ (X * (pr * (1 - pr))[:, None])
        b = b + np.linalg.solve(H, X.T @ (y - pr))
    pr = 1.0 / (1.0 + np.exp(-(X @ b)))
    return b, np.linalg.inv(X.T @ (X * (pr * (1 - pr))[:, None]))


def check_logistic(backend, N, p, chains, niter, seed=1):
    """C2: with N >> p the posterior of beta is the MLE's normal approximation (the hierarchical prior is weak against the
    likelihood): posterior mean within half a standard error of the MLE, posterior sd within 20 % of the asymptotic one,
    and the generator's beta inside 4.5 posterior sd."""
    m, conf = configs.logistic(N=N, p=p, seed=1001)
    conf.samplerConfs[-1].control["scale"] = 0.1     # start RW_block near the posterior's scale: short test runs
    inits = np.stack([m.init_vector(i) for i in configs.logistic_inits(p, chains)])
    cm, s = _run(backend, m, conf, inits, niter, niter // 2, seed)
    beta = _cols(cm, s, "beta")
    mle, cov = logistic_mle(np.asarray(m.constants["X"]), np.asarray(m.data["y"]))
    se = np.sqrt(np.diag(cov))
    assert np.all(np.abs(beta.mean(0) - mle) < 0.5 * se), (beta.mean(0) - mle) / se
    assert np.all(np.abs(beta.std(0) / se - 1) < 0.2), beta.std(0) / se
    assert np.all(np.abs(beta.mean(0) - m.truth["beta"]) < 4.5 * beta.std(0))


def _zero_mean(diff, abs_tol, nsig=6.0, nbatch=40):
    """diff[sample, k]: per-sample differences whose expectation is zero under a correct sampler (a draw minus its
    Rao-Blackwellised conditional mean).  Batch-means Monte Carlo error per column; chains are concatenated."""
    L = diff.shape[0] // nbatch
    bm = diff[:L * nbatch].reshape(nbatch, L, -1).mean(1)
    mcse = bm.std(0, ddof=1) / np.sqrt(nbatch)
    bad = np.abs(diff.mean(0)) > nsig * mcse + abs_tol
    assert not np.any(bad), (int(bad.sum()), np.abs(diff.mean(0))[bad][:5], mcse[bad][:5])


def check_lmm(backend, N, G, chains, niter, seed=2, permute=False):
    """C3: mu, sig_e, sig_u against the generator's values, and every u[g] against its conditional mean given the other
    sampled parameters, computed HERE from the data and the group memberships (n_g, ybar_g): E[u_g | mu, sig_e, sig_u, y] =
    n_g (ybar_g - mu) / (n_g + sig_e^2 / sig_u^2).  A node attributed to the wrong group moves that mean visibly when
    groups are small."""
    m, conf = configs.lmm(N=N, G=G, seed=1003, permute=permute)
    conf.addMonitors("u")
    inits = np.tile(m.init_vector(), (chains, 1))
    cm, s = _run(backend, m, conf, inits, niter, niter // 2, seed)
    mu, se_, su = (_cols(cm, s, k)[:, 0] for k in ("mu", "sig_e", "sig_u"))
    u = _cols(cm, s, "u")
    t = m.truth
    assert abs(se_.mean() - t["sig_e"]) < 4.5 * se_.std() + 0.01, (se_.mean(), se_.std())
    assert abs(su.mean() - np.std(t["u"])) < 4.5 * su.std() + 0.02, (su.mean(), su.std())
    assert abs(mu.mean() - (t["mu"] + t["u"].mean())) < 4.5 * mu.std() + 0.05, (mu.mean(), mu.std())
    y, grp = np.asarray(m.data["y"]), np.asarray(m.constants["grp"]).astype(int) - 1
    n_g = np.bincount(grp, minlength=G)
    ybar = np.bincount(grp, weights=y, minlength=G) / n_g
    # u[g] is updated last in the iteration (mu, sig_u, sig_e before it), so the recorded u is a draw given the recorded others
    w = n_g[None, :] / (n_g[None, :] + (se_ ** 2 / su ** 2)[:, None])
    _zero_mean(u - w * (ybar[None, :] - mu[:, None]), 0.01)


def check_occupancy(backend, S, J, chains, niter, seed=3):
    """C4: the four logit coefficients against the generator's values, and every latent z[i] against its conditional
    probability given the sampled coefficients, computed here from x, w and y: 1 if any visit detected, else
    psi_i prod_j (1 - p_ij) / (psi_i prod_j (1 - p_ij) + 1 - psi_i)."""
    m, conf = configs.occupancy(S=S, J=J, seed=1004)
    conf.addMonitors("z")
    inits = np.tile(m.init_vector(), (chains, 1))
    cm, s = _run(backend, m, conf, inits, niter, niter // 2, seed)
    for k in ("b0", "b1", "a0", "a1"):
        v = _cols(cm, s, k)[:, 0]
        assert abs(v.mean() - m.truth[k]) < 4.5 * v.std() + 0.01, (k, v.mean(), v.std())
        assert v.std() < 0.5   # the data are informative: a sampler that ignores them would sit on the prior (sd 1.8 - 2.5)
    # sampler order: b0, a0, b1, a1 (RW) and then the z sweep, so the recorded z is a draw given the recorded coefficients
    b0, b1, a0, a1 = (_cols(cm, s, k)[:, 0] for k in ("b0", "b1", "a0", "a1"))
    z = _cols(cm, s, "z")
    x, w, y = np.asarray(m.constants["x"]), np.asarray(m.constants["w"]), np.asarray(m.data["y"])
    sub = np.arange(0, s.shape[0], max(1, s.shape[0] // 2000))   # a thinned set of samples keeps this cheap
    psi = 1 / (1 + np.exp(-(b0[sub, None] + b1[sub, None] * x[None, :])))
    q = np.ones_like(psi)
    for j in range(J):
        q *= 1 - 1 / (1 + np.exp(-(a0[sub, None] + a1[sub, None] * w[None, :, j])))
    pz = np.where(y.max(1)[None, :] > 0, 1.0, psi * q / (psi * q + 1 - psi))
    _zero_mean(z[sub] - pz, 0.01)


def check_glmm(backend, N, G, chains, niter, seed=4):
    """C5: b1 and the identified products lam[g] * exp(b0) against the generator's values, and every lam[g] against its
    conditional mean (a + sum_g y) / (b + sum_g exp(b0 + b1 x)) computed here from the data and the group memberships."""
    m, conf = configs.glmm(N=N, G=G, seed=1005)
    conf.addMonitors("lam")
    inits = np.tile(m.init_vector(), (chains, 1))
    cm, s = _run(backend, m, conf, inits, niter, niter // 2, seed)
    a, b, b0, b1 = (_cols(cm, s, k)[:, 0] for k in ("a", "b", "b0", "b1"))
    lam = _cols(cm, s, "lam")
    assert abs(b1.mean() - m.truth["b1"]) < 4.5 * b1.std() + 0.005, (b1.mean(), b1.std())
    rate = lam * np.exp(b0)[:, None]
    want = m.truth["lam"] * np.exp(m.truth["b0"])
    z = (rate.mean(0) - want) / rate.std(0)
    assert np.all(np.abs(z) < 5.0) and abs(z.mean()) < 1.0, (np.abs(z).max(), z.mean())
    # lam is the last sampler of the iteration: the recorded lam[g] is a gamma draw given the recorded a, b, b0, b1
    x, y, grp = np.asarray(m.constants["x"]), np.asarray(m.data["y"]), np.asarray(m.constants["grp"]).astype(int) - 1
    sub = np.arange(0, s.shape[0], max(1, s.shape[0] // 4000))
    ysum = np.bincount(grp, weights=y, minlength=G)
    e = np.exp(b0[sub, None] + b1[sub, None] * x[None, :])
    esum = np.stack([np.bincount(grp, weights=e[k], minlength=G) for k in range(sub.size)])
    cond = (a[sub, None] + ysum[None, :]) / (b[sub, None] + esum)
    _zero_mean(lam[sub] - cond, 0.01 * cond.mean())


# ------------------------------------------------------------------------------------------------ golden density vectors
def golden_models():
    """One model per distribution whose data declaration evaluates every golden vector of that distribution as one node
    (tests/golden/density_vectors.json: 50-digit log-densities).  Yields (dist, model, conf, expected[n])."""
    vec = json.load(open(os.path.join(HERE, "golden", "density_vectors.json")))
    by = {}
    for v in vec:
        by.setdefault(v["dist"], []).append(v)
    decl = {
        "dnorm": "y[i] ~ dnorm(a[i], sd = b[i])", "dgamma": "y[i] ~ dgamma(a[i], scale = b[i])", "dpois": "y[i] ~ dpois(a[i])",
        "dbinom": "y[i] ~ dbin(b[i], a[i])", "dbeta": "y[i] ~ dbeta(a[i], b[i])", "dlogis": "y[i] ~ dlogis(a[i], scale = b[i])",
        "dexp": "y[i] ~ dexp(a[i])", "dlnorm": "y[i] ~ dlnorm(a[i], sdlog = b[i])",
    }
    for dist, rows in by.items():
        want = np.array([float(r["logp"]) for r in rows])
        if dist == "dcat":
            for K in sorted({len(r["args"]) - 1 for r in rows}):
                sub = [r for r in rows if len(r["args"]) - 1 == K]
                P = np.array([r["args"][1:] for r in sub])
                y = np.array([r["args"][0] for r in sub])
                code = "{ zz ~ dnorm(0, 1)\n for(i in 1:n) { y[i] ~ dcat(P[i, 1:%d]) } }" % K
                m = nb.nimbleModel(code, constants={"n": len(sub), "P": P}, data={"y": y}, inits={"zz": 0.0})
                yield "dcat%d" % K, m, nb.configureMCMC(m), np.array([float(r["logp"]) for r in sub])
            continue
        args = np.array([r["args"] for r in rows])
        cst = {"n": len(rows), "a": args[:, 1]}
        if args.shape[1] > 2:
            cst["b"] = args[:, 2]
        code = "{ zz ~ dnorm(0, 1)\n for(i in 1:n) { %s } }" % decl[dist]
        m = nb.nimbleModel(code, constants=cst, data={"y": args[:, 0]}, inits={"zz": 0.0})
        yield dist, m, nb.configureMCMC(m), want


def check_golden_vectors(backend, tol):
    seen = set()
    for dist, m, conf, want in golden_models():
        cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
        dm = cm._ensure(2)
        dm.set_values(np.tile(m.init_vector(), (2, 1)))
        di = [i for i, d in enumerate(m.decls) if d.var == "y"][0]
        got = dm.calculate_nodes(di)
        assert got.shape == (want.size, 2)
        err = np.abs(got - want[:, None]) / np.maximum(1.0, np.abs(want[:, None]))
        assert err.max() < tol, (dist, err.max(), int(err.argmax()) // 2)
        seen.add(dist.rstrip("0123456789"))
        dm.close()
    assert seen == {"dnorm", "dgamma", "dpois", "dbinom", "dbeta", "dlogis", "dexp", "dcat", "dlnorm"}
