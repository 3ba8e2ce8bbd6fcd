"""CPU: pin the oracle's samplers and driver against the known-answer fixtures the reference's own MCMC tests
hold (P/tests/testthat/test-mcmc.R), with the reference's tolerances.  These are the only reference fixtures
for this path that do not depend on R's Mersenne-Twister stream."""
import numpy as np
import pytest

import nimble_b200 as nb
from tests import small_models as sm
from tests.helpers import compile_on, pooled, run


def test_pump_posterior_means(oracle):
    # test-mcmc.R:164-179 (means +- tolerance); JAGS table P/inst/classic-bugs/vol1/pump/bench-test1.R
    m, conf = sm.SMALL["pump"]()
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 10000, 1000, seed=11))
    mean = dict(zip(cm.colnames, s.mean(0)))
    for name, want, tol in (("theta[1]", 0.06, 0.01), ("theta[2]", 0.10, 0.01), ("theta[9]", 1.58, 0.05),
                            ("theta[10]", 1.97, 0.05), ("alpha", 0.73, 0.1), ("beta", 0.98, 0.1)):
        assert abs(mean[name] - want) < tol, (name, mean[name])
    sd = dict(zip(cm.colnames, s.std(0)))
    assert abs(sd["alpha"] - 0.2721) < 0.03 and abs(sd["beta"] - 0.5488) < 0.06


def test_pump_default_samplers(oracle):
    # mcmcTestLog_Correct.Rout pump block: RW on alpha, conjugate_dgamma_dgamma on beta, conjugate on theta[i]
    m, conf = sm.SMALL["pump"]()
    kinds = [(s.name, s.var) for s in conf.samplerConfs]
    assert kinds == [("RW", "alpha"), ("conjugate_dgamma", "beta"), ("conjugate_dgamma", "theta")]


def test_very_simple(oracle):
    # test-mcmc.R:209-227
    m, conf = sm.very_simple()
    assert [s.name for s in conf.samplerConfs][0] == "conjugate_dnorm"
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 10000, 500, seed=3))
    x, z = s[:, cm.colnames.index("x")], s[:, cm.colnames.index("z")]
    assert abs(x.mean() - 6 / 5) < 0.1 and abs(z.mean() - 5) < 0.1
    assert abs(x.std() - 1 / np.sqrt(5)) < 0.05 and abs(z.std() - 0.5) < 0.05


@pytest.mark.parametrize("kind", ["default", "block_scalars", "block_vector", "rw"])
def test_basic_block(oracle, kind):
    # test-mcmc.R:284-332
    m, conf = sm.basic_block(kind)
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 10000, 1000, seed=5))
    assert np.all(np.abs(s.mean(0) - np.array([-2 / 3, 0, 2 / 3])) < 0.1)
    assert np.all(np.abs(s.var(0) - 1 / 3) < 0.05)


def test_binary_sampler(oracle):
    # test-mcmc.R:1379-1443
    m, conf = sm.binary_model()
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 50000, 0, seed=9))
    want = dict(a=.5, b=.6, c=.05, d=.2, e=.9, f=.9525, g=.0475, h=.5)
    for k, v in want.items():
        assert abs(s[:, cm.colnames.index(k)].mean() - v) < 0.004, k


def test_binary_out_of_bounds(oracle):
    # test-mcmc.R:1445-1469: NaN / -Inf logProbs must never be accepted
    m, conf = sm.binary_out_of_bounds()
    assert [s.name for s in conf.samplerConfs] == ["binary", "binary"]
    cm = compile_on(oracle, conf)
    s = run(cm, np.tile(m.init_vector(), (3, 1)), 100, 0, seed=0)
    assert np.all(s == 1.0)


def test_rw_block_adapts_to_correlated_target(oracle):
    # spirit of test-mcmc.R:597-666 (RW_block recovers a non-trivial covariance), on independent normals with
    # very different scales: mean (1,2,3), var (1,2,3)
    code = "{ x[1] ~ dnorm(1, var = 1)\n x[2] ~ dnorm(2, var = 2)\n x[3] ~ dnorm(3, var = 3) }"
    m = nb.nimbleModel(code, inits={"x": [0.0, 0.0, 0.0]})
    conf = nb.configureMCMC(m, nodes=[], monitors=["x"])
    conf.addSampler("x[1:3]", type="RW_block")
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 30000, 5000, seed=21))
    assert np.all(np.abs(s.mean(0) - np.array([1, 2, 3])) < 0.15)
    assert np.all(np.abs(s.var(0) - np.array([1, 2, 3])) < 0.35)
    pc = cm.dm.sampler_state(0, 2)[:, 0].reshape(3, 3)
    assert pc[0, 0] < pc[1, 1] < pc[2, 2]


def test_reset_false_continues_and_appends(oracle):
    # MCMC_build.R:243-263: reset=FALSE keeps sampler state and appends to mvSamples
    m, conf = sm.SMALL["pump"]()
    cm = compile_on(oracle, conf)
    init = np.tile(m.init_vector(), (2, 1))
    a = run(cm, init, 300, 0, seed=4).copy()
    b = run(cm, init, 200, 0, seed=4, reset=False)
    assert b.shape[1] == 500 and np.array_equal(b[:, :300], a)
    cm2 = compile_on(oracle, conf)
    c = run(cm2, init, 500, 0, seed=4)
    assert np.array_equal(b, c)  # continuation == one long run (same counter-based draws)


def test_thin_and_burnin_rows(oracle):
    m, conf = sm.SMALL["pump"]()
    cm = compile_on(oracle, conf)
    s = run(cm, np.tile(m.init_vector(), (1, 1)), 103, 13, thin=4, seed=1)
    assert s.shape[1] == (103 - 13) // 4
    full = run(compile_on(oracle, conf), np.tile(m.init_vector(), (1, 1)), 103, 0, seed=1)
    assert np.array_equal(s[0], full[0][13 + 3::4][: s.shape[1]])


def test_run_argument_errors(oracle):
    # MCMC_build.R:231-240
    m, conf = sm.SMALL["pump"]()
    cm = compile_on(oracle, conf)
    init = np.tile(m.init_vector(), (1, 1))
    for kw, msg in ((dict(niter=-1), "niter < 0"), (dict(niter=5, nburnin=-1), "nburnin < 0"),
                    (dict(niter=5, nburnin=6), "nburnin > niter"), (dict(niter=5, thin=0), "thin < 1")):
        with pytest.raises(ValueError, match=msg):
            cm.run_chains(init, **kw)


def test_beta_binomial_and_gamma_exponential_conjugacy(oracle):
    """Conjugate relationships beyond the two north_star names: every iteration is an exact posterior draw."""
    from nimble_b200 import table as T
    m, conf, (a, b) = sm.beta_binomial()
    assert [s.type for s in conf.samplerConfs] == [T.SAMPLER_CONJ_BETA] and conf.samplerConfs[0].name == "conjugate_dbeta"
    s = pooled(run(compile_on(oracle, conf), np.tile(m.init_vector(), (4, 1)), 5000, 0, seed=3))[:, 0]
    mean, var = a / (a + b), a * b / ((a + b) ** 2 * (a + b + 1))
    assert abs(s.mean() - mean) < 4 * np.sqrt(var / s.size) and abs(s.var() - var) < 0.05 * var
    m, conf, (shape, rate) = sm.gamma_exponential()
    assert [s.type for s in conf.samplerConfs] == [T.SAMPLER_CONJ_GAMMA]
    s = pooled(run(compile_on(oracle, conf), np.tile(m.init_vector(), (4, 1)), 5000, 0, seed=4))[:, 0]
    assert abs(s.mean() - shape / rate) < 4 * np.sqrt(shape / rate ** 2 / s.size)
    assert abs(s.var() - shape / rate ** 2) < 0.06 * shape / rate ** 2
    # a beta prior seen through anything but the identity link is not conjugate (falls back to RW)
    m2 = nb.nimbleModel("{ p ~ dbeta(1, 1)\n y ~ dbin(0.5 * p, 10) }", data={"y": 3.0}, inits={"p": 0.5})
    assert [s.type for s in nb.configureMCMC(m2).samplerConfs] == [T.SAMPLER_RW]


def test_posterior_predictive_sampler(oracle):
    """SURVEY.md 8f rank 1 (first part): terminal nodes get sampler_posterior_predictive (MCMC_configuration.R:283-285, 368),
    listed last (:236); each draw is the node's own distribution given the current parents (MCMC_samplers.R:97-101)."""
    from nimble_b200 import table as T
    m, conf = sm.very_simple()                      # test-mcmc.R:209-227: z is the predictive node
    assert [(s.name, s.var) for s in conf.samplerConfs] == [("conjugate_dnorm", "x"), ("posterior_predictive", "z")]
    m, conf = sm.predictive_nodes()
    assert [s.name for s in conf.samplerConfs] == ["RW"] + ["posterior_predictive"] * 8
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 30000, 1000, seed=5))
    col = {n: s[:, i] for i, n in enumerate(cm.colnames)}
    a = col["a"]
    se = lambda x: 6 * x.std() / np.sqrt(x.size / 20)          # a's chain is autocorrelated: generous Monte Carlo band
    assert abs(col["pn"].mean() - a.mean()) < se(col["pn"]) and abs(col["pn"].var() - (4 + a.var())) < 0.4
    assert abs(col["pg"].mean() - (a / 2).mean()) < se(col["pg"])
    assert abs(col["pb"].mean() - (2 / (2 + a)).mean()) < se(col["pb"]) and col["pb"].min() > 0 and col["pb"].max() <= 1
    assert abs(col["pe"].mean() - (1 / a).mean()) < 0.08 * (1 / a).mean()
    assert abs(col["pl"].mean() - a.mean()) < se(col["pl"]) and abs(col["pl"].var() - (np.pi ** 2 / 12 + a.var())) < 0.2  # dlogis(location, rate = 2)
    assert set(np.unique(col["pz"])) == {0.0, 1.0} and abs(col["pz"].mean() - 0.3) < 0.01
    assert abs(col["pp"].mean() - 3 * a.mean()) < se(col["pp"]) and np.all(col["pp"] == np.floor(col["pp"]))
    assert abs(col["pk"].mean() - 3.6) < 0.03 and abs(col["pk"].var() - 12 * 0.3 * 0.7) < 0.06 and col["pk"].max() <= 12
    # explicit assignment and its guard
    conf2 = nb.configureMCMC(m, nodes=["a"])
    conf2.addSampler("pg", type="posterior_predictive")
    assert conf2.samplerConfs[-1].type == T.SAMPLER_POSTERIOR_PREDICTIVE
    with pytest.raises(ValueError, match="non-posterior-predictive"):
        conf2.addSampler("a", type="posterior_predictive")


def test_slice_sampler(oracle):
    """sampler_slice (MCMC_samplers.R:864-972), the default of non-binary discrete nodes."""
    m, conf = sm.slice_targets()
    assert [(s.name, s.var) for s in conf.samplerConfs] == [("slice", "x"), ("slice", "g")]
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 20000, 1000, seed=1))
    assert np.all(s[:, 0] == np.floor(s[:, 0])) and s[:, 0].min() >= 0
    assert abs(s[:, 0].mean() - 3.5) < 0.05 and abs(s[:, 0].var() - 3.5) < 0.15
    assert abs(s[:, 1].mean() - 1.5) < 0.02 and abs(s[:, 1].var() - 0.75) < 0.04
    w = cm.dm.sampler_state(1, 0)                       # adapted width -> 2 x mean jump (:950-958)
    assert np.all((w > 1.2) & (w < 2.2)) and np.all(cm.dm.sampler_state(1, 1) == 20000 // 200)
    m, conf = sm.latent_counts()
    assert [(s.name, s.var) for s in conf.samplerConfs] == [("conjugate_dgamma", "lam"), ("conjugate_dbeta", "p"), ("slice", "N")]
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 6000, 1000, seed=2))
    N = s[:, 2:]
    assert np.all(N == np.floor(N)) and 3 < s[:, 0].mean() < 12 and 0.2 < s[:, 1].mean() < 0.8
    with pytest.raises(ValueError, match="more than one target"):
        conf.addSampler("N[1:3]", type="slice")


def test_categorical_sampler_and_dynamic_indexing(oracle):
    """sampler_categorical (MCMC_samplers.R:162-226) on latent mixture labels; the dependents read
    mu[z[i]] / s[z[i]] through the dynamic-index ops, and mu[k] stays conjugate (offset / coefficient found by probing)."""
    from nimble_b200 import table as T
    m, conf, (mu_t, zt) = sm.gaussian_mixture()
    assert [(s.name, s.var, len(s.positions)) for s in conf.samplerConfs] == [("conjugate_dnorm", "mu", 3), ("categorical", "z", 60)]
    cm = compile_on(oracle, conf)
    ops = [o[0] for o in cm.builder.ops]
    assert T.OP_CVAR_IDX in ops and T.OP_DATA_IDX in ops
    assert [(s.type, s.n_targets) for s in cm.builder.samplers] == [(T.SAMPLER_CONJ_NORMAL, 1)] * 3 + [(T.SAMPLER_CATEGORICAL, 60)]
    s = run(cm, np.tile(m.init_vector(), (2, 1)), 1500, 500, seed=1)
    for ch in s:                                             # per chain: the labelling is only defined up to a permutation
        mu, z = ch[:, :3].mean(0), np.round(np.median(ch[:, 3:], axis=0)).astype(int)
        order = np.argsort(mu)
        assert np.allclose(mu[order], mu_t, atol=0.5)
        relabel = np.empty(3, dtype=int)
        relabel[order] = np.arange(1, 4)
        assert (relabel[z - 1] == zt).mean() > 0.95
        assert set(np.unique(ch[:, 3:])) <= {1.0, 2.0, 3.0}
    with pytest.raises(ValueError, match="dcat"):
        conf.addSampler("mu[1]", type="categorical")


def _missing_inits_model():
    code = """{ mu ~ dnorm(3, sd = 0.01)
      s ~ dgamma(400, 200)
      q ~ dbeta(200, 200)
      for (g in 1:5) { u[g] ~ dnorm(mu, sd = 0.01) }
      for (i in 1:20) { y[i] ~ dnorm(u[grp[i]], sd = s) }
      k ~ dbern(q) }"""
    m = nb.nimbleModel(code, constants={"grp": np.repeat(np.arange(1, 6), 4)}, data={"y": np.linspace(2, 4, 20)},
                       inits={"s": 2.0})
    return m, nb.configureMCMC(m, monitors=["mu", "s", "q", "u", "k"])


def test_missing_initial_values_are_simulated(oracle):
    """initializeModel (P/R/initializeModel.R:78-139): NA non-data nodes are simulated from their distributions, parents first;
    given values are kept; a node without a variate generator must be given."""
    m, conf = _missing_inits_model()
    cm = compile_on(oracle, conf)
    init = np.tile(m.init_vector(), (200, 1))
    assert np.isnan(init[0, m.var_slot["mu"]]) and init[0, m.var_slot["s"]] == 2.0
    run(cm, init, 0, 0, seed=3)
    v = cm.dm.get_values()
    mu, u = v[:, m.var_slot["mu"]], v[:, m.var_slot["u"]:m.var_slot["u"] + 5]
    assert np.all(np.isfinite(v)) and np.all(v[:, m.var_slot["s"]] == 2.0)
    assert abs(mu.mean() - 3) < 0.005 and 0.005 < mu.std() < 0.02
    assert np.abs(u - mu[:, None]).max() < 0.06 and np.abs(u - 3).max() > np.abs(u - mu[:, None]).mean()   # children follow their own chain's parent
    assert abs(v[:, m.var_slot["q"]].mean() - 0.5) < 0.01 and set(np.unique(v[:, m.var_slot["k"]])) == {0.0, 1.0}
    again = compile_on(oracle, conf)
    run(again, init, 0, 0, seed=3)
    assert np.array_equal(again.dm.get_values(), v)                   # same seed, same simulated inits
    m2 = nb.nimbleModel("{ lam ~ dgamma(2, 1)\n n ~ dcat(w[1:3])\n y ~ dnorm(lam, sd = n) }", constants={"w": [0.2, 0.3, 0.5]},
                        data={"y": 2.0}, inits={"lam": 3.0})
    with pytest.raises(nb.ModelError, match="initial values are required for n"):
        run(compile_on(oracle, nb.configureMCMC(m2)), np.tile(m2.init_vector(), (2, 1)), 5, 0, seed=1)


def test_monitors2_and_thin2(oracle):
    """A second monitor set with its own thinning interval (configureMCMC(monitors2 =, thin2 =), runMCMC(thin2 =);
    MCMC_build.R:294-303, MCMC_run.R:145-151): each set's rows are those of a run that records that set alone."""
    m, _ = sm.SMALL["pump"]()
    conf = nb.configureMCMC(m, monitors=["alpha", "beta"], monitors2=["theta", "beta"], thin=3, thin2=5)
    cm = compile_on(oracle, conf)
    out = nb.runMCMC(cm, niter=103, nburnin=13, nchains=2, setSeed=9, summary=True)
    assert set(out) == {"samples", "samples2", "summary"}
    s1, s2 = out["samples"], out["samples2"]
    assert s1["chain1"].shape == (30, 2) and s2["chain2"].shape == (18, 11)
    assert list(s2["chain1"].colnames)[:2] == ["theta[1]", "theta[2]"] and list(s2["chain1"].colnames)[-1] == "beta"
    a = nb.runMCMC(compile_on(oracle, nb.configureMCMC(m, monitors=["alpha", "beta"], thin=3)), niter=103, nburnin=13, nchains=2, setSeed=9)
    b = nb.runMCMC(compile_on(oracle, nb.configureMCMC(m, monitors=["theta", "beta"], thin=5)), niter=103, nburnin=13, nchains=2, setSeed=9)
    for ch in ("chain1", "chain2"):
        assert np.array_equal(np.asarray(s1[ch]), np.asarray(a[ch])) and np.array_equal(np.asarray(s2[ch]), np.asarray(b[ch]))
    assert "theta[3]" in out["summary"] and "alpha" in out["summary"]
    one = nb.runMCMC(cm, niter=20, nchains=1, setSeed=1, thin2=4)
    assert np.asarray(one["samples2"]).shape == (5, 11) and np.asarray(one["samples"]).shape == (6, 2)
    # reset = FALSE with a second monitor set: both sample sets grow, thinning restarts with the call (MCMC_build.R:257-263, 279-303)
    inits = np.tile(m.init_vector(), (2, 1))
    cm.run_chains(inits, 30, 0, 3, True, 4, thin2=5)
    first1, first2 = cm.mvSamples.copy(), cm.mvSamples2.copy()
    cm.run_chains(inits, 21, 0, 3, False, 4, thin2=5)
    assert cm.mvSamples.shape[1] == 10 + 7 and cm.mvSamples2.shape[1] == 6 + 4
    assert np.array_equal(cm.mvSamples[:, :10], first1) and np.array_equal(cm.mvSamples2[:, :6], first2)
    solo = compile_on(oracle, nb.configureMCMC(m, monitors=["theta", "beta"], thin=5))
    solo.run_chains(inits, 30, 0, 5, True, 4)
    solo.run_chains(inits, 21, 0, 5, False, 4)
    assert np.array_equal(cm.mvSamples2, solo.mvSamples)


def _waic_case():
    m, _ = sm.SMALL["lmm"]()
    conf = nb.configureMCMC(m, monitors=["mu", "sig_u", "sig_e", "u"], enableWAIC=True)
    for v in ("sig_u", "sig_e"):
        conf.removeSamplers(v)
        conf.addSampler(v, type="RW", control={"log": True})
    return m, conf


def test_online_waic_equals_the_offline_formula(oracle):
    """online conditional WAIC (MCMC_WAIC.R:209-282), accumulated over every post-burn-in iteration of every
    chain, against lppd / pWAIC recomputed offline from the saved samples (the reference's calculateWAIC, MCMC_WAIC.R:49-75)."""
    import scipy.stats as st
    from scipy.special import logsumexp
    m, conf = _waic_case()
    cm = compile_on(oracle, conf)
    out = nb.runMCMC(cm, niter=400, nburnin=100, nchains=2, setSeed=3, WAIC=True)
    w = out["WAIC"]
    S = np.concatenate([np.asarray(out["samples"]["chain1"]), np.asarray(out["samples"]["chain2"])])
    cn, y, grp = list(cm.colnames), m.decls[-1].values, np.repeat(np.arange(30), 200)
    mu, se = S[:, cn.index("mu")], S[:, cn.index("sig_e")]
    u = S[:, [cn.index("u[%d]" % (g + 1)) for g in range(30)]]
    lp = st.norm.logpdf(y[None, :], mu[:, None] + u[:, grp], se[:, None])
    lppd, pw = (logsumexp(lp, axis=0) - np.log(lp.shape[0])).sum(), lp.var(axis=0, ddof=1).sum()
    assert w["n"] == 600 and abs(w["lppd"] - lppd) < 1e-8 * abs(lppd) and abs(w["pWAIC"] - pw) < 1e-8 * pw
    assert abs(w["WAIC"] + 2 * (lppd - pw)) < 1e-8 * abs(w["WAIC"])
    plain = compile_on(oracle, nb.configureMCMC(m))
    with pytest.raises(ValueError, match="enableWAIC = TRUE"):
        nb.runMCMC(plain, niter=10, WAIC=True)
