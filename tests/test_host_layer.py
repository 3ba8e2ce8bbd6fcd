"""CPU: the host layer (BUGS parsing, lowering to the node table, default sampler assignment, conjugacy
detection, error behaviour).  Mirrors the reference's configuration tests, P/tests/testthat/test-mcmc.R:1684-1900."""
import numpy as np
import pytest

import nimble_b200 as nb
from nimble_b200 import table as T
from nimble_b200.mcmc import _Lowering
from tests import small_models as sm


def lower(conf):
    return _Lowering(conf.model, conf).run()


def test_parse_nested_loops_and_links():
    m, conf = sm.SMALL["occupancy"]()
    names = [d.var for d in m.decls]
    assert names == ["b0", "a0", "b1", "a1", "z", "y"]
    y = m.decls[-1]
    assert y.n == 300 * 4 and y.dist == T.DBINOM and y.is_data
    # logit(p[i,j]) <- ... was inverted and inlined: prob = z[i] * ilogit(a0 + a1*w[i,j])
    prob = y.params[1]
    assert prob[0] == "bin" and prob[1] == T.OP_MUL and prob[3][0] == "un" and prob[3][1] == T.OP_ILOGIT


def test_default_samplers_per_config():
    kinds = {k: [(s.name, s.var) for s in f()[1].samplerConfs] for k, f in sm.SMALL.items()}
    assert kinds["pump"] == [("RW", "alpha"), ("conjugate_dgamma", "beta"), ("conjugate_dgamma", "theta")]
    assert kinds["logistic"] == [("RW", "tau_b"), ("conjugate_dnorm", "mu_b"), ("RW_block", "beta")]
    assert kinds["lmm"] == [("conjugate_dnorm", "mu"), ("RW", "sig_u"), ("RW", "sig_e"), ("conjugate_dnorm", "u")]
    assert kinds["occupancy"] == [("RW", "b0"), ("RW", "a0"), ("RW", "b1"), ("RW", "a1"), ("binary", "z")]
    assert kinds["glmm"] == [("RW", "a"), ("conjugate_dgamma", "b"), ("RW", "b0"), ("RW", "b1"),
                             ("conjugate_dgamma", "lam")]


def test_conjugacy_links():
    # MCMC_conjugacy.R:80-93,109-115; link classes as in cc_checkLinearity (:1397-1524)
    m, _ = sm.SMALL["pump"]()
    t, info = nb.check_conjugacy(m, m.decl_of("theta")[0])
    assert t == T.SAMPLER_CONJ_GAMMA and list(info.values()) == [(T.LINK_MULTIPLICATIVE, 0)]
    t, info = nb.check_conjugacy(m, m.decl_of("beta")[0])
    assert list(info.values()) == [(T.LINK_IDENTITY, T.PARAM_RATE)]
    assert nb.check_conjugacy(m, m.decl_of("alpha")[0]) is None  # dexp prior: not in the table
    m, _ = sm.SMALL["lmm"]()
    t, info = nb.check_conjugacy(m, m.decl_of("mu")[0])
    assert t == T.SAMPLER_CONJ_NORMAL and list(info.values()) == [(T.LINK_ADDITIVE, 0)]
    m, _ = sm.very_simple()
    t, info = nb.check_conjugacy(m, m.decl_of("x")[0])
    assert list(info.values()) == [(T.LINK_ADDITIVE, 0)]


def test_conjugacy_not_detected_through_nonlinear_link():
    # test-mcmc.R:1684-1793: conjugacy is lost through exp()
    code = "{ a ~ dnorm(0, 1)\n for (i in 1:3) { y[i] ~ dnorm(exp(a), 1) } }"
    m = nb.nimbleModel(code, data={"y": [0.1, 0.2, 0.3]}, inits={"a": 0.0})
    assert nb.check_conjugacy(m, m.decl_of("a")[0]) is None
    code = "{ a ~ dnorm(0, 1)\n for (i in 1:3) { y[i] ~ dnorm(3 + 2 * a * x[i], 1) } }"
    m = nb.nimbleModel(code, constants={"x": [1.0, 2.0, 3.0]}, data={"y": [0.1, 0.2, 0.3]}, inits={"a": 0.0})
    assert list(nb.check_conjugacy(m, m.decl_of("a")[0])[1].values()) == [(T.LINK_LINEAR, 0)]
    # target in two parameters of the dependent -> not conjugate
    code = "{ a ~ dgamma(1, 1)\n y ~ dnorm(a, tau = a) }"
    m = nb.nimbleModel(code, data={"y": 0.5}, inits={"a": 1.0})
    assert nb.check_conjugacy(m, m.decl_of("a")[0]) is None


def test_selections_and_vectorisation():
    m, conf = sm.SMALL["glmm"]()
    b = lower(conf)
    lam = [s for s in b.samplers if s.type == T.SAMPLER_CONJ_GAMMA and s.n_targets == 40][0]
    deps = [b.deps[lam.dep_off + e] for e in range(lam.n_dep)]
    assert [d.stage for d in deps] == [0, 1]
    assert deps[0].sel == T.SEL_SAME and deps[1].sel == T.SEL_STRIDE and deps[1].a == 100
    assert deps[1].link == T.LINK_MULTIPLICATIVE
    m, conf = sm.SMALL["lmm_perm"]()
    b = lower(conf)
    u = [s for s in b.samplers if s.n_targets == 30][0]
    dep = b.deps[u.dep_off + 1]
    assert dep.sel == T.SEL_CSR
    off, idx = b.iarr[dep.a], b.iarr[dep.b]
    grp = m.constants["grp"].astype(int)
    for g in (0, 7, 29):
        assert np.all(grp[idx[off[g]:off[g + 1]]] == g + 1)
    m, conf = sm.SMALL["occupancy"]()
    b = lower(conf)
    z = [s for s in b.samplers if s.type == T.SAMPLER_BINARY][0]
    assert z.n_targets == 300 and b.deps[z.dep_off + 1].sel == T.SEL_STRIDE and b.deps[z.dep_off + 1].a == 4


def test_fused_family_recognised_and_monitors():
    m, conf = sm.SMALL["logistic"]()
    cm_b = lower(conf)
    y = cm_b.decls[-1]
    assert y.family == T.FAM_BERN_LOGIT_INPROD and y.fam_arg[2] == 7
    assert len(cm_b.monitors) == 2 + 7
    blk = cm_b.samplers[-1]
    assert blk.type == T.SAMPLER_RW_BLOCK and blk.d == 7 and blk.n_dep == 2
    assert cm_b.deps[blk.dep_off].sel == T.SEL_RANGE and cm_b.deps[blk.dep_off + 1].sel == T.SEL_ALL


def test_non_separable_targets_fall_back_to_sequential():
    code = "{ for (j in 1:3) { b[j] ~ dnorm(0, 1) }\n for (i in 1:5) { y[i] ~ dpois(exp(inprod(X[i,1:3], b[1:3]))) } }"
    X = np.arange(15.0).reshape(5, 3) / 10
    m = nb.nimbleModel(code, constants={"X": X}, data={"y": [1, 0, 2, 3, 1]}, inits={"b": [0.0, 0.0, 0.0]})
    b = lower(nb.configureMCMC(m))
    assert [s.n_targets for s in b.samplers] == [1, 1, 1]
    assert [s.target_slot for s in b.samplers] == [0, 1, 2]


def test_errors_match_reference_behaviour():
    m, conf = sm.SMALL["pump"]()
    with pytest.raises(ValueError, match="more than one target"):      # MCMC_samplers.R:348
        conf.addSampler("theta", type="RW")
    with pytest.raises(NotImplementedError, match="no CPU fallback"):
        conf.addSampler("alpha", type="AF_slice")
    with pytest.raises(nb.ModelError):
        conf.addSampler("nosuch", type="RW")
    with pytest.raises(RuntimeError, match="compileNimble"):
        nb.buildMCMC(conf).run(10)
    with pytest.raises(TypeError):
        nb.runMCMC(nb.buildMCMC(conf), niter=10)                        # MCMC_run.R:100 (is.Cnf)
    with pytest.raises(ValueError, match="log=TRUE and reflective=TRUE"):  # MCMC_samplers.R:350
        conf.addSampler("alpha", type="RW", control={"log": True, "reflective": True})
        lower(conf)
    assert np.isnan(nb.nimbleModel("{ a ~ dnorm(0,1) }").init_vector()).all()   # missing inits stay NA until initializeModel
    with pytest.raises(nb.ModelError, match="outside the supported set"):
        nb.nimbleModel("{ a ~ dweib(1,1) }", inits={"a": 1})


def test_runmcmc_shapes_and_names(oracle):
    m, conf = sm.SMALL["pump"]()
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=oracle)
    one = nb.runMCMC(cm, niter=50, nburnin=10, nchains=1, setSeed=3)
    assert one.shape == (40, 12) and one.colnames[:3] == ["alpha", "beta", "theta[1]"]
    three = nb.runMCMC(cm, niter=50, nburnin=10, nchains=3, setSeed=3, inits=lambda: {"alpha": 2.0})
    assert list(three) == ["chain1", "chain2", "chain3"]              # MCMC_run.R:183-189
    assert not np.array_equal(three["chain1"], three["chain2"])
    out = nb.runMCMC(cm, niter=50, nchains=2, setSeed=3, summary=True)
    assert "Mean" in out["summary"]["alpha"]
    with pytest.raises(ValueError, match="length nchains"):
        nb.runMCMC(cm, niter=5, nchains=3, inits=[{}, {}])


def test_ess_estimator():
    rng = np.random.default_rng(0)
    x = rng.standard_normal(4000)
    assert 3000 < nb.effective_size(x) < 5000
    ar = np.empty(4000)
    ar[0] = 0
    for i in range(1, 4000):
        ar[i] = 0.9 * ar[i - 1] + x[i]
    ess = nb.effective_size(ar)
    assert 100 < ess < 400  # theory: n (1-rho)/(1+rho) = 210


def test_initialize_model_warnings(oracle):
    # initializeModel.R:97-139: NaN / -Inf logProbs of the initial state are reported, sampling continues
    code = "{ s ~ dgamma(1, 1)\n for (i in 1:3) { y[i] ~ dnorm(0, sd = s) } }"
    m = nb.nimbleModel(code, data={"y": [0.1, 0.2, 0.3]}, inits={"s": -1.0})
    cm = nb.compileNimble(nb.buildMCMC(nb.configureMCMC(m)), backend=oracle)
    with pytest.warns(RuntimeWarning, match="logProb is"):
        nb.runMCMC(cm, niter=5, setSeed=1)


def test_batched_ess_equals_the_scalar_estimator():
    # coda::effectiveSize restated twice (nimble_b200/ess.py): the batched form used by bench.py must give the scalar one's values
    from nimble_b200.ess import effective_size, effective_size_batch, min_ess
    rng = np.random.default_rng(0)
    phi = np.array([0.0, 0.5, 0.9, 0.99, -0.5, 0.0, 0.7])
    x = np.zeros((7, 1200))
    e = rng.normal(size=x.shape)
    for t in range(1, x.shape[1]):
        x[:, t] = phi * x[:, t - 1] + e[:, t]
    x[5] = 3.0                                        # a constant column has zero ESS and is dropped from the minimum
    want = np.array([effective_size(r) for r in x])
    got = effective_size_batch(x)
    assert np.allclose(got, want, rtol=1e-9) and got[5] == 0.0
    assert 1100 < want[0] < 1300 and want[3] < 40      # white noise ~ n, phi = 0.99 ~ n / 199
    s = np.stack([x.T, x.T[::-1]])                     # [2 chains][rows][7 cols]
    m, tot = min_ess(s)
    assert abs(m - min(t for t in tot if t > 0)) < 1e-9 and tot[5] == 0.0


def test_waic_from_rank_partials():
    """Data-sharded WAIC: the ranks' partial sums (log-sum part, sum of squares, nodes, count) add up to the unsharded
    statistics (MCMC_WAIC.R:276-279)."""
    from nimble_b200.parallel import waic_from_partials
    rng = np.random.default_rng(3)
    n, cnt = 37, 50
    lp = rng.normal(-1.0, 0.3, (cnt, n))                      # per-sample, per-node log-densities
    logmeanexp = np.log(np.exp(lp).mean(0))
    want_lppd, want_p = logmeanexp.sum(), lp.var(0, ddof=1).sum()
    logsumexp, ssq = np.log(np.exp(lp).sum(0)), ((lp - lp.mean(0)) ** 2).sum(0)
    parts = [[logsumexp[:20].sum(), ssq[:20].sum(), 20, cnt], [logsumexp[20:].sum(), ssq[20:].sum(), 17, cnt]]
    w = waic_from_partials(parts)
    assert abs(w["lppd"] - want_lppd) < 1e-10 and abs(w["pWAIC"] - want_p) < 1e-10
    assert abs(w["WAIC"] + 2 * (want_lppd - want_p)) < 1e-10
    assert np.isnan(waic_from_partials([[0.0, 0.0, 3, 1]])["WAIC"])


def test_init_matrix_equals_per_chain_init_vectors():
    """runMCMC flattens a list of per-chain inits per variable (Model.init_matrix): same rows as init_vector chain by chain --
    scalars, vectors, a scalar broadcast over a vector, missing entries (NaN: simulated by initializeModel), ragged key sets."""
    from nimble_b200 import configs
    m, _ = configs.pump()
    rng = np.random.default_rng(0)
    cases = [
        [{"alpha": 1.0 + k, "beta": 2.0, "theta": rng.random(10)} for k in range(7)],
        [{"theta": 0.5 + k} for k in range(3)],                      # scalar broadcast over theta[1:10]
        [None, {"alpha": 2.0}, {"alpha": 3.0, "beta": 1.0}],         # ragged: falls back to the per-chain path
        [{} for _ in range(4)],
    ]
    for inits in cases:
        want = np.stack([m.init_vector(i) for i in inits])
        assert np.array_equal(m.init_matrix(inits), want, equal_nan=True)
