"""GPU edge cases of the sampling path against the CPU oracle: groups with no dependents, one-node declarations, one chain,
the largest supported RW_block, degenerate run lengths.  Same tolerances as test_gpu_parity.py."""
import numpy as np
import pytest

import nimble_b200 as nb
from nimble_b200 import configs
from nimble_b200 import table as T
from tests import small_models as sm
from tests.helpers import compile_on, rel_err, run

pytestmark = pytest.mark.gpu
RTOL_LP = 1e-12
RTOL_TRAJ = 1e-9


def _lmm_with_groups(grp, G, seed):
    rng = np.random.default_rng(seed)
    grp = np.asarray(grp)
    y = 2.0 + rng.standard_normal(G)[grp - 1] + 1.5 * rng.standard_normal(grp.size)
    m = nb.nimbleModel(configs.LMM_CODE, constants={"N": grp.size, "G": G, "grp": grp}, data={"y": y},
                       inits={"mu": 0.0, "sig_u": 1.0, "sig_e": 1.0, "u": np.zeros(G)})
    conf = nb.configureMCMC(m, monitors=["mu", "sig_u", "sig_e", "u"])
    for v in ("sig_u", "sig_e"):
        conf.removeSamplers(v)
        conf.addSampler(v, type="RW", control={"log": True})
    return m, conf


def _glmm_with_groups(grp, G, seed):
    rng = np.random.default_rng(seed)
    grp = np.asarray(grp)
    x = rng.standard_normal(grp.size)
    y = rng.poisson(rng.gamma(2.0, 0.5, G)[grp - 1] * np.exp(0.5 + 0.3 * x)).astype(np.float64)
    m = nb.nimbleModel(configs.GLMM_CODE, constants={"N": grp.size, "G": G, "grp": grp, "x": x}, data={"y": y},
                       inits={"a": 1.0, "b": 1.0, "b0": 0.0, "b1": 0.0, "lam": np.ones(G)})
    conf = nb.configureMCMC(m, monitors=["a", "b", "b0", "b1", "lam"])
    conf.removeSamplers("a")
    conf.addSampler("a", type="RW", control={"log": True})
    return m, conf


def _groups(G, empty, per, rng=None):
    live = [g for g in range(1, G + 1) if g not in empty]
    grp = np.repeat(live, per)
    return rng.permutation(grp) if rng is not None else grp


@pytest.mark.parametrize("build,permute,nchains", [(_lmm_with_groups, False, 5), (_lmm_with_groups, True, 5),
                                                   (_lmm_with_groups, True, 40), (_glmm_with_groups, False, 5),
                                                   (_glmm_with_groups, True, 40)])
def test_groups_without_dependents(oracle, b200, build, permute, nchains):
    """Random effects whose group holds no observation (first, middle, last and two adjacent groups): their conjugate
    update has no contributions and draws from the prior.  (The reference would hand such end nodes to
    posterior_predictive, MCMC_configuration.R:368, which draws from the same distribution.)"""
    G, empty = 14, {1, 6, 7, 14}
    grp = _groups(G, empty, 37, np.random.default_rng(3) if permute else None)
    m, conf = build(grp, G, 11)
    inits = sm.random_inits(m, nchains, 7, spread=0.1)
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, inits, 230, 20, seed=5), run(cg, inits, 230, 20, seed=5)
    assert np.all(np.isfinite(so))
    err = rel_err(sg, so)
    assert err.max() < RTOL_TRAJ, (err.max(), np.unravel_index(err.argmax(), err.shape))
    y = len(cg.builder.decls) - 1
    assert rel_err(cg.dm.calculate(y), co.dm.calculate(y)).max() < RTOL_LP


@pytest.mark.parametrize("nchains", [1, 33])
def test_single_node_declarations(oracle, b200, nchains):
    """N = 1, G = 1 (every declaration has one instance) and a one-observation logistic regression with p = 1."""
    for m, conf in (_lmm_with_groups([1], 1, 2), _glmm_with_groups([1], 1, 3), configs.logistic(N=1, p=1, seed=4)):
        inits = sm.random_inits(m, nchains, 9, spread=0.1)
        co, cg = compile_on(oracle, conf), compile_on(b200, conf)
        so, sg = run(co, inits, 230, 0, seed=2), run(cg, inits, 230, 0, seed=2)
        assert so.shape == sg.shape and np.all(np.isfinite(so))
        assert rel_err(sg, so).max() < RTOL_TRAJ


@pytest.mark.parametrize("name", ["pump", "logistic", "lmm", "lmm_perm", "occupancy", "glmm"])
def test_one_chain(oracle, b200, name):
    """nchains = 1 is the reference's own default (runMCMC(nchains = 1)): one live lane per warp everywhere."""
    m, conf = sm.SMALL[name]()
    inits = sm.random_inits(m, 1, 13, spread=0.1)
    if name == "occupancy":
        inits[:, m.var_slot["z"]:m.var_slot["z"] + 300] = 1.0
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, inits, 220, 10, seed=9), run(cg, inits, 220, 10, seed=9)
    assert so.shape == sg.shape == (1, 210, len(co.colnames))
    assert rel_err(sg, so).max() < RTOL_TRAJ
    assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < RTOL_TRAJ


def _block_model(d):
    code = "{ for (i in 1:d) { x[i] ~ dnorm(0, sd = 2)\n y[i] ~ dnorm(x[i], sd = s[i]) } }"
    rng = np.random.default_rng(d)
    m = nb.nimbleModel(code, constants={"d": d, "s": 0.5 + rng.random(d)}, data={"y": rng.standard_normal(d)},
                       inits={"x": np.zeros(d)})
    conf = nb.configureMCMC(m, nodes=[], monitors=["x"])
    return m, conf


def test_largest_block_and_one_beyond(oracle, b200):
    """RW_block at the largest supported dimension (64: the in-place Cholesky and the 16 x 64 proposal tile are sized for
    it), across adaptation steps, then one dimension beyond, which must be refused before anything runs."""
    m, conf = _block_model(T.MAX_BLOCK_D)
    conf.addSampler("x[1:%d]" % T.MAX_BLOCK_D, type="RW_block", control={"adaptInterval": 50, "scale": 0.3})
    inits = sm.random_inits(m, 18, 3, spread=0.2)
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, inits, 180, 0, seed=6), run(cg, inits, 180, 0, seed=6)
    assert rel_err(sg, so).max() < RTOL_TRAJ
    assert rel_err(cg.dm.sampler_state(0, 2), co.dm.sampler_state(0, 2)).max() < 1e-8
    assert np.array_equal(cg.dm.sampler_state(0, 1), co.dm.sampler_state(0, 1))
    assert len(np.unique(so[0, :, 0])) > 5                     # the block does move
    m, conf = _block_model(T.MAX_BLOCK_D + 1)
    conf.addSampler("x[1:%d]" % (T.MAX_BLOCK_D + 1), type="RW_block")
    with pytest.raises(ValueError, match="exceeds the supported maximum"):
        compile_on(b200, conf)._ensure(2)


def test_block_dimension_is_checked_by_the_library(b200):
    """The same limit at the C ABI: a table whose RW_block has d = 65 is rejected by nimb200_model_create."""
    m, conf = _block_model(4)
    conf.addSampler("x[1:4]", type="RW_block")
    cg = compile_on(b200, conf)
    cg.builder.samplers[0].d = T.MAX_BLOCK_D + 1
    with pytest.raises(Exception, match="RW_block dimension"):
        cg._ensure(2)


def test_degenerate_run_lengths(oracle, b200):
    """niter = 0, nburnin = niter and thin > niter - nburnin give zero rows and leave (resp. advance) the state as the
    reference's run() does (MCMC_build.R:231-263); argument errors carry the reference's messages."""
    m, conf = sm.SMALL["pump"]()
    init = sm.random_inits(m, 3, 1, 0.1)
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    s = run(cg, init, 0, 0, seed=1)
    assert s.shape == (3, 0, len(cg.colnames)) and np.array_equal(cg.dm.get_values(), init)
    for niter, nburnin, thin in ((7, 7, 1), (7, 2, 6), (1, 0, 1)):
        so, sg = run(co, init, niter, nburnin, thin, seed=2), run(cg, init, niter, nburnin, thin, seed=2)
        assert so.shape == sg.shape == (3, (niter - nburnin) // thin, len(cg.colnames))
        assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < RTOL_TRAJ
    for kw, msg in ((dict(niter=-1), "niter < 0"), (dict(niter=5, nburnin=-1), "nburnin < 0"),
                    (dict(niter=5, nburnin=6), "nburnin > niter"), (dict(niter=5, thin=0), "thin < 1")):
        with pytest.raises(ValueError, match=msg):
            cg.run_chains(init, **kw)


@pytest.mark.parametrize("nchains", [3, 40])
def test_beta_and_gamma_exponential_conjugacy_match_oracle(oracle, b200, nchains):
    """dbeta prior with dbern / dbin dependents and dgamma prior with dexp dependents: draws bit-compatible with the oracle
    (two Marsaglia-Tsang gammas on disjoint call ranges of the same counter for the beta variate)."""
    for build in (sm.beta_binomial, sm.gamma_exponential):
        m, conf, _ = build()
        inits = np.tile(m.init_vector(), (nchains, 1))
        so, sg = run(compile_on(oracle, conf), inits, 300, 0, seed=8), run(compile_on(b200, conf), inits, 300, 0, seed=8)
        assert rel_err(sg, so).max() < RTOL_TRAJ
        assert len(np.unique(so[:, :, 0])) == so[:, :, 0].size


def test_beta_conjugacy_wide_path(oracle, b200):
    """One beta target with more dependents than a sweep thread takes (> 2048): contributions by k_eval<2> + k_conj_finish."""
    rng = np.random.default_rng(5)
    N = 5000
    z = (rng.random(N) < 0.37).astype(float)
    m = nb.nimbleModel("{ p ~ dbeta(1.5, 2.5)\n for (i in 1:N) { z[i] ~ dbern(p) } }", constants={"N": N}, data={"z": z},
                       inits={"p": 0.5})
    conf = nb.configureMCMC(m)
    assert conf.samplerConfs[0].type == T.SAMPLER_CONJ_BETA
    inits = np.tile(m.init_vector(), (6, 1))
    so, sg = run(compile_on(oracle, conf), inits, 120, 0, seed=2), run(compile_on(b200, conf), inits, 120, 0, seed=2)
    assert rel_err(sg, so).max() < RTOL_TRAJ
    assert abs(so.mean() - (1.5 + z.sum()) / (4.0 + N)) < 0.002


@pytest.mark.parametrize("nchains", [2, 37])
def test_posterior_predictive_matches_oracle(oracle, b200, nchains):
    """Terminal nodes simulated on the device from the same counter-based variates as the oracle: bit-compatible draws for
    dnorm, dgamma, dbeta, dexp, dlogis and dbern targets, and the reference's own predictive fixture (test-mcmc.R:209-227)."""
    for build in (sm.predictive_nodes, sm.very_simple):
        m, conf = build()
        assert conf.samplerConfs[-1].type == T.SAMPLER_POSTERIOR_PREDICTIVE
        inits = np.tile(m.init_vector(), (nchains, 1))
        so, sg = run(compile_on(oracle, conf), inits, 300, 0, seed=4), run(compile_on(b200, conf), inits, 300, 0, seed=4)
        assert np.all(np.isfinite(so))
        assert rel_err(sg, so).max() < RTOL_TRAJ


@pytest.mark.parametrize("nchains", [3, 35])
def test_slice_sampler_matches_oracle(oracle, b200, nchains):
    """sampler_slice on the device (one thread per (target, chain), data-dependent stepping out and shrinkage) against the oracle:
    same uniforms, same candidate sequence, same adapted widths -- for a discrete and a continuous target and for latent counts
    under a binomial observation."""
    for build in (sm.slice_targets, sm.latent_counts):
        m, conf = build()
        inits = np.tile(m.init_vector(), (nchains, 1))
        co, cg = compile_on(oracle, conf), compile_on(b200, conf)
        so, sg = run(co, inits, 450, 0, seed=4), run(cg, inits, 450, 0, seed=4)
        assert np.all(np.isfinite(so))
        assert rel_err(sg, so).max() < RTOL_TRAJ
        for s, sp in enumerate(co.builder.samplers):
            if sp.type == T.SAMPLER_SLICE:
                assert rel_err(cg.dm.sampler_state(s, 0), co.dm.sampler_state(s, 0)).max() < RTOL_TRAJ
                assert np.array_equal(cg.dm.sampler_state(s, 1), co.dm.sampler_state(s, 1))


@pytest.mark.parametrize("nchains", [2, 33])
def test_categorical_sampler_and_dynamic_indexing_match_oracle(oracle, b200, nchains):
    """Mixture labels drawn by the device categorical sampler equal the oracle's draw for draw; so do the conjugate updates of
    the component means, whose dependents are found through mu[z[i]] at run time."""
    m, conf, _ = sm.gaussian_mixture()
    inits = np.tile(m.init_vector(), (nchains, 1))
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, inits, 300, 0, seed=6), run(cg, inits, 300, 0, seed=6)
    assert np.array_equal(sg[:, :, 3:], so[:, :, 3:])          # labels: exact
    assert rel_err(sg, so).max() < RTOL_TRAJ
    y = len(cg.builder.decls) - 1
    assert rel_err(cg.dm.calculate(y), co.dm.calculate(y)).max() < RTOL_LP
    bad = inits.copy()
    bad[0, m.var_slot["z"]] = 7.0                               # label outside 1..K: the indexed read is NaN on both sides
    co.dm.set_values(bad)
    cg.dm.set_values(bad)
    a, b = cg.dm.calculate(y), co.dm.calculate(y)
    assert np.isnan(a[0]) and np.isnan(b[0]) and rel_err(a[1:], b[1:]).max() < RTOL_LP


def test_missing_initial_values_simulated_like_the_oracle(oracle, b200):
    """initializeModel on the device: NaN slots are simulated from the same counter-based variates, parents first, and the run
    that follows is the oracle's run."""
    from tests.test_oracle_known_answers import _missing_inits_model
    m, conf = _missing_inits_model()
    init = np.tile(m.init_vector(), (37, 1))
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, init, 120, 0, seed=11), run(cg, init, 120, 0, seed=11)
    assert np.all(np.isfinite(so)) and rel_err(sg, so).max() < RTOL_TRAJ
    run(co, init, 0, 0, seed=5), run(cg, init, 0, 0, seed=5)
    assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < 1e-15


def test_online_waic_matches_oracle(oracle, b200):
    """Online WAIC on the device (one thread per data node walking the chains of each post-burn-in iteration) against the
    oracle's chain-by-chain accumulation: same statistics up to summation order, also across a reset = FALSE continuation."""
    from tests.test_oracle_known_answers import _waic_case
    m, conf = _waic_case()
    inits = sm.random_inits(m, 5, 3, 0.1)
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    run(co, inits, 260, 60, seed=8), run(cg, inits, 260, 60, seed=8)
    wo, wg = co.getWAIC(), cg.getWAIC()
    assert wo["n"] == wg["n"] == 200 * 5
    for k in ("WAIC", "lppd", "pWAIC"):
        assert abs(wg[k] - wo[k]) < 1e-9 * abs(wo[k]), (k, wg[k], wo[k])
    run(co, inits, 50, 0, seed=8, reset=False), run(cg, inits, 50, 0, seed=8, reset=False)
    wo, wg = co.getWAIC(), cg.getWAIC()
    assert wo["n"] == wg["n"] == 250 * 5 and abs(wg["WAIC"] - wo["WAIC"]) < 1e-9 * abs(wo["WAIC"])


def test_sweep_by_products_keep_trajectories(oracle, b200, monkeypatch):
    """The conjugate gamma sweep (Poisson family) and the binary sweep (occupancy family) refresh the cached block sums of the
    declarations they touch; the MH steps that follow start from those sums.  Sizes chosen so that the coefficient samplers are
    wide (> 2048 dependents) and the by-products are live; with them switched off the trajectories must stay the same."""
    cases = [configs.glmm(N=4200, G=30, seed=14), configs.occupancy(S=2100, J=2, seed=15)]
    for m, conf in cases:
        inits = sm.random_inits(m, 40, 8, 0.1)
        if "z" in m.var_slot:
            inits[:, m.var_slot["z"]:m.var_slot["z"] + 2100] = 1.0
        so = run(compile_on(oracle, conf), inits, 150, 0, seed=3)
        sg = run(compile_on(b200, conf), inits, 150, 0, seed=3)
        assert rel_err(sg, so).max() < RTOL_TRAJ
        monkeypatch.setenv("NIMB200_NO_BYPRODUCT", "1")
        plain = run(compile_on(b200, conf), inits, 150, 0, seed=3)
        monkeypatch.delenv("NIMB200_NO_BYPRODUCT")
        assert rel_err(plain, so).max() < RTOL_TRAJ and rel_err(plain, sg).max() < RTOL_TRAJ


def test_sampler_timing_mode(b200):
    """run(time = TRUE) / getTimes() (MCMC_build.R:281-285, 323-326): per-sampler device times, one entry per configured sampler;
    the timed run takes the ordinary launch path (no CUDA graph) and must produce the same samples."""
    m, conf = sm.SMALL["glmm"]()
    inits = sm.random_inits(m, 6, 2, 0.1)
    cg = compile_on(b200, conf)
    plain = cg.run_chains(inits, 60, 0, 1, True, 4).copy()
    with pytest.raises(RuntimeError, match="time = True"):
        cg.getTimes()
    timed = cg.run_chains(inits, 60, 0, 1, True, 4, time=True)
    assert np.array_equal(plain, timed)
    t = cg.getTimes()
    assert t.shape == (len(conf.samplerConfs),) and np.all(t > 0)
    total_ms, _ = cg.dm.last_timing()
    assert t.sum() * 1e3 <= total_ms * 1.05


def test_captured_iteration_is_replayed_by_continuation_runs(b200):
    """A `reset = FALSE` continuation replays the CUDA graph an earlier call captured (row offset, iteration offset, burn-in,
    thinning and the sample buffer are read from device memory): split runs, thinning changes and a grown sample buffer give
    the bits of one ungraphed run."""
    import os
    m, conf = sm.SMALL["logistic"]()
    inits = sm.random_inits(m, 9, 3, 0.1)
    cm = compile_on(b200, conf)
    run(cm, inits, 40, 10, seed=5)
    a1 = cm.dm.samples().copy()
    run(cm, inits, 25, 0, seed=5, reset=False)           # replayed from its first iteration
    run(cm, inits, 30, 0, thin=3, seed=5, reset=False)   # other thinning, buffer grown twice by now
    got = cm.dm.samples()
    assert cm.dm.last_allreduces()[1], "the last run should have gone through the captured graph"
    os.environ["NIMB200_GRAPH"] = "0"
    try:
        ref = compile_on(b200, conf)
        run(ref, inits, 40, 10, seed=5)
        run(ref, inits, 25, 0, seed=5, reset=False)
        run(ref, inits, 30, 0, thin=3, seed=5, reset=False)
        want = ref.dm.samples()
    finally:
        del os.environ["NIMB200_GRAPH"]
    assert got.shape == want.shape == (9, 30 + 25 + 10, got.shape[2])
    assert np.array_equal(got, want) and np.array_equal(got[:, :30], a1)
    # another seed is another graph: same answer as a fresh model run with that seed
    run(cm, inits, 30, 0, seed=6)
    fresh = compile_on(b200, conf)
    run(fresh, inits, 30, 0, seed=6)
    assert np.array_equal(cm.dm.samples(), fresh.dm.samples())


def test_samples_summary_on_device_equals_numpy(b200):
    """samplesSummary (P/R/MCMC_utils.R:315-341) computed on the device -- exact order statistics by radix selection, type-7
    quantiles, two-stage moments -- against numpy on the samples read back: per chain, pooled, with ties (a binary column)
    and with very few rows."""
    m, conf = sm.binary_model()
    cm = compile_on(b200, conf)
    run(cm, np.tile(m.init_vector(), (5, 1)), 400, 0, seed=2)
    s = cm.dm.samples()
    for lo, hi in ((0, 5), (2, 3)):
        want = s[lo:hi].reshape(-1, s.shape[-1])
        got = cm.dm.samples_summary(lo, hi)
        ref = np.stack([want.mean(0), np.median(want, 0), want.std(0, ddof=1), np.quantile(want, 0.025, 0), np.quantile(want, 0.975, 0)])
        np.testing.assert_allclose(got, ref, rtol=1e-12, atol=1e-14)
    m, conf = sm.SMALL["lmm"]()
    conf.addMonitors("u")
    cm = compile_on(b200, conf)
    for niter in (3, 257):
        run(cm, sm.random_inits(m, 9, 1, 0.1), niter, 0, seed=4)
        s = cm.dm.samples()
        want = s.reshape(-1, s.shape[-1])
        got = cm.dm.samples_summary()
        ref = np.stack([want.mean(0), np.median(want, 0), want.std(0, ddof=1), np.quantile(want, 0.025, 0), np.quantile(want, 0.975, 0)])
        np.testing.assert_allclose(got, ref, rtol=1e-11, atol=1e-13)
    out = nb.runMCMC(cm, niter=50, nchains=3, setSeed=1, summary=True)
    pooled = np.concatenate([np.asarray(out["samples"]["chain%d" % k]) for k in (1, 2, 3)])
    assert abs(out["summary"]["mu"]["Median"] - np.median(pooled[:, cm.colnames.index("mu")])) < 1e-12


def test_sampler_histories(b200):
    """nimbleOptions(MCMCsaveHistory = TRUE) (P/R/MCMC_samplers.R:394-399, 726-739): entry k of a history is what the sampler used
    during adaptation interval k -- the scale / propCov read back after interval k - 1 -- and the interval's acceptance rate."""
    m, conf = sm.basic_block("block_vector")            # RW_block on x[1:3], adaptInterval 500
    conf.samplerConfs[0].control["adaptInterval"] = 40
    conf.addSampler("x[2]", type="RW", control={"adaptInterval": 25})
    nb.nimbleOptions(MCMCsaveHistory=True)
    try:
        cm = compile_on(b200, conf)
        inits = sm.random_inits(m, 7, 2, 0.1)
        scales, covs = [], []
        run(cm, inits, 40, 0, seed=9)
        for k in range(3):
            scales.append(cm.dm.sampler_state(0, 0)[0].copy())
            covs.append(cm.dm.sampler_state(0, 2).copy())
            if k < 2:
                run(cm, inits, 40, 0, seed=9, reset=False)
        sh, ah, ch = cm.getScaleHistory(0), cm.getAcceptanceHistory(0), cm.getPropCovHistory(0)
        assert sh.shape == (3, 1, 7) and ah.shape == (3, 1, 7) and ch.shape == (3, 3, 3, 7)
        assert np.all(sh[0] == 1.0) and np.array_equal(sh[1, 0], scales[0]) and np.array_equal(sh[2, 0], scales[1])
        assert np.all((ah >= 0) & (ah <= 1)) and np.all(ah * 40 == np.round(ah * 40))
        assert np.all(ch[0] == np.eye(3)[:, :, None])
        assert np.array_equal(ch[2].reshape(9, 7), covs[1].reshape(3, 3, 7).transpose(1, 0, 2).reshape(9, 7))
        rw = cm.getScaleHistory(1)
        assert rw.shape == (4, 1, 7) and np.all(rw[0] == 1.0)   # 120 iterations / 25
        run(cm, inits, 10, 0, seed=9)                        # reset: histories start over
        assert cm.getScaleHistory(0).shape[0] == 0
    finally:
        nb.nimbleOptions(MCMCsaveHistory=False)
    with pytest.raises(RuntimeError, match="MCMCsaveHistory"):
        compile_on(b200, conf).getScaleHistory(0)
