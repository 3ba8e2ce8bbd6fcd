"""GPU: the CUDA path (through the C ABI, libnimb200.so) against the CPU oracle on identical lowered tables,
identical initial values and identical counter-based draws.

Tolerances (north_star): log-densities and block sums within 1e-12 relative; MCMC trajectories are compared
at 1e-9 relative over a few hundred iterations (draws are bit-identical, densities differ by <= 1-2 ulp through
libm vs CUDA exp/log/pow/lgamma, block sums by summation order), which also proves every accept/reject decision
agreed -- one flipped decision would move a chain by O(1)."""
import os

import numpy as np
import pytest

import nimble_b200 as nb
from nimble_b200 import configs
from nimble_b200 import table as T
from tests import small_models as sm
from tests.helpers import compile_on, pooled, rel_err, run

pytestmark = pytest.mark.gpu
RTOL_LP = 1e-12
RTOL_TRAJ = 1e-9


def pair(oracle, b200, conf, nchains, inits):
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    do, dg = co._ensure(nchains), cg._ensure(nchains)
    do.set_values(inits)
    dg.set_values(inits)
    return co, cg, do, dg


def test_set_get_values_roundtrip(b200):
    m, conf = sm.SMALL["glmm"]()
    cg = compile_on(b200, conf)
    v = sm.random_inits(m, 7, 1)
    d = cg._ensure(7)
    d.set_values(v)
    assert np.array_equal(d.get_values(), v)


@pytest.mark.parametrize("nchains", [1, 9, 70])
def test_per_node_logprob_every_distribution(oracle, b200, nchains):
    m, conf = sm.all_dists()
    inits = sm.random_inits(m, nchains, 42)
    co, cg, do, dg = pair(oracle, b200, conf, nchains, inits)
    for di, d in enumerate(co.builder.decls):
        a, b = dg.calculate_nodes(di), do.calculate_nodes(di)
        assert a.shape == (d.n, nchains)
        assert np.all(np.isfinite(b)), T.DIST_NAMES[d.dist]
        assert rel_err(a, b).max() < RTOL_LP, (T.DIST_NAMES[d.dist], rel_err(a, b).max())


def test_special_values_match(oracle, b200):
    # -Inf / NaN / +Inf outcomes must agree exactly (they drive early rejection in the samplers)
    code = """{ s ~ dgamma(1,1)
      for (i in 1:6) { yn[i] ~ dnorm(0, sd = s - c[i])
                       yp[i] ~ dpois(s * c[i])
                       yb[i] ~ dbern(s * c[i] / 3)
                       yg[i] ~ dgamma(c[i], rate = s) } }"""
    consts = {"c": [0.0, 0.5, 1.0, 2.0, 3.0, 4.0]}
    data = {"yn": [0, 0, 1, 0, 0, 0], "yp": [0, 1, 2.5, -1, 3, 0], "yb": [0, 1, 1, 0, 1, 2], "yg": [0, 0, 0, -1, 1, 2]}
    m = nb.nimbleModel(code, constants=consts, data=data, inits={"s": 1.0})
    conf = nb.configureMCMC(m, useConjugacy=False)
    co, cg, do, dg = pair(oracle, b200, conf, 2, np.array([[1.0], [2.0]]))
    for di in range(len(co.builder.decls)):
        a, b = dg.calculate_nodes(di), do.calculate_nodes(di)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b)), di
        assert rel_err(a, b).max() < RTOL_LP


def test_calculate_calculateDiff_getLogProb_semantics(oracle, b200):
    m, conf = sm.SMALL["glmm"]()
    co, cg, do, dg = pair(oracle, b200, conf, 5, sm.random_inits(m, 5, 3))
    y = len(co.builder.decls) - 1
    s_g, s_o = dg.calculate(y), do.calculate(y)
    assert rel_err(s_g, s_o).max() < RTOL_LP
    assert rel_err(dg.getLogProb(y), s_g).max() == 0.0              # stored block logProb
    assert rel_err(dg.calculate(y, 100, 1300), do.calculate(y, 100, 1300)).max() < RTOL_LP
    v2 = sm.random_inits(m, 5, 4)
    dg.set_values(v2)
    do.set_values(v2)
    d_g, d_o = dg.calculateDiff(y), do.calculateDiff(y)             # new - stored, stored <- new
    assert np.max(np.abs(d_g - d_o)) < 1e-9 * np.max(np.abs(s_o))
    assert rel_err(dg.getLogProb(y), do.getLogProb(y)).max() < RTOL_LP
    assert np.max(np.abs(dg.calculateDiff(y))) == 0.0
    with pytest.raises(Exception, match="range"):
        dg.calculate(y, 10, 10 ** 9)


@pytest.mark.parametrize("name,nchains", [("pump", 4), ("pump", 33), ("logistic", 5), ("lmm", 5), ("lmm_perm", 3),
                                          ("occupancy", 6), ("glmm", 5)])
def test_mcmc_trajectories_match_oracle(oracle, b200, name, nchains):
    m, conf = sm.SMALL[name]()
    inits = sm.random_inits(m, nchains, 7, spread=0.1)
    if name == "occupancy":
        inits[:, m.var_slot["z"]:m.var_slot["z"] + 300] = 1.0
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    niter = 450  # crosses the adaptation steps at 200 and 400
    so, sg = run(co, inits, niter, 50, seed=5), run(cg, inits, niter, 50, seed=5)
    assert so.shape == sg.shape == (nchains, 400, len(co.colnames))
    err = rel_err(sg, so)
    assert err.max() < RTOL_TRAJ, (name, err.max(), np.unravel_index(err.argmax(), err.shape))
    assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < RTOL_TRAJ
    for s, sp in enumerate(co.builder.samplers):
        if sp.type in (T.SAMPLER_RW, T.SAMPLER_RW_BLOCK):
            assert rel_err(cg.dm.sampler_state(s, 0), co.dm.sampler_state(s, 0)).max() < RTOL_TRAJ
            assert np.array_equal(cg.dm.sampler_state(s, 1), co.dm.sampler_state(s, 1))
        if sp.type == T.SAMPLER_RW_BLOCK:
            assert rel_err(cg.dm.sampler_state(s, 2), co.dm.sampler_state(s, 2)).max() < 1e-8


def test_logistic_generic_kernel_matches_fused(oracle, b200):
    m, conf = configs.logistic(N=2000, p=50, seed=8)
    inits = sm.random_inits(m, 70, 9, spread=0.2)
    cg = compile_on(b200, conf)
    dg = cg._ensure(70)
    dg.set_values(inits)
    y = len(cg.builder.decls) - 1
    fused = dg.calculate(y)
    os.environ["NIMB200_NO_FUSED"] = "1"
    try:
        cg2 = compile_on(b200, conf)
        d2 = cg2._ensure(70)
        d2.set_values(inits)
        generic = d2.calculate(y)
    finally:
        del os.environ["NIMB200_NO_FUSED"]
    co = compile_on(oracle, conf)
    do = co._ensure(70)
    do.set_values(inits)
    ref = do.calculate(y)
    assert rel_err(generic, ref).max() < RTOL_LP
    assert rel_err(fused, ref).max() < RTOL_LP


@pytest.mark.parametrize("N,p,C", [(1, 1, 1), (127, 3, 2), (129, 50, 65), (5000, 50, 130), (4097, 64, 64)])
def test_fused_logistic_block_sum_shapes(oracle, b200, N, p, C):
    m, conf = configs.logistic(N=N, p=max(p, 2), seed=N)
    rng = np.random.default_rng(N)
    inits = np.tile(m.init_vector(), (C, 1))
    bs = m.var_slot["beta"]
    inits[:, bs:bs + max(p, 2)] = rng.standard_normal((C, max(p, 2))) * 1.5
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    dg, do = cg._ensure(C), co._ensure(C)
    dg.set_values(inits)
    do.set_values(inits)
    y = len(cg.builder.decls) - 1
    assert rel_err(dg.calculate(y), do.calculate(y)).max() < RTOL_LP


def test_fused_logistic_extreme_linear_predictors(oracle, b200):
    # |eta| up to ~40: p underflows / q == 0 branches of dbinom_raw must agree, including -Inf
    m, conf = configs.logistic(N=600, p=4, seed=3)
    C = 8
    inits = np.tile(m.init_vector(), (C, 1))
    bs = m.var_slot["beta"]
    inits[:, bs] = np.array([-45, -20, -8, -1, 1, 8, 20, 45.0])
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    dg, do = cg._ensure(C), co._ensure(C)
    dg.set_values(inits)
    do.set_values(inits)
    y = len(cg.builder.decls) - 1
    a, b = dg.calculate(y), do.calculate(y)
    assert np.array_equal(np.isinf(a), np.isinf(b))
    assert rel_err(a, b).max() < 1e-9  # ill-conditioned 1-p for |eta| > 9: see DESIGN.md "conditioning"


def test_known_answers_on_gpu(b200):
    # the reference's own fixtures, on the CUDA path: test-mcmc.R:164-179 and :1379-1443
    m, conf = sm.SMALL["pump"]()
    cg = compile_on(b200, conf)
    s = pooled(run(cg, np.tile(m.init_vector(), (64, 1)), 3000, 500, seed=2))
    mean = dict(zip(cg.colnames, s.mean(0)))
    for name, want, tol in (("theta[1]", 0.06, 0.01), ("theta[2]", 0.10, 0.01), ("theta[9]", 1.58, 0.05),
                            ("theta[10]", 1.97, 0.05), ("alpha", 0.73, 0.1), ("beta", 0.98, 0.1)):
        assert abs(mean[name] - want) < tol, (name, mean[name])
    m, conf = sm.binary_model()
    cg = compile_on(b200, conf)
    s = pooled(run(cg, np.tile(m.init_vector(), (256, 1)), 2000, 0, seed=9))
    for k, v in dict(a=.5, b=.6, c=.05, d=.2, e=.9, f=.9525, g=.0475, h=.5).items():
        assert abs(s[:, cg.colnames.index(k)].mean() - v) < 0.0025, k
    m, conf = sm.binary_out_of_bounds()
    assert np.all(run(compile_on(b200, conf), np.tile(m.init_vector(), (5, 1)), 100, 0) == 1.0)


def test_reset_false_continuation_on_gpu(b200):
    m, conf = sm.SMALL["logistic"]()
    init = sm.random_inits(m, 3, 1, 0.1)
    cg = compile_on(b200, conf)
    a = run(cg, init, 250, 0, seed=4).copy()
    b = run(cg, init, 210, 0, seed=4, reset=False)
    c = run(compile_on(b200, conf), init, 460, 0, seed=4)
    assert np.array_equal(b[:, :250], a) and np.array_equal(b, c)


def test_first_chain_offsets_the_streams(b200):
    # chain sharding: rank r runs chains [r*K, (r+1)*K) with first_chain = r*K and must reproduce them exactly
    m, conf = sm.SMALL["pump"]()
    init = np.tile(m.init_vector(), (6, 1))
    full = run(compile_on(b200, conf), init, 120, 0, seed=8)
    part = nb.compileNimble(nb.buildMCMC(conf), backend=b200, first_chain=4)
    assert np.array_equal(run(part, init[:2], 120, 0, seed=8), full[4:6])


def test_bad_tables_fail_loudly(b200):
    m, conf = sm.SMALL["pump"]()
    cg = compile_on(b200, conf)
    cg.builder.samplers[0].type = 99
    with pytest.raises(Exception, match="no CPU fallback"):
        cg._ensure(2)
    cg = compile_on(b200, conf)
    cg.builder.decls[0].dist = 77
    with pytest.raises(Exception, match="unknown distribution"):
        cg._ensure(2)


def test_full_size_logistic_properties(oracle, b200):
    # BASELINE.json configs[1] at full size (N = 1e6, p = 50): size-independent checks.
    m, conf = configs.logistic()
    C = 128
    inits = np.stack([m.init_vector(i) for i in configs.logistic_inits(50, C)])
    inits[0, m.var_slot["beta"]:m.var_slot["beta"] + 50] = 0.0
    cg = compile_on(b200, conf)
    dg = cg._ensure(C)
    dg.set_values(inits)
    y = len(cg.builder.decls) - 1
    total = dg.calculate(y)
    assert abs(total[0] - 1_000_000 * np.log(0.5)) < 1e-12 * 1_000_000        # beta = 0: every node is log(1/2)
    parts = sum(dg.calculate(y, lo, lo + 250_000) for lo in range(0, 1_000_000, 250_000))  # generic range kernel
    assert rel_err(parts, total).max() < RTOL_LP                                 # additivity across node ranges
    co = compile_on(oracle, conf)
    do = co._ensure(2)
    do.set_values(inits[:2])
    lo, hi = 431_000, 441_000
    d2 = compile_on(b200, conf)._ensure(2)
    d2.set_values(inits[:2])
    assert rel_err(d2.calculate(y, lo, hi), do.calculate(y, lo, hi)).max() < RTOL_LP


def test_fastmath_accuracy(b200, oracle):
    """The fused kernels' own exp / log / reciprocal against libm (numpy), and the fused Bernoulli-logit
    log-density against the oracle's dbinom(y, 1, ilogit(eta)) over the whole range of linear predictors."""
    import ctypes as C
    f = b200.lib.nimb200_debug_math
    f.argtypes = [C.c_int32, C.POINTER(C.c_double), C.c_int32, C.POINTER(C.c_double)]
    dp = C.POINTER(C.c_double)

    def dev(which, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty_like(x)
        assert f(which, x.ctypes.data_as(dp), x.size, out.ctypes.data_as(dp)) == 0
        return out
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-745, 709, 200000), rng.uniform(-40, 40, 200000), rng.normal(0, 1e-3, 1000),
                        [0.0, -0.0, 709.7, -745.1, -760.0, 720.0, 1e-300]])
    with np.errstate(over="ignore", under="ignore"):
        want = np.exp(x)
    got = dev(0, x)
    ok = np.isfinite(want) & (want > 1e-300)
    assert np.max(np.abs(got[ok] - want[ok]) / want[ok]) < 4e-16
    assert np.all(got[want == np.inf] == np.inf) and np.all(got[x < -750] == 0.0)
    t = np.concatenate([rng.uniform(0, 1, 200000), np.exp(rng.uniform(-700, 0, 200000)), 1 - np.exp(rng.uniform(-36, 0, 50000)),
                        [1.0, 0.5, 2.0 ** -1074, 2.0 ** -1022, 0.7071067811865476, 0.7071067811865475]])
    got, want = dev(1, t), np.log(t)
    assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)) < 4e-16
    assert dev(1, np.array([0.0]))[0] == -np.inf and np.isnan(dev(1, np.array([np.nan]))[0])
    d = np.concatenate([1 + np.exp(rng.uniform(-40, 700, 200000)), [1.0, 2.0, 3.0]])
    assert np.max(np.abs(dev(2, d) * d - 1.0)) < 4e-16
    # the fused log-density against the oracle's nmath path, node by node
    eta = np.concatenate([rng.uniform(-8, 8, 100000), rng.uniform(-40, 40, 20000), [-800.0, -745.0, -37.0, 36.7, 37.5, 800.0, 0.0]])
    ref = oracle.lib
    for which, yv in ((3, 0.0), (4, 1.0)):
        got = dev(which, eta)
        with np.errstate(over="ignore"):
            p = 1.0 / (1.0 + np.exp(-eta))
        want = np.array([ref.ref_dbinom(yv, 1.0, float(pp)) for pp in p])
        assert np.array_equal(np.isinf(got), np.isinf(want)), which
        fin = np.isfinite(want)
        # well-conditioned range: 1e-12 relative; beyond |eta| ~ 9 the reference value itself moves by more than that per ulp of exp
        wc = fin & (np.abs(eta) <= 8)
        assert np.max(np.abs(got[wc] - want[wc]) / np.maximum(np.abs(want[wc]), 1e-300)) < 1e-12
        assert np.max(np.abs(got[fin] - want[fin])) < 2e-2      # noise level of log(1 - p) itself near p = 1 - 1e-13
        assert np.isnan(dev(which, np.array([np.nan]))[0])


@pytest.mark.parametrize("name", ["logistic_big", "lmm", "glmm", "occupancy"])
def test_data_shards_add_up_to_the_whole(b200, name):
    """Data-node sharding, arithmetic only (no communicator): the block sums of the R per-rank models add up to the
    unsharded block sum, for the fused kernel (tile ranges) and the generic kernel (node ranges)."""
    m, conf = (configs.logistic(N=5000, p=50, seed=6) if name == "logistic_big" else sm.SMALL[name]())
    C, R = 6, 3
    inits = sm.random_inits(m, C, 11, 0.1)
    whole = compile_on(b200, conf)._ensure(C)
    whole.set_values(inits)
    for di, d in enumerate(whole.builder.decls):
        if d.value_kind != 0:
            continue
        total = whole.calculate(di)
        parts = np.zeros(C)
        covered = 0
        for r in range(R):
            dm = compile_on(b200, conf)._ensure(C)
            dm.enable_data_sharding(R, r, None, min_nodes=256)
            dm.set_values(inits)
            parts += dm.calculate(di)
        assert rel_err(parts, total).max() < RTOL_LP, (name, di)


def test_pois_family_block_sums_and_edges(oracle, b200):
    """Fused Poisson log-linear family (C >= 32 chains) against the oracle, including chains that must take the exact
    per-element path: lam = 0, lam < 0, overflowing and underflowing linear predictors."""
    m, conf = sm.SMALL["glmm"]()
    C = 70
    inits = sm.random_inits(m, C, 5, 0.3)
    lam0, b0s = m.var_slot["lam"], m.var_slot["b0"]
    inits[3, lam0 + 2] = 0.0
    inits[4, lam0 + 5] = -1.0
    inits[5, b0s] = 800.0
    inits[6, b0s] = -800.0
    inits[7, b0s] = np.nan
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    dg, do = cg._ensure(C), co._ensure(C)
    dg.set_values(inits)
    do.set_values(inits)
    y = len(cg.builder.decls) - 1
    assert cg.builder.decls[y].family == T.FAM_POIS_LOGLIN
    a, b = dg.calculate(y), do.calculate(y)
    assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b))
    assert rel_err(a, b).max() < RTOL_LP, rel_err(a, b).max()
    assert rel_err(dg.calculate(y, 256, 3072), do.calculate(y, 256, 3072)).max() < RTOL_LP


def test_pois_moment_series_and_node_pass_agree(oracle, b200, monkeypatch):
    """Group sums S_g(b1) = sum exp(b1 x): the data-only moment series (|b1| max|x| <= 2) and the node pass give the same block
    sums as the oracle; chains on both sides of the switch share warps, some far outside it."""
    m, conf = sm.SMALL["glmm"]()
    C = 96
    inits = sm.random_inits(m, C, 9, 0.2)
    b1s = m.var_slot["b1"]
    inits[:, b1s] = np.linspace(-3.0, 3.0, C)          # max|x| ~ 3 in the small model: |u| runs from 0 to ~9
    inits[5, b1s], inits[40, b1s], inits[41, b1s] = 150.0, -400.0, 0.0
    co = compile_on(oracle, conf)
    do = co._ensure(C)
    do.set_values(inits)
    y = len(co.builder.decls) - 1
    want = do.calculate(y)
    got = {}
    for label, env in (("series", None), ("nodes", "1")):
        if env:
            monkeypatch.setenv("NIMB200_NO_SERIES", env)
        dg = compile_on(b200, conf)._ensure(C)
        dg.set_values(inits)
        got[label] = dg.calculate(y)
        assert np.array_equal(np.isnan(got[label]), np.isnan(want)) and np.array_equal(np.isinf(got[label]), np.isinf(want)), label
        assert rel_err(got[label], want).max() < RTOL_LP, (label, rel_err(got[label], want).max())
    assert rel_err(got["series"], got["nodes"]).max() < 1e-13


def test_gamma_layer_statistics_path(oracle, b200):
    """An i.i.d. gamma layer with chain-level parameters (>= 256 nodes: lam[g] ~ dgamma(a, b)) is evaluated from (sum log lam,
    sum lam): block sums, the conjugate update of the rate and whole trajectories agree with the oracle's node-by-node path,
    including chains whose values or parameters fall outside the closed form (zero / negative / infinite lam, shape > 64)."""
    m, conf = configs.glmm(N=3000, G=300, seed=12)
    C = 40
    inits = sm.random_inits(m, C, 3, 0.2)
    lam0, a_s, b_s = m.var_slot["lam"], m.var_slot["a"], m.var_slot["b"]
    inits[2, lam0 + 7] = 0.0
    inits[3, lam0 + 299] = -0.5
    inits[4, lam0] = np.inf
    inits[5, a_s] = 70.0
    inits[6, a_s] = -1.0
    inits[7, b_s] = 0.0
    inits[8, lam0 + 100] = np.nan
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    dg, do = cg._ensure(C), co._ensure(C)
    dg.set_values(inits)
    do.set_values(inits)
    ld = [i for i, d in enumerate(cg.builder.decls) if d.dist == T.DGAMMA and d.n == 300]
    assert len(ld) == 1
    for _ in range(2):   # the second call is served from the cached statistics
        a, b = dg.calculate(ld[0]), do.calculate(ld[0])
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b)), (a[:10], b[:10])
        assert rel_err(a, b).max() < RTOL_LP, rel_err(a, b).max()
    good = sm.random_inits(m, C, 8, 0.1)
    so, sg = run(co, good, 150, 10, seed=5), run(cg, good, 150, 10, seed=5)
    assert rel_err(sg, so).max() < RTOL_TRAJ
    assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < RTOL_TRAJ


def test_pois_family_trajectories(oracle, b200):
    m, conf = sm.SMALL["glmm"]()
    inits = sm.random_inits(m, 40, 8, 0.1)
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, inits, 260, 10, seed=3), run(cg, inits, 260, 10, seed=3)
    assert rel_err(sg, so).max() < RTOL_TRAJ
    assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < RTOL_TRAJ


def test_norm_family_block_sums_and_edges(oracle, b200):
    """Streaming normal family against the oracle, including chains with sd <= 0 / Inf / NaN and non-finite means."""
    for name in ("lmm", "lmm_perm"):
        m, conf = sm.SMALL[name]()
        C = 13
        inits = sm.random_inits(m, C, 5, 0.3)
        se, mu = m.var_slot["sig_e"], m.var_slot["mu"]
        inits[3, se] = 0.0
        inits[4, se] = -1.0
        inits[5, se] = np.inf
        inits[6, se] = np.nan
        inits[7, mu] = np.inf
        inits[8, mu] = np.nan
        cg, co = compile_on(b200, conf), compile_on(oracle, conf)
        dg, do = cg._ensure(C), co._ensure(C)
        dg.set_values(inits)
        do.set_values(inits)
        y = len(cg.builder.decls) - 1
        assert cg.builder.decls[y].family == T.FAM_NORM_GATHER
        a, b = dg.calculate(y), do.calculate(y)
        assert np.array_equal(np.isnan(a), np.isnan(b)), (a, b)
        assert np.array_equal(np.isinf(a), np.isinf(b)) and np.array_equal(np.sign(a[np.isinf(a)]), np.sign(b[np.isinf(b)]))
        assert rel_err(a, b).max() < RTOL_LP, rel_err(a, b).max()
        assert rel_err(dg.calculate(y, 100, 4211), do.calculate(y, 100, 4211)).max() < RTOL_LP
    # sd == 0 with every residual exactly zero -> +Inf, as dnorm(x, x, 0)
    m = nb.nimbleModel("{ s ~ dgamma(1,1)\n a ~ dnorm(0, 1)\n for (i in 1:5) { y[i] ~ dnorm(a, sd = s) } }", data={"y": [2.0] * 5},
                       inits={"s": 0.0, "a": 2.0})
    conf = nb.configureMCMC(m, useConjugacy=False)
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    dg, do = cg._ensure(2), co._ensure(2)
    v = np.array([[0.0, 2.0], [0.0, 2.5]])
    dg.set_values(v)
    do.set_values(v)
    assert np.array_equal(dg.calculate(2), do.calculate(2)) and dg.calculate(2)[0] == np.inf and dg.calculate(2)[1] == -np.inf


def test_norm_family_group_statistics_path(oracle, b200):
    """After the first sweep over the gathered vector the device keeps (sum y, n, centred sum of squares) per group and forms
    whole-declaration block sums from them, G terms per chain instead of a pass over the nodes: same values as the oracle's
    node-by-node sum, also for sd <= 0 / Inf / NaN, non-finite means, permuted group indices and groups without nodes."""
    for name in ("lmm", "lmm_perm"):
        m, conf = sm.SMALL[name]()
        C = 13
        inits = sm.random_inits(m, C, 5, 0.3)
        cg, co = compile_on(b200, conf), compile_on(oracle, conf)
        run(cg, inits, 5, 0, seed=3)          # the u sweep has run: group statistics are in place
        state = cg.dm.get_values()
        se, mu = m.var_slot["sig_e"], m.var_slot["mu"]
        state[3, se], state[4, se], state[5, se], state[6, se] = 0.0, -1.0, np.inf, np.nan
        state[7, mu], state[8, mu] = np.inf, np.nan
        state[9, m.var_slot["u"] + 2] = 1e6
        cg.dm.set_values(state)
        do = co._ensure(C)
        do.set_values(state)
        y = len(cg.builder.decls) - 1
        a, b = cg.dm.calculate(y), do.calculate(y)
        assert np.array_equal(np.isnan(a), np.isnan(b)), (a, b)
        assert np.array_equal(np.isinf(a), np.isinf(b)) and np.array_equal(np.sign(a[np.isinf(a)]), np.sign(b[np.isinf(b)]))
        assert rel_err(a, b).max() < RTOL_LP, rel_err(a, b).max()
        # and the streaming kernel on the same state (a node range never takes the group path)
        n = cg.builder.decls[y].n
        parts = cg.dm.calculate(y, 0, n // 2) + cg.dm.calculate(y, n // 2, n)
        ok = np.isfinite(b)
        assert rel_err(parts[ok], a[ok]).max() < RTOL_LP


def test_bern_family_block_sums_and_edges(oracle, b200):
    """Bernoulli-logit-linear family (occupancy model) against the oracle for >= 32 chains: the data declaration with its
    latent gate, the latent declaration itself, and chains that must take the exact path (gate / value not 0 or 1, NaN)."""
    m, conf = sm.SMALL["occupancy"]()
    C = 40
    inits = sm.random_inits(m, C, 4, 0.3)
    z0 = m.var_slot["z"]
    rng = np.random.default_rng(1)
    inits[:, z0:z0 + 300] = (rng.random((C, 300)) < 0.6).astype(float)
    inits[3, z0 + 5] = 0.5      # gate / value that is neither 0 nor 1
    inits[4, m.var_slot["a0"]] = 60.0
    inits[5, m.var_slot["a0"]] = -800.0
    inits[6, m.var_slot["b1"]] = np.nan
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    dg, do = cg._ensure(C), co._ensure(C)
    dg.set_values(inits)
    do.set_values(inits)
    fams = [d.family for d in cg.builder.decls]
    assert fams[-2:] == [T.FAM_BERN_LOGIT_LIN, T.FAM_BERN_LOGIT_LIN]
    for di in (len(fams) - 2, len(fams) - 1):
        a, b = dg.calculate(di), do.calculate(di)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b)), di
        assert rel_err(a, b).max() < RTOL_LP, (di, rel_err(a, b).max())


def test_bern_data_kernel_agrees_with_the_general_kernel(oracle, b200, monkeypatch):
    """The data-declaration block-sum kernel (value checks per tile, predictor checks by the moderate path's vote) against the
    general kernel and the oracle, with odd gates and wild predictors -- and once more on data that hold a value that is neither
    0 nor 1 (its tile goes through the scalar form; every chain's sum is then -Inf, as dbinom(0.5, 1, p) is)."""
    from nimble_b200.configs import OCCUPANCY_CODE
    for odd in (False, True):
        rng = np.random.default_rng(5)
        S, J = 700, 4
        x, w = rng.standard_normal(S), rng.standard_normal((S, J))
        y = (rng.random((S, J)) < 0.4).astype(np.float64)
        if odd:
            y[123, 2] = 0.5
        m = nb.nimbleModel(OCCUPANCY_CODE, constants={"S": S, "J": J, "x": x, "w": w}, data={"y": y},
                           inits={"b0": 0.0, "b1": 0.0, "a0": 0.0, "a1": 0.0, "z": np.ones(S)}, name="occ_odd")
        conf = nb.configureMCMC(m, monitors=["a0"])
        C = 40
        inits = sm.random_inits(m, C, 4, 0.3)
        z0 = m.var_slot["z"]
        inits[:, z0:z0 + S] = np.maximum((rng.random((C, S)) < 0.6).astype(float), y.max(axis=1) >= 1.0)   # z = 1 where detected
        inits[3, z0 + 5] = 0.5
        inits[4, m.var_slot["a0"]] = 60.0
        inits[5, m.var_slot["a0"]] = -800.0
        inits[6, m.var_slot["a1"]] = np.nan
        inits[7, m.var_slot["a0"]] = np.inf
        inits[8, z0 + 9] = 0.0                                                # may contradict a detection: -Inf for that chain
        co = compile_on(oracle, conf)
        do = co._ensure(C)
        do.set_values(inits)
        yd = len(co.builder.decls) - 1
        assert co.builder.decls[yd].family == T.FAM_BERN_LOGIT_LIN and co.builder.decls[yd].value_kind == 0
        want = do.calculate(yd)
        if not odd:
            assert np.isfinite(want).sum() > C // 2
        got = {}
        for label, env in (("data", None), ("general", "1")):
            if env:
                monkeypatch.setenv("NIMB200_BERN_GENERAL", env)
            else:
                monkeypatch.delenv("NIMB200_BERN_GENERAL", raising=False)
            dg = compile_on(b200, conf)._ensure(C)
            dg.set_values(inits)
            got[label] = dg.calculate(yd)
            assert np.array_equal(np.isnan(got[label]), np.isnan(want)), (odd, label)
            assert np.array_equal(np.isinf(got[label]), np.isinf(want)), (odd, label)
            assert rel_err(got[label], want).max() < RTOL_LP, (odd, label, rel_err(got[label], want).max())
        assert rel_err(got["data"], got["general"]).max() < 1e-13


def test_bern_family_trajectories(oracle, b200):
    m, conf = sm.SMALL["occupancy"]()
    inits = sm.random_inits(m, 40, 8, 0.1)
    inits[:, m.var_slot["z"]:m.var_slot["z"] + 300] = 1.0
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    so, sg = run(co, inits, 260, 10, seed=3), run(cg, inits, 260, 10, seed=3)
    assert rel_err(sg, so).max() < RTOL_TRAJ
    assert np.array_equal(cg.dm.get_values()[:, m.var_slot["z"]:], co.dm.get_values()[:, m.var_slot["z"]:])


def test_posteriors_agree_within_monte_carlo_error(oracle, b200):
    """north_star, third criterion: independent GPU and CPU runs (different seeds, different chains) give posterior means and
    variances that agree within Monte Carlo standard error (ESS by the coda-style estimator applied to both sides)."""
    m, conf = sm.SMALL["logistic"]()
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    sg = run(cg, sm.random_inits(m, 64, 1, 0.1), 3000, 1000, seed=101)
    so = run(co, sm.random_inits(m, 8, 2, 0.1), 3000, 1000, seed=202)
    for j, name in enumerate(cg.colnames):
        eg = sum(nb.effective_size(sg[c, :, j]) for c in range(sg.shape[0]))
        eo = sum(nb.effective_size(so[c, :, j]) for c in range(so.shape[0]))
        vg, vo = sg[:, :, j].var(), so[:, :, j].var()
        mcse = np.sqrt(vg / max(eg, 1.0) + vo / max(eo, 1.0))
        assert abs(sg[:, :, j].mean() - so[:, :, j].mean()) < 5 * mcse, (name, sg[:, :, j].mean(), so[:, :, j].mean(), mcse)
        assert abs(np.log(vg / vo)) < 0.6, (name, vg, vo)
