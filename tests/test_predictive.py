"""Posterior-predictive networks: assignment of sampler_posterior_predictive to the root of an entirely non-data network, which
then simulates everything downstream (P/R/MCMC_configuration.R:247-287, 368; P/R/MCMC_samplers.R:86-105), and the reference's
default of leaving predictive nodes out of the other samplers' calculations (P/R/options.R:190, MCMC_build.R:156-176).
Fixtures: P/tests/testthat/test-mcmc.R:2341-2364 (assignment), :2367-2392 (calcNodes of the root's sampler)."""
import numpy as np
import pytest

import nimble_b200 as nb
from nimble_b200 import table as T
from tests.helpers import compile_on, pooled, rel_err, run

ABC = "{ a ~ dnorm(0, 1)\n y ~ dnorm(a, 1)\n b[1] ~ dnorm(a, 1)\n b[2] ~ dnorm(a, 1)\n c ~ dnorm(b[1] + b[2], 1) }"
INITS = {"a": 0.0, "b": [0.0, 0.0], "c": 0.0}


def _names(conf):
    return [(s.name, s.var) for s in conf.samplerConfs]


def test_assignment_without_data_everything_is_predictive():
    # test-mcmc.R:2341-2350: no data at all -> one posterior_predictive sampler on the root `a`
    m = nb.nimbleModel(ABC.replace("y ~ dnorm(a, 1)\n", ""), inits=INITS)
    conf = nb.configureMCMC(m)
    assert _names(conf) == [("posterior_predictive", "a")]
    assert [m.decls[i].var for i in conf.samplerConfs[0].network] == ["b", "b", "c"]


def test_assignment_with_data():
    # test-mcmc.R:2351-2355: y observed -> a conjugate, b[1] and b[2] roots of predictive networks that share c
    m = nb.nimbleModel(ABC, data={"y": 1.0}, inits=INITS)
    conf = nb.configureMCMC(m)
    assert _names(conf) == [("conjugate_dnorm", "a"), ("posterior_predictive", "b"), ("posterior_predictive", "b")]
    assert all([m.decls[i].var for i in sc.network] == ["c"] for sc in conf.samplerConfs[1:])
    # a's conjugate update sees y only: predictive b[1], b[2] are left out of its dependents
    low = nb.compileNimble(nb.buildMCMC(conf), backend=_NoBackend()).lowering
    s0 = low.b.samplers[0]
    deps = [low.b.deps[s0.dep_off + k] for k in range(s0.n_dep)]
    assert [model_var(m, low, d.decl) for d in deps if d.stage == 1] == ["y"]


def model_var(m, low, table_decl):
    inv = {v: k for k, v in low.decl_ids.items()}
    return m.decls[inv[table_decl]].var


class _NoBackend:
    """compileNimble only lowers; no device is touched until a run starts."""
    name = "none"


def test_root_sampler_covers_its_network_and_parent_ignores_it():
    # test-mcmc.R:2367-2392: a RW (y ~ dexp(a^2+1) is not conjugate), b posterior_predictive with calcNodes (b, c); c gets nothing
    code = "{ a ~ dnorm(0, 1)\n y ~ dexp(a^2 + 1)\n b ~ dnorm(a, 1)\n c ~ dnorm(b, 1) }"
    m = nb.nimbleModel(code, data={"y": 3.0}, inits={"a": 0.0, "b": 1.0, "c": 1.0})
    conf = nb.configureMCMC(m)
    assert _names(conf) == [("RW", "a"), ("posterior_predictive", "b")]
    assert [m.decls[i].var for i in conf.samplerConfs[1].network] == ["c"]
    with pytest.raises(ValueError, match="non-posterior-predictive"):
        conf.addSampler("a", type="posterior_predictive")


def _network_model():
    code = ("{ mu ~ dnorm(0, 1)\n for(i in 1:n) { y[i] ~ dnorm(mu, 1) }\n a ~ dnorm(mu, 1)\n b ~ dnorm(2*a, sd = 0.5)\n"
            "for(k in 1:3) { c[k] ~ dnorm(a + b, 1) } }")
    y = np.array([0.3, 1.1, 0.8, 1.6, 0.2])
    m = nb.nimbleModel(code, constants={"n": 5}, data={"y": y}, inits={"mu": 0.0, "a": 0.0, "b": 0.0, "c": np.zeros(3)})
    return m, nb.configureMCMC(m, monitors=["mu", "a", "b", "c"]), y


def test_predictive_network_moments_on_oracle(oracle):
    m, conf, y = _network_model()
    assert _names(conf) == [("conjugate_dnorm", "mu"), ("posterior_predictive", "a")]
    cm = compile_on(oracle, conf)
    s = pooled(run(cm, np.tile(m.init_vector(), (4, 1)), 20000, 500, seed=17))
    col = lambda n_: s[:, cm.colnames.index(n_)]
    pm, pv = y.sum() / 6.0, 1.0 / 6.0                # mu | y, untouched by the predictive nodes
    assert abs(col("mu").mean() - pm) < 0.02 and abs(col("mu").var() - pv) < 0.01
    assert abs(col("a").mean() - pm) < 0.03 and abs(col("a").var() - (pv + 1)) < 0.05
    assert abs(col("b").mean() - 2 * pm) < 0.06 and abs(col("b").var() - (4 * (pv + 1) + 0.25)) < 0.25
    assert abs(col("c[2]").mean() - 3 * pm) < 0.09 and abs(col("c[2]").var() - (9 * (pv + 1) + 0.25 + 1)) < 0.6
    # each c[k] is drawn from the a and b of the SAME iteration (the network is simulated parents first)
    assert abs(np.var(col("c[1]") - col("a") - col("b")) - 1.0) < 0.05


@pytest.mark.gpu
def test_predictive_network_matches_oracle(b200, oracle):
    m, conf, _ = _network_model()
    inits = np.tile(m.init_vector(), (9, 1))
    sg, so = run(compile_on(b200, conf), inits, 300, 20, seed=4), run(compile_on(oracle, conf), inits, 300, 20, seed=4)
    assert rel_err(sg, so).max() < 1e-9
