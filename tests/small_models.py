"""Small models used by the parity tests: the reference's own known-answer MCMC fixtures
(P/tests/testthat/test-mcmc.R) plus scaled-down versions of the BASELINE.json configurations."""
import numpy as np

import nimble_b200 as nb
from nimble_b200 import configs


def very_simple():
    """test-mcmc.R:209-227: E x = 6/5, sd x = 1/sqrt(5), E z = 5, sd z = 1/2."""
    code = "{ x ~ dnorm(0, 2)\n y ~ dnorm(x+1, 3)\n z ~ dnorm(y+2, 4) }"
    m = nb.nimbleModel(code, data={"y": 3.0}, inits={"x": 0.0, "z": 0.0})
    return m, nb.configureMCMC(m, monitors=["x", "z"])


def basic_block(kind):
    """test-mcmc.R:284-332: x[i] ~ dnorm(0,1), y[i] ~ dnorm(x[i],2), y = -1:1 -> mean (-2/3,0,2/3), var 1/3."""
    code = "{ for(i in 1:3) { x[i] ~ dnorm(0, 1)\n y[i] ~ dnorm(x[i], 2) } }"
    m = nb.nimbleModel(code, data={"y": [-1.0, 0.0, 1.0]}, inits={"x": [0.0, 0.0, 0.0]})
    conf = nb.configureMCMC(m, monitors=["x"])
    if kind == "block_scalars":
        conf.removeSamplers()
        for i in (1, 2, 3):
            conf.addSampler("x[%d]" % i, type="RW_block")
    elif kind == "block_vector":
        conf.removeSamplers()
        conf.addSampler("x", type="RW_block", control={"adaptInterval": 500})
    elif kind == "rw":
        conf.removeSamplers()
        for i in (1, 2, 3):
            conf.addSampler("x[%d]" % i, type="RW")
    return m, conf


def binary_model():
    """test-mcmc.R:1379-1443: means a=.5 b=.6 c=.05 d=.2 e=.9 f=.9525 g=.0475 h=.5 (tol .0025 at 1e5 its)."""
    code = """{
        a ~ dbern(0.5)
        b ~ dbern(0.6)
        c ~ dbern(0.05)
        d ~ dbin(prob=0.2, size=1)
        e ~ dbinom(prob=0.9, size=1)
        f ~ dbern(0.5)
        g ~ dbern(0.5)
        h ~ dbern(0.5)
        for(i in 1:10) { yf[i] ~ dnorm(f, sd = 1) }
        for(i in 1:10) { yg[i] ~ dnorm(g, sd = 1) }
        for(i in 1:10) { yh[i] ~ dnorm(h, sd = 1) }
    }"""
    data = {"yf": [0] * 2 + [1] * 8, "yg": [0] * 8 + [1] * 2, "yh": [0] * 5 + [1] * 5}
    inits = {k: 0.0 for k in "abcdefgh"}
    m = nb.nimbleModel(code, data=data, inits=inits)
    conf = nb.configureMCMC(m, nodes=[], monitors=list("abcdefgh"))
    for k in "abcdefgh":
        conf.addSampler(k, type="binary")
    return m, conf


def binary_out_of_bounds():
    """test-mcmc.R:1445-1469: every sample stays at 1."""
    code = "{ px ~ dbern(0.5)\n py ~ dbern(0.5)\n x ~ dnorm(0, sd = px - 0.5)\n y ~ dnorm(0, tau = py) }"
    m = nb.nimbleModel(code, data={"x": 0.0, "y": 0.0}, inits={"px": 1.0, "py": 1.0})
    return m, nb.configureMCMC(m)


def all_dists():
    """One declaration per supported distribution, latent parameters so every chain differs."""
    code = """{
      mu ~ dnorm(0, sd = 3)
      s ~ dgamma(2, 1)
      r ~ dgamma(3, rate = 2)
      pp ~ dbeta(2.5, 3.5)
      lo ~ dlogis(0, 1)
      ex ~ dexp(2)
      for (i in 1:n) {
        yn[i] ~ dnorm(mu + lo, sd = s)
        yg[i] ~ dgamma(r, rate = s)
        yp[i] ~ dpois(r * e[i])
        yb[i] ~ dbin(pp, m[i])
        ybe[i] ~ dbeta(r, s + 1)
        yl[i] ~ dlogis(mu, scale = s)
        ye[i] ~ dexp(ex + r)
        yc[i] ~ dcat(w[1:4])
        yt[i] ~ dnorm(mu, tau = r)
        yv[i] ~ dnorm(mu, var = s)
      }
    }"""
    rng = np.random.default_rng(5)
    n = 37
    data = {"yn": rng.normal(0, 2, n), "yg": rng.gamma(2, 1, n), "yp": rng.poisson(3, n).astype(float),
            "yb": rng.binomial(12, 0.4, n).astype(float), "ybe": rng.beta(2, 3, n), "yl": rng.logistic(0, 1, n),
            "ye": rng.exponential(1, n), "yc": rng.integers(1, 5, n).astype(float), "yt": rng.normal(0, 1, n),
            "yv": rng.normal(0, 1, n)}
    consts = {"n": n, "e": rng.gamma(2, 1, n), "m": np.full(n, 12.0), "w": [0.1, 0.4, 0.2, 0.9]}
    inits = {"mu": 0.3, "s": 1.2, "r": 1.7, "pp": 0.45, "lo": -0.2, "ex": 0.8}
    m = nb.nimbleModel(code, constants=consts, data=data, inits=inits)
    return m, nb.configureMCMC(m, useConjugacy=False)


def beta_binomial():
    """dbeta prior with dbern and dbin dependents (identity link), MCMC_conjugacy.R:4-15; test-mcmc.R:400-447 uses the same
    relationship.  Posterior p ~ Beta(2 + sum y + sum z, 3 + sum(n - y) + sum(1 - z)): every iteration is an exact draw."""
    rng = np.random.default_rng(12)
    n = rng.integers(1, 9, 7).astype(float)
    y = rng.binomial(n.astype(int), 0.3).astype(float)
    z = (rng.random(5) < 0.3).astype(float)
    code = """{ p ~ dbeta(2, 3)
      for (i in 1:7) { y[i] ~ dbin(p, n[i]) }
      for (j in 1:5) { z[j] ~ dbern(p) } }"""
    m = nb.nimbleModel(code, constants={"n": n}, data={"y": y, "z": z}, inits={"p": 0.5})
    post = (2 + y.sum() + z.sum(), 3 + (n - y).sum() + (1 - z).sum())
    return m, nb.configureMCMC(m), post


def gamma_exponential():
    """dgamma prior on the rate of dexp dependents, multiplicative link (MCMC_conjugacy.R:89): posterior
    lam ~ Gamma(2 + n, rate 1 + sum(c_i t_i))."""
    rng = np.random.default_rng(13)
    c = 0.5 + rng.random(9)
    t = rng.exponential(0.7, 9)
    code = "{ lam ~ dgamma(2, 1)\n for (i in 1:9) { t[i] ~ dexp(c[i] * lam) } }"
    m = nb.nimbleModel(code, constants={"c": c}, data={"t": t}, inits={"lam": 1.0})
    return m, nb.configureMCMC(m), (2 + 9, 1 + float((c * t).sum()))


def predictive_nodes():
    """One terminal (posterior predictive) node per distribution the device can simulate, all driven by one latent a."""
    code = """{ a ~ dgamma(2, 1)
      y ~ dnorm(a, 1)
      pn ~ dnorm(a, sd = 2)
      pg ~ dgamma(a, 2)
      pb ~ dbeta(2, a)
      pe ~ dexp(a)
      pl ~ dlogis(a, 2)
      pz ~ dbern(0.3)
      pp ~ dpois(3 * a)
      pk ~ dbin(0.3, 12) }"""
    m = nb.nimbleModel(code, data={"y": 1.5}, inits={"a": 1.0, "pn": 0.0, "pg": 1.0, "pb": 0.5, "pe": 1.0, "pl": 0.0, "pz": 0.0,
                                                     "pp": 1.0, "pk": 3.0})
    return m, nb.configureMCMC(m, monitors=["a", "pn", "pg", "pb", "pe", "pl", "pz", "pp", "pk"])


def slice_targets():
    """A discrete and a continuous node with known marginals under the slice sampler: x ~ Poisson(3.5), g ~ Gamma(3, rate 2)."""
    m = nb.nimbleModel("{ x ~ dpois(3.5)\n g ~ dgamma(3, 2) }", inits={"x": 3.0, "g": 1.0})
    conf = nb.configureMCMC(m, nodes=[])
    conf.addSampler("x", type="slice")
    conf.addSampler("g", type="slice")
    return m, conf


def latent_counts():
    """N-mixture style model: latent Poisson counts (slice, the default for non-binary discrete nodes, MCMC_configuration.R:428)
    under a binomial observation; gamma-Poisson and beta-binomial conjugacy on the latent layer."""
    rng = np.random.default_rng(3)
    Nt = rng.poisson(6.0, 12)
    y = rng.binomial(Nt, 0.4).astype(float)
    code = """{ lam ~ dgamma(2, 0.2)
      p ~ dbeta(2, 2)
      for (i in 1:12) { N[i] ~ dpois(lam)
        y[i] ~ dbin(p, N[i]) } }"""
    m = nb.nimbleModel(code, data={"y": y}, inits={"lam": 5.0, "p": 0.5, "N": y + 2})
    return m, nb.configureMCMC(m, monitors=["lam", "p", "N"])


def gaussian_mixture():
    """Finite mixture with latent labels: z[i] ~ dcat (categorical sampler, MCMC_configuration.R:425), dynamic indexing mu[z[i]]
    into a latent vector and s[z[i]] into a constant one, conjugate normal updates of mu[k] whose dependents change with z."""
    rng = np.random.default_rng(1)
    zt = rng.integers(1, 4, 60)
    mu_t = np.array([-4.0, 0.0, 5.0])
    y = mu_t[zt - 1] + 0.7 * rng.standard_normal(60)
    code = """{ for (k in 1:3) { mu[k] ~ dnorm(0, sd = 10) }
      for (i in 1:60) { z[i] ~ dcat(w[1:3])
        y[i] ~ dnorm(mu[z[i]], sd = s[z[i]]) } }"""
    m = nb.nimbleModel(code, constants={"w": [0.3, 0.3, 0.4], "s": [0.7, 0.6, 0.8]}, data={"y": y},
                       inits={"mu": [-1.0, 0.0, 1.0], "z": np.ones(60)})
    return m, nb.configureMCMC(m, monitors=["mu", "z"]), (mu_t, zt)


def random_inits(model, nchains, seed, spread=0.3):
    """Positive, chain-specific perturbations of the model's inits (works for the models above)."""
    rng = np.random.default_rng(seed)
    base = model.init_vector()
    out = np.tile(base, (nchains, 1))
    out *= np.exp(spread * rng.standard_normal(out.shape))
    out += (base == 0) * spread * rng.standard_normal(out.shape)
    return out


SMALL = {
    "pump": lambda: configs.pump(),
    "logistic": lambda: configs.logistic(N=3000, p=7, seed=1),
    "lmm": lambda: configs.lmm(N=6000, G=30, seed=2),
    "lmm_perm": lambda: configs.lmm(N=6000, G=30, seed=2, permute=True),
    "occupancy": lambda: configs.occupancy(S=300, J=4, seed=3),
    "glmm": lambda: configs.glmm(N=4000, G=40, seed=4),
}
