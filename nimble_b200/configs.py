"""The five BASELINE.json configurations as BUGS models on synthetic data.

Every generator draws with numpy `default_rng(seed)` in the order listed in SURVEY.md 8d; sizes can be scaled
down for tests.  The BUGS text is what a NIMBLE user would write, so the same models run in the reference."""
import numpy as np

from .mcmc import buildMCMC, configureMCMC
from .model import nimbleModel

PUMP_CODE = """
model {
  for (i in 1:N){
      theta[i] ~ dgamma(alpha,beta);
      lambda[i] <- theta[i]*t[i];
      x[i] ~ dpois(lambda[i])
  }
  alpha ~ dexp(1.0);
  beta ~ dgamma(0.1,1.0);
}
"""
# P/inst/classic-bugs/vol1/pump/pump-data.R, pump-init.R
PUMP_CONSTANTS = {"N": 10, "t": [94.3, 15.7, 62.9, 126, 5.24, 31.4, 1.05, 1.05, 2.1, 10.5]}
PUMP_DATA = {"x": [5, 1, 5, 14, 3, 19, 1, 1, 4, 22]}
PUMP_INITS = {"alpha": 1.0, "beta": 1.0, "theta": [0.1] * 10}


def pump():
    """C1: BUGS 'pump' (P/inst/classic-bugs/vol1/pump/pump.bug), default samplers."""
    m = nimbleModel(PUMP_CODE, constants=PUMP_CONSTANTS, data=PUMP_DATA, inits=PUMP_INITS, name="pump")
    conf = configureMCMC(m)
    conf.addMonitors("theta")
    return m, conf


LOGISTIC_CODE = """
{
  tau_b ~ dgamma(2, 1)
  mu_b ~ dnorm(0, sd = 10)
  for (j in 1:p) { beta[j] ~ dnorm(mu_b, tau = tau_b) }
  for (i in 1:N) {
    logit(pr[i]) <- inprod(X[i, 1:p], beta[1:p])
    y[i] ~ dbern(pr[i])
  }
}
"""


def logistic(N=1_000_000, p=50, seed=1001):
    """C2: hierarchical logistic regression, RW_block on beta[1:p], RW(log) on tau_b, conjugate mu_b."""
    rng = np.random.default_rng(seed)
    X = np.empty((N, p))
    X[:, 0] = 1.0
    X[:, 1:] = rng.standard_normal((N, p - 1)) / np.sqrt(p)
    beta_true = rng.standard_normal(p)
    eta = X @ beta_true
    y = (rng.random(N) < 1.0 / (1.0 + np.exp(-eta))).astype(np.float64)
    m = nimbleModel(LOGISTIC_CODE, constants={"N": N, "p": p, "X": X}, data={"y": y},
                    inits={"tau_b": 1.0, "mu_b": 0.0, "beta": np.zeros(p)}, name="logistic")
    m.truth = {"beta": beta_true}
    conf = configureMCMC(m, monitors=["tau_b", "mu_b", "beta"])
    conf.removeSamplers("beta")
    conf.removeSamplers("tau_b")
    conf.addSampler("tau_b", type="RW", control={"log": True})
    conf.addSampler("beta[1:%d]" % p, type="RW_block", control={"adaptInterval": 200, "scale": 1.0})
    # execution order as configureMCMC would list them: top-level nodes first, then the block
    order = {"tau_b": 0, "mu_b": 1, "beta": 2}
    conf.samplerConfs.sort(key=lambda s: order[s.var])
    return m, conf


def logistic_inits(p, nchains, seed=1001):
    """Per-chain inits beta ~ N(0, 0.1^2)."""
    rng = np.random.default_rng(seed + 7919)
    return [{"tau_b": 1.0, "mu_b": 0.0, "beta": 0.1 * rng.standard_normal(p)} for _ in range(nchains)]


LMM_CODE = """
{
  mu ~ dnorm(0, sd = 100)
  sig_u ~ dgamma(2, 1)
  sig_e ~ dgamma(2, 1)
  for (g in 1:G) { u[g] ~ dnorm(0, sd = sig_u) }
  for (i in 1:N) { y[i] ~ dnorm(mu + u[grp[i]], sd = sig_e) }
}
"""


def lmm(N=10_000_000, G=10_000, seed=1003, permute=False):
    """C3: linear mixed model, conjugate u[g] and mu, RW(log) on sig_u / sig_e."""
    rng = np.random.default_rng(seed)
    grp = np.repeat(np.arange(1, G + 1), N // G)
    if grp.size < N:
        grp = np.concatenate([grp, np.full(N - grp.size, G)])
    if permute:
        grp = rng.permutation(grp)
    u_true = rng.standard_normal(G)
    y = 2.0 + u_true[grp - 1] + 1.5 * rng.standard_normal(N)
    m = nimbleModel(LMM_CODE, constants={"N": N, "G": G, "grp": grp}, data={"y": y},
                    inits={"mu": 0.0, "sig_u": 1.0, "sig_e": 1.0, "u": np.zeros(G)}, name="lmm")
    m.truth = {"mu": 2.0, "sig_e": 1.5, "sig_u": 1.0, "u": u_true}
    conf = configureMCMC(m, monitors=["mu", "sig_u", "sig_e"])
    for v in ("sig_u", "sig_e"):
        conf.removeSamplers(v)
        conf.addSampler(v, type="RW", control={"log": True})
    order = {"mu": 0, "sig_u": 1, "sig_e": 2, "u": 3}
    conf.samplerConfs.sort(key=lambda s: order[s.var])
    return m, conf


OCCUPANCY_CODE = """
{
  b0 ~ dlogis(0, 1)
  a0 ~ dlogis(0, 1)
  b1 ~ dnorm(0, sd = 2.5)
  a1 ~ dnorm(0, sd = 2.5)
  for (i in 1:S) {
    logit(psi[i]) <- b0 + b1 * x[i]
    z[i] ~ dbern(psi[i])
    for (j in 1:J) {
      logit(p[i, j]) <- a0 + a1 * w[i, j]
      y[i, j] ~ dbern(z[i] * p[i, j])
    }
  }
}
"""


def occupancy(S=100_000, J=10, seed=1004):
    """C4: occupancy model, binary sampler on z[i] (one parallel sweep), RW on the four logit coefficients."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(S)
    w = rng.standard_normal((S, J))
    psi = 1 / (1 + np.exp(-(0.3 + 1.0 * x)))
    z = (rng.random(S) < psi).astype(np.float64)
    pdet = 1 / (1 + np.exp(-(-0.5 + 0.7 * w)))
    y = ((rng.random((S, J)) < pdet) * z[:, None]).astype(np.float64)
    zinit = np.maximum(y.max(axis=1), 0.0)
    zinit[zinit == 0] = 1.0  # start every site occupied: always a valid state
    m = nimbleModel(OCCUPANCY_CODE, constants={"S": S, "J": J, "x": x, "w": w}, data={"y": y},
                    inits={"b0": 0.0, "b1": 0.0, "a0": 0.0, "a1": 0.0, "z": zinit}, name="occupancy")
    m.truth = {"b0": 0.3, "b1": 1.0, "a0": -0.5, "a1": 0.7, "z": z}
    conf = configureMCMC(m, monitors=["b0", "b1", "a0", "a1"])
    return m, conf


GLMM_CODE = """
{
  a ~ dgamma(1, 1)
  b ~ dgamma(0.1, 1)
  b0 ~ dnorm(0, sd = 10)
  b1 ~ dnorm(0, sd = 10)
  for (g in 1:G) { lam[g] ~ dgamma(a, b) }
  for (i in 1:N) {
    log(m[i]) <- b0 + b1 * x[i]
    y[i] ~ dpois(lam[grp[i]] * m[i])
  }
}
"""


def glmm(N=1_000_000, G=10_000, seed=1005):
    """C5: Poisson GLMM, conjugate gamma-Poisson lam[g], conjugate gamma-gamma b, RW(log) a, RW b0/b1."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(N)
    grp = np.repeat(np.arange(1, G + 1), N // G)
    if grp.size < N:
        grp = np.concatenate([grp, np.full(N - grp.size, G)])
    lam_true = rng.gamma(2.0, 1.0 / 2.0, G)
    y = rng.poisson(lam_true[grp - 1] * np.exp(0.5 + 0.3 * x)).astype(np.float64)
    m = nimbleModel(GLMM_CODE, constants={"N": N, "G": G, "grp": grp, "x": x}, data={"y": y},
                    inits={"a": 1.0, "b": 1.0, "b0": 0.0, "b1": 0.0, "lam": np.ones(G)}, name="glmm")
    m.truth = {"b0": 0.5, "b1": 0.3, "lam": lam_true}
    conf = configureMCMC(m, monitors=["a", "b", "b0", "b1"])
    conf.removeSamplers("a")
    conf.addSampler("a", type="RW", control={"log": True})
    order = {"a": 0, "b": 1, "b0": 2, "b1": 3, "lam": 4}
    conf.samplerConfs.sort(key=lambda s: order[s.var])
    return m, conf


def evals_per_iteration(cmcmc):
    """logProb evaluations one MCMC iteration costs per chain, counted the way the reference does its work
    (SURVEY.md 8d): MH samplers evaluate each dependent once (calculateDiff against stored logProbs), conjugate
    samplers recalculate the target and its dependents once, the binary sampler calculates the flipped state."""
    b = cmcmc.builder
    total = 0
    for s in b.samplers:
        Tn = s.n_targets
        for e in range(s.n_dep):
            dp = b.deps[s.dep_off + e]
            n = b.decls[dp.decl].n
            if dp.sel == 0:
                cnt = n * Tn
            elif dp.sel == 1:
                cnt = Tn
            elif dp.sel == 2:
                cnt = dp.a * Tn
            elif dp.sel in (3, 4):
                off = b.iarr[dp.a]
                cnt = int(off[Tn] - off[0])
            else:
                cnt = (dp.b - dp.a) * Tn
            total += cnt * (s.tries if s.type == 2 else 1)
    return total
