"""Randomly assembled small models for the device-versus-oracle fuzz test: every case draws sizes, groupings (sorted, permuted,
with empty groups), link structure, priors and sampler options from a seeded generator, so that selections (ALL / SAME / STRIDE /
SEG / CSR), index modes, conjugate links and sampler kinds meet in combinations no hand-written fixture covers."""
import numpy as np

import nimble_b200 as nb


def _groups(rng, N, G):
    kind = rng.integers(0, 3)
    if kind == 0:                                   # contiguous blocks
        g = np.sort(rng.integers(1, G + 1, N))
    elif kind == 1:                                 # arbitrary order
        g = rng.integers(1, G + 1, N)
    else:                                           # some groups empty
        live = rng.choice(np.arange(1, G + 1), size=max(1, G // 2), replace=False)
        g = rng.choice(live, N)
    return g.astype(np.int64)


def normal_hierarchy(rng):
    N, G = int(rng.integers(1, 40)), int(rng.integers(1, 7))
    g, x = _groups(rng, N, G), rng.standard_normal(N)
    mean = rng.choice(["mu + u[g[i]]", "u[g[i]] + b * x[i]", "mu + b * x[i] + u[g[i]]", "a[g[i]] * x[i] + mu"])
    lines = ["mu ~ dnorm(0, sd = 5)", "b ~ dnorm(0, 0.5)", "su ~ dgamma(2, 1)",
             "for (k in 1:G) { u[k] ~ dnorm(0, sd = su)\n a[k] ~ dnorm(1, sd = 2) }"]
    if rng.random() < 0.5:
        lines += ["tau ~ dgamma(2, 2)", "for (i in 1:N) { y[i] ~ dnorm(%s, tau) }" % mean]
        inits = {"tau": 1.0}
    else:
        lines += ["se ~ dgamma(2, 1)", "for (i in 1:N) { y[i] ~ dnorm(%s, sd = se) }" % mean]
        inits = {"se": 1.0}
    inits.update({"mu": 0.1, "b": 0.2, "su": 1.0, "u": 0.1 * rng.standard_normal(G), "a": np.ones(G)})
    code = "{ " + "\n".join(lines) + " }"
    m = nb.nimbleModel(code, constants={"N": N, "G": G, "g": g, "x": x}, data={"y": 1 + x + rng.standard_normal(N)}, inits=inits)
    return m


def poisson_hierarchy(rng):
    N, G = int(rng.integers(1, 40)), int(rng.integers(1, 7))
    g, x, e = _groups(rng, N, G), rng.standard_normal(N), rng.gamma(2, 1, N)
    rate = rng.choice(["lam[g[i]] * e[i]", "lam[g[i]] * exp(b0 + b1 * x[i])", "exp(b0 + b1 * x[i]) * e[i]", "lam[g[i]]"])
    code = """{ a ~ dgamma(2, 1)
      r ~ dgamma(1, 1)
      b0 ~ dnorm(0, sd = 2)
      b1 ~ dnorm(0, sd = 2)
      for (k in 1:G) { lam[k] ~ dgamma(a, r) }
      for (i in 1:N) { y[i] ~ dpois(%s) } }""" % rate
    m = nb.nimbleModel(code, constants={"N": N, "G": G, "g": g, "x": x, "e": e}, data={"y": rng.poisson(2.0, N).astype(float)},
                       inits={"a": 1.0, "r": 1.0, "b0": 0.1, "b1": -0.1, "lam": np.ones(G)})
    return m


def bernoulli_layers(rng):
    S, J = int(rng.integers(1, 15)), int(rng.integers(1, 5))
    x, w = rng.standard_normal(S), rng.standard_normal((S, J))
    gate = rng.random() < 0.6
    code = """{ b0 ~ dlogis(0, 1)
      b1 ~ dnorm(0, sd = 2)
      a0 ~ dnorm(0, sd = 2)
      a1 ~ dnorm(0, sd = 2)
      for (i in 1:S) {
        logit(psi[i]) <- b0 + b1 * x[i]
        z[i] ~ dbern(psi[i])
        for (j in 1:J) {
          logit(p[i, j]) <- a0 + a1 * w[i, j]
          y[i, j] ~ dbern(%s) } } }""" % ("z[i] * p[i, j]" if gate else "p[i, j]")
    z = (rng.random(S) < 0.6).astype(float)
    y = ((rng.random((S, J)) < 0.5) * (z[:, None] if gate else 1.0)).astype(float)
    m = nb.nimbleModel(code, constants={"S": S, "J": J, "x": x, "w": w}, data={"y": y},
                       inits={"b0": 0.1, "b1": 0.2, "a0": -0.1, "a1": 0.3, "z": np.maximum(z, y.max(axis=1))})
    return m


def counts_and_labels(rng):
    n, K = int(rng.integers(2, 25)), int(rng.integers(2, 5))
    w = rng.random(K) + 0.2
    y = rng.standard_normal(n) * 2
    size = rng.integers(1, 9, n).astype(float)
    c = rng.binomial(size.astype(int), 0.4).astype(float)
    code = """{ p ~ dbeta(2, 2)
      lam ~ dgamma(2, 0.3)
      for (k in 1:K) { mu[k] ~ dnorm(0, sd = 5) }
      for (i in 1:n) {
        z[i] ~ dcat(w[1:K])
        y[i] ~ dnorm(mu[z[i]], sd = s[z[i]])
        M[i] ~ dpois(lam)
        c[i] ~ dbin(p, M[i]) } }"""
    m = nb.nimbleModel(code, constants={"n": n, "K": K, "w": w, "s": 0.5 + rng.random(K)}, data={"y": y, "c": c},
                       inits={"p": 0.5, "lam": 5.0, "mu": np.linspace(-2, 2, K), "z": np.ones(n), "M": c + 1})
    return m


def mixed_bag(rng):
    """Every remaining distribution as a likelihood or prior (dbeta, dexp, dlogis, dcat data, dbin with constant size), priors
    that are not conjugate to them, and hand-assigned samplers: slice on a continuous node, RW on the log scale, reflective RW."""
    n, K = int(rng.integers(1, 30)), int(rng.integers(2, 5))
    size = rng.integers(1, 7, n).astype(float)
    code = """{ p ~ dbeta(2, 3)
      r ~ dexp(1.5)
      loc ~ dlogis(0, 1)
      sh ~ dgamma(3, 2)
      q ~ dbeta(1.5, 1.5)
      for (i in 1:n) {
        k[i] ~ dbin(p, size[i])
        t[i] ~ dexp(r * e[i])
        l[i] ~ dlogis(loc, scale = sh)
        b[i] ~ dbeta(sh, 2)
        d[i] ~ dcat(w[1:K])
        g[i] ~ dgamma(sh, rate = r)
        f[i] ~ dbern(q * p) } }"""
    data = {"k": rng.binomial(size.astype(int), 0.4).astype(float), "t": rng.exponential(1.0, n), "l": rng.logistic(0, 1, n),
            "b": rng.beta(2, 2, n), "d": rng.integers(1, K + 1, n).astype(float), "g": rng.gamma(2, 1, n),
            "f": (rng.random(n) < 0.3).astype(float)}
    m = nb.nimbleModel(code, constants={"n": n, "K": K, "size": size, "e": 0.5 + rng.random(n), "w": rng.random(K) + 0.1}, data=data,
                       inits={"p": 0.4, "r": 1.0, "loc": 0.1, "sh": 1.5, "q": 0.5})
    return m


TEMPLATES = [normal_hierarchy, poisson_hierarchy, bernoulli_layers, counts_and_labels, mixed_bag]


def make_case(seed):
    rng = np.random.default_rng(seed)
    m = TEMPLATES[seed % len(TEMPLATES)](rng)
    conf = nb.configureMCMC(m, useConjugacy=bool(rng.random() < 0.7),
                            control={"adaptInterval": int(rng.choice([20, 50, 200]))})
    if "loc" in m.var_slot:                         # mixed_bag: hand-assigned samplers on some of the scalars
        choice = int(rng.integers(0, 4))
        if choice == 0:
            conf.removeSamplers("loc")
            conf.addSampler("loc", type="slice", control={"sliceWidth": 0.7, "adaptInterval": 25})
        elif choice == 1:
            conf.removeSamplers("sh")
            conf.addSampler("sh", type="RW", control={"log": True, "adaptInterval": 30})
        elif choice == 2:
            conf.removeSamplers("q")
            conf.addSampler("q", type="RW", control={"reflective": True, "lower": 0.0, "upper": 1.0, "scale": 0.4, "adaptInterval": 30})
    if rng.random() < 0.3:                          # an RW_block over a vector parameter instead of its scalar samplers
        for var in ("u", "lam", "mu"):
            if var in m.var_slot and m.var_shape[var] and 1 < m.var_shape[var][0] <= 6:
                conf.removeSamplers(var)
                conf.addSampler("%s[1:%d]" % (var, m.var_shape[var][0]), type="RW_block",
                                control={"adaptInterval": 40, "scale": 0.5})
                break
    return m, conf, int(rng.integers(1, 50))
