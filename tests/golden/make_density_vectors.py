#!/usr/bin/env python3
"""Generate tests/golden/density_vectors.json: (distribution, arguments) -> log-density to 50 digits.

The reference's tests hold no literal values for dnorm/dgamma/dpois/dbinom/dbeta/dlogis/dexp (they compare
with the running R's own nmath, P/tests/testthat/test-dsl_dists.R:61-66), and neither R nor Rmath exists in
the build container.  These vectors pin the oracle (oracle/nmath_ref.c) to an independent evaluation of the
textbook formulas in 50-digit arithmetic instead.  dcat follows P/inst/CppCode/dists.cpp:914-934 and the
invariance checked in P/tests/testthat/test-distributions.R:259-271.  Run: python tests/golden/make_density_vectors.py"""
import json
import os

import mpmath as mp
import numpy as np

mp.mp.dps = 50
rng = np.random.default_rng(20240611)
F = mp.mpf


def dnorm(x, mu, sd):
    z = (F(x) - F(mu)) / F(sd)
    return -mp.log(F(sd)) - mp.log(2 * mp.pi) / 2 - z * z / 2


def dgamma(x, a, s):
    return (F(a) - 1) * mp.log(F(x)) - F(x) / F(s) - F(a) * mp.log(F(s)) - mp.loggamma(F(a))


def dpois(x, lam):
    return F(x) * mp.log(F(lam)) - F(lam) - mp.loggamma(F(x) + 1)


def dbinom(x, n, p):
    x, n, p = F(x), F(n), F(p)
    lc = mp.loggamma(n + 1) - mp.loggamma(x + 1) - mp.loggamma(n - x + 1)
    return lc + (x * mp.log(p) if x > 0 else 0) + ((n - x) * mp.log(1 - p) if n - x > 0 else 0)


def dbeta(x, a, b):
    x, a, b = F(x), F(a), F(b)
    return (a - 1) * mp.log(x) + (b - 1) * mp.log(1 - x) - (mp.loggamma(a) + mp.loggamma(b) - mp.loggamma(a + b))


def dlogis(x, loc, s):
    z = abs(F(x) - F(loc)) / F(s)
    return -z - mp.log(F(s)) - 2 * mp.log(1 + mp.exp(-z))


def dexp(x, rate):
    return mp.log(F(rate)) - F(rate) * F(x)


def dlnorm(x, ml, sl):
    z = (mp.log(F(x)) - F(ml)) / F(sl)
    return -mp.log(F(x)) - mp.log(F(sl)) - mp.log(2 * mp.pi) / 2 - z * z / 2


rows = []


def add(name, args, val):
    rows.append({"dist": name, "args": [float(a) for a in args], "logp": mp.nstr(val, 30)})


for _ in range(60):
    x, mu, sd = rng.normal(0, 3), rng.normal(0, 2), float(np.exp(rng.normal(0, 1.5)))
    add("dnorm", (x, mu, sd), dnorm(x, mu, sd))
for shape in [0.1, 0.5, 0.7, 1.0, 1.3, 2.0, 2.5, 7.3, 15.5, 16.2, 33.0, 36.5, 90.1, 600.0]:
    for _ in range(5):
        scale = float(np.exp(rng.normal(0, 1)))
        x = float(rng.gamma(shape, scale)) + 1e-3
        add("dgamma", (x, shape, scale), dgamma(x, shape, scale))
for lam in [0.01, 0.3, 1.0, 4.7, 20.0, 150.5, 3000.0]:
    for x in sorted(set([0, 1, 2, int(lam), int(lam) + 1, int(lam * 1.5) + 3, int(lam * 0.5)])):
        add("dpois", (x, lam), dpois(x, lam))
for n in [1, 2, 5, 17, 40, 1000]:
    for p in [1e-6, 0.003, 0.05, 0.0999, 0.1, 0.3, 0.5, 0.77, 0.9, 0.901, 0.97, 0.9999]:
        for x in sorted(set([0, 1, n // 2, n - 1, n])):
            if 0 <= x <= n:
                add("dbinom", (x, n, p), dbinom(x, n, p))
for a, b in [(0.5, 0.5), (1.0, 1.0), (1.5, 3.0), (2.0, 2.0), (2.5, 2.5), (3.0, 7.5), (12.0, 4.0), (0.7, 15.0), (30.0, 50.0)]:
    for x in [0.01, 0.2, 0.5, 0.83, 0.999]:
        add("dbeta", (x, a, b), dbeta(x, a, b))
for _ in range(40):
    x, loc, s = rng.normal(0, 4), rng.normal(0, 2), float(np.exp(rng.normal(0, 1)))
    add("dlogis", (x, loc, s), dlogis(x, loc, s))
for _ in range(20):
    x, rate = float(rng.exponential(2.0)), float(np.exp(rng.normal(0, 1)))
    add("dexp", (x, rate), dexp(x, rate))
for K in [2, 5, 11]:
    prob = rng.random(K) * 3
    for x in range(1, K + 1):
        add("dcat", [x] + list(prob), mp.log(F(prob[x - 1])) - mp.log(mp.fsum([F(q) for q in prob])))
for _ in range(30):   # appended last: the vectors above keep their values
    ml, sl = rng.normal(0, 1.5), float(np.exp(rng.normal(0, 1)))
    x = float(np.exp(rng.normal(ml, 2 * sl)))
    add("dlnorm", (x, ml, sl), dlnorm(x, ml, sl))

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "density_vectors.json")
with open(out, "w") as f:
    json.dump(rows, f, indent=0)
print(len(rows), "vectors ->", out)
