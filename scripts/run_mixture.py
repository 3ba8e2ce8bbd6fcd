"""Time a finite Gaussian mixture with latent labels (categorical sampler + dynamic indexing + conjugate component means):
python scripts/run_mixture.py [n] [K] [chains] [iters]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import _capi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 3
C = int(sys.argv[3]) if len(sys.argv) > 3 else 256
its = int(sys.argv[4]) if len(sys.argv) > 4 else 30
rng = np.random.default_rng(1)
zt = rng.integers(1, K + 1, n)
mu_t = np.linspace(-4, 4, K)
y = mu_t[zt - 1] + 0.7 * rng.standard_normal(n)
code = """{ for (k in 1:K) { mu[k] ~ dnorm(0, sd = 10) }
  for (i in 1:n) { z[i] ~ dcat(w[1:K])
    y[i] ~ dnorm(mu[z[i]], sd = 0.7) } }"""
m = nb.nimbleModel(code, constants={"n": n, "K": K, "w": np.full(K, 1.0 / K)}, data={"y": y},
                   inits={"mu": np.linspace(-1, 1, K), "z": rng.integers(1, K + 1, n).astype(float)})
conf = nb.configureMCMC(m, monitors=["mu"])
cm = nb.compileNimble(nb.buildMCMC(conf), backend=_capi.default_backend())
dm = cm._ensure(C)
dm.set_values(np.tile(m.init_vector(), (C, 1)))
dm.run(3, 0, 1, True, 1)
dm.run(its, 0, 1, False, 1)
ms, launches = dm.last_timing()
s = dm.samples()
print("mixture n=%d K=%d chains=%d: %.3f ms/iter, %d launches/iter, label draws/s %.3g, mu (chain 0, last) %s"
      % (n, K, C, ms / its, launches // its, n * C * its / (ms * 1e-3), np.sort(s[0, -1]).round(2)))
