"""Time the streaming normal block sum (N = 1e7, 8 chains) for sorted / permuted / index-free means:
python scripts/probe_norm.py [reps]   (NIMB200_NORM_FLAGS is switched between timings)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import configs, _capi
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
C = 8
N = int(float(os.environ.get("PROBE_N", "1e7")))
be = _capi.default_backend()


def noindex():
    rng = np.random.default_rng(1)
    m = nb.nimbleModel("{ mu ~ dnorm(0, sd = 100)\n s ~ dgamma(2, 1)\n for (i in 1:N) { y[i] ~ dnorm(mu, sd = s) } }",
                       constants={"N": N}, data={"y": 2 + rng.standard_normal(N)}, inits={"mu": 0.0, "s": 1.0})
    return m, nb.configureMCMC(m, useConjugacy=False)


models = {"sorted": lambda: configs.lmm(N=N, G=N // 1000), "permuted": lambda: configs.lmm(N=N, G=N // 1000, permute=True), "noindex": noindex}
if N > 10_000_000:
    del models["permuted"]
for name, mk in models.items():
    m, conf = mk()
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=be)
    dm = cm._ensure(C)
    dm.set_values(np.tile(m.init_vector(), (C, 1)))
    y = len(cm.builder.decls) - 1
    b = N * (12 if name != "noindex" else 8)
    for flags in (0,):
        os.environ["NIMB200_NORM_FLAGS"] = str(flags)
        ref = dm.calculate(y).copy()
        ms = min(dm.time_calculate(y, reps) for _ in range(3))
        print("N=%g %-9s flags=%d: %.2f us per block sum (2 launches), %.0f GB/s algorithmic, sum[0]=%.6f" % (N, name, flags, ms * 1e3, b / ms / 1e6, ref[0]), flush=True)
