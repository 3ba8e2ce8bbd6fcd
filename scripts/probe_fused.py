"""Probe: fused logistic block-sum kernel time as a function of the spread of the linear predictor."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import configs, _capi

N = int(os.environ.get("PN", 1_000_000)); C = int(os.environ.get("PC", 1024))
m, conf = configs.logistic(N=N, p=50)
cm = nb.compileNimble(nb.buildMCMC(conf), backend=_capi.default_backend())
dm = cm._ensure(C)
y = len(cm.builder.decls) - 1
rng = np.random.default_rng(0)
bs = m.var_slot["beta"]
for s in (0.0, 0.1, 0.5, 1.0, 2.0, 4.0):
    init = np.tile(m.init_vector(), (C, 1))
    init[:, bs:bs + 50] = s * rng.standard_normal((C, 50))
    dm.set_values(init)
    ms = dm.time_calculate(y, 5)
    print("beta sd %.1f  kernel %.2f ms  %.1f Gevals/s" % (s, ms, N * C / ms / 1e6), flush=True)
