"""Time one BASELINE.json configuration on the GPU: python scripts/run_config.py <pump|logistic|lmm|occupancy|glmm> [chains] [iters] [scale]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import configs, _capi

name = sys.argv[1]
defaults = {"pump": (4, 2000), "logistic": (1024, 20), "lmm": (8, 400), "occupancy": (64, 100), "glmm": (8192, 5)}
C = int(sys.argv[2]) if len(sys.argv) > 2 else defaults[name][0]
K = int(sys.argv[3]) if len(sys.argv) > 3 else defaults[name][1]
scale = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
t0 = time.time()
if name == "pump": m, conf = configs.pump()
elif name == "logistic": m, conf = configs.logistic(N=int(1e6 * scale))
elif name == "lmm": m, conf = configs.lmm(N=int(1e7 * scale), G=int(1e4 * scale))
elif name == "occupancy": m, conf = configs.occupancy(S=int(1e5 * scale))
else: m, conf = configs.glmm(N=int(1e6 * scale), G=int(1e4 * scale))
t1 = time.time()
cm = nb.compileNimble(nb.buildMCMC(conf), backend=_capi.default_backend())
ev = configs.evals_per_iteration(cm)
t2 = time.time()
inits = np.tile(m.init_vector(), (C, 1))
dm = cm._ensure(C)
dm.set_values(inits)
t3 = time.time()
dm.run(3, 0, 1, True, 1)
dm.run(K, 0, 1, False, 1)
ms, launches = dm.last_timing()
print("%s: chains %d, model %.1fs lower %.1fs upload %.1fs | %.3f ms/iter, %d launches/iter, %.3g evals/iter/chain, %.3g evals/s"
      % (name, C, t1 - t0, t2 - t1, t3 - t2, ms / K, launches // K, ev, ev * C * K / (ms * 1e-3)))
