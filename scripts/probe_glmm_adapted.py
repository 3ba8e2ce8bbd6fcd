"""GLMM (configs[4]) with RW scales of posterior width from the first iteration (every chain inside the moment series): a few
iterations, for `ncu` launch lists of the adapted regime without 2500 warm-up iterations."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nimble_b200 as nb
from nimble_b200 import _capi, configs
C = 8192
m, conf = configs.glmm()
for v, sc, lg in (("a", 0.03, True), ("b0", 0.002, False), ("b1", 0.001, False)):
    conf.removeSamplers(v)
    conf.addSampler(v, type="RW", control={"log": lg, "scale": sc})
order = {"a": 0, "b": 1, "b0": 2, "b1": 3, "lam": 4}
conf.samplerConfs.sort(key=lambda s: order[s.var])
cm = nb.compileNimble(nb.buildMCMC(conf), backend=_capi.default_backend())
dm = cm._ensure(C)
init = np.tile(m.init_vector(), (C, 1))
init[:, m.var_slot["b0"]] = 0.158
init[:, m.var_slot["b1"]] = 0.3
init[:, m.var_slot["a"]] = 2.0
init[:, m.var_slot["b"]] = 1.5
dm.set_values(init)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
dm.run(n, 0, 1, True, 1)
ms, l = dm.last_timing()
print("glmm, posterior-width proposals: %.3f ms/iteration, %.1f launches" % (ms / n, l / n))
