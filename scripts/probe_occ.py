"""Occupancy (configs[3]) iteration time and the y-declaration block sum on one GPU."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nimble_b200 as nb
from nimble_b200 import _capi, configs
backend = _capi.default_backend()
m, conf = configs.occupancy()
C = 64
cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
dm = cm._ensure(C)
init = np.tile(m.init_vector(), (C, 1))
rng = np.random.default_rng(4)
for k, v in (("b0", 0.3), ("b1", 1.0), ("a0", -0.5), ("a1", 0.7)):
    init[:, m.var_slot[k]] = v + 0.1 * rng.standard_normal(C)
dm.set_values(init)
dm.run(10, 0, 1, True, 1)
dm.run(100, 0, 1, False, 1)
ms, l = dm.last_timing()
yd = [cm.lowering.decl_ids[d.index] for d in m.decls if d.var == "y"][0]
print("occupancy: %.4f ms/iteration, %.1f launches; y block sum %.1f us" % (ms / 100, l / 100, 1e3 * min(dm.time_calculate(yd, 20) for _ in range(3))), flush=True)
dm.set_timing(True)
dm.run(50, 0, 1, False, 1)
print("   per-sampler ms:", np.round(dm.sampler_times() / 50, 4))
