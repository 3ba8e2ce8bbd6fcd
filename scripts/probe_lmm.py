"""LMM (configs[2]) iteration time and launch count on one GPU, with and without the group-statistics block sums."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nimble_b200 as nb
from nimble_b200 import _capi, configs

backend = _capi.default_backend()
N = int(os.environ.get("LMM_N", "10000000"))
m, conf = configs.lmm(N=N)
for label, env in (("group statistics", None), ("streaming passes (NIMB200_NO_BYPRODUCT)", "1")):
    if env:
        os.environ["NIMB200_NO_BYPRODUCT"] = env
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    dm = cm._ensure(8)
    dm.set_values(np.tile(m.init_vector(), (8, 1)))
    dm.run(20, 0, 1, True, 1)
    dm.run(500, 0, 1, False, 1)
    ms, l = dm.last_timing()
    s = dm.samples()[:, -500:, :]
    print("%s: %.4f ms/iteration, %.1f launches/iteration; posterior means %s" % (label, ms / 500, l / 500, np.round(s.mean((0, 1)), 4)), flush=True)
    dm.set_timing(True)
    dm.run(100, 0, 1, False, 1)
    print("   per-sampler (ungraphed) ms/iteration:", np.round(dm.sampler_times() / 100, 4))
    dm.close()
