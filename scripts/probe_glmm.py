"""GLMM (configs[4]) iteration time on one GPU, with and without the group-sum path.  GLMM_WARM=2000 measures the adapted
regime (proposals of posterior width: every chain inside the moment series), the default 5 the first iterations (unit-scale proposals)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nimble_b200 as nb
from nimble_b200 import _capi, configs

backend = _capi.default_backend()
C = int(os.environ.get("GLMM_C", "8192"))
m, conf = configs.glmm()
cases = (("group sums", None),) if os.environ.get("GLMM_ONLY_GROUPS") else (("group sums", None), ("node kernels (NIMB200_NO_GROUPSUMS)", "1"))
for label, env in cases:
    if env:
        os.environ["NIMB200_NO_GROUPSUMS"] = env
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    dm = cm._ensure(C)
    dm.set_values(np.tile(m.init_vector(), (C, 1)))
    dm.run(int(os.environ.get("GLMM_WARM", "5")), 0, 1, True, 1)   # 2000+: past the adaptation of the RW scales
    dm.run(30, 0, 1, False, 1)
    ms, l = dm.last_timing()
    s = dm.samples()[:, -30:, :]
    print("%s: %.3f ms/iteration, %.1f launches/iteration; means %s" % (label, ms / 30, l / 30, np.round(s.mean((0, 1)), 4)), flush=True)
    dm.set_timing(True)
    dm.run(10, 0, 1, False, 1)
    print("   per-sampler (ungraphed) ms/iteration:", np.round(dm.sampler_times() / 10, 3))
    dm.close()
