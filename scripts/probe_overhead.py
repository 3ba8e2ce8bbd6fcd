"""Where does the non-kernel time of a C2 iteration go?  Small-N models isolate the chain-level kernels and the launch gaps."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nimble_b200 as nb
from nimble_b200 import _capi, configs

backend = _capi.default_backend()
for N, C in ((1280, 1024), (1280, 128), (1_000_000, 1024), (1_000_000, 128)):
    m, conf = configs.logistic(N=N, p=50, seed=1001)
    inits = np.stack([m.init_vector(i) for i in configs.logistic_inits(50, C)])
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    dm = cm._ensure(C)
    dm.set_values(inits)
    dm.run(10, 0, 1, True, 1)
    dm.run(100, 0, 1, False, 1)
    ms, l = dm.last_timing()
    y = len(cm.builder.decls) - 1
    k = dm.time_calculate(y, 10)
    print("N=%d C=%d: %.4f ms/iteration, %d launches/iteration, fused kernel alone %.4f ms -> rest %.4f ms" % (N, C, ms / 100, l // 100, k, ms / 100 - k), flush=True)
    dm.set_timing(True)
    dm.run(50, 0, 1, False, 1)
    print("   per-sampler (ungraphed) ms/iteration:", np.round(dm.sampler_times() / 50, 4))
    dm.close()
