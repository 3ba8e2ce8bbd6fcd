"""Per-sampler device time of every BASELINE.json configuration (run(time = TRUE) / getTimes()): python scripts/sampler_times.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import configs, _capi
be = _capi.default_backend()
cases = [("pump", configs.pump, 4, 400), ("logistic", configs.logistic, 1024, 20), ("lmm", configs.lmm, 8, 100),
         ("occupancy", configs.occupancy, 64, 30), ("glmm", configs.glmm, 8192, 5)]
for name, mk, C, K in cases:
    m, conf = mk()
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=be)
    cm.check_initial_logprobs = False
    inits = np.tile(m.init_vector(), (C, 1))
    cm.run_chains(inits, 3, 0, 1, True, 1)
    cm.run_chains(inits, K, 0, 1, True, 1, time=True)
    t = cm.getTimes() * 1e3 / K
    parts = ", ".join("%s(%s) %.3f" % (sc.name, sc.var, x) for sc, x in zip(conf.samplerConfs, t))
    print("%s, %d chains, ms per iteration by sampler (ungraphed launch path): %s; sum %.3f" % (name, C, parts, t.sum()), flush=True)
    cm.dm.close()
