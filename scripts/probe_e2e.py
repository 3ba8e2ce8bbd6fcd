"""Where the end-to-end time of one runMCMC call goes (C2, host buffers): python scripts/probe_e2e.py [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import nimble_b200 as nb
from nimble_b200 import configs, _capi
K = int(sys.argv[1]) if len(sys.argv) > 1 else 210
C = 1024
m, conf = configs.logistic()
be = _capi.default_backend()
warm = nb.compileNimble(nb.buildMCMC(conf), backend=be)
warm._ensure(8); warm.dm.close()
cm = nb.compileNimble(nb.buildMCMC(conf), backend=be)
inits = configs.logistic_inits(50, C)
t = [time.perf_counter()]
rows = np.stack([m.init_vector(i) for i in inits]); t.append(time.perf_counter())
dm = cm._ensure(C); torch.cuda.synchronize(); t.append(time.perf_counter())
dm.set_values(rows); torch.cuda.synchronize(); t.append(time.perf_counter())
cm._initialize_model_checks(dm); torch.cuda.synchronize(); t.append(time.perf_counter())
dm.run(K, 10, 1, True, 1); torch.cuda.synchronize(); t.append(time.perf_counter())
s = dm.samples(); t.append(time.perf_counter())
mats = [nb.mcmc.SampleMatrix(s[c], cm.colnames) for c in range(C)]; t.append(time.perf_counter())
names = ["init vectors", "model create + upload", "set_values", "initializeModel checks", "run", "samples read-back", "per-chain matrices"]
for n_, a, b in zip(names, t[:-1], t[1:]):
    print("%-26s %8.1f ms" % (n_, (b - a) * 1e3))
ms, _ = dm.last_timing()
print("device-timed iterations     %8.1f ms; total %.1f ms" % (ms, (t[-1] - t[0]) * 1e3))
