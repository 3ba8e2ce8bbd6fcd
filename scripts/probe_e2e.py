"""Where does the end-to-end time of one runMCMC() call on host buffers go (C2, 1024 chains, 25 iterations)?"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nimble_b200 as nb
from nimble_b200 import _capi, configs

backend = _capi.default_backend()
m, conf = configs.logistic()
C = 1024
inits_all = configs.logistic_inits(50, C)
for rep in range(2):
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rows = np.stack([m.init_vector(i) for i in inits_all])
    t1 = time.perf_counter()
    dm = cm._ensure(C)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    dm.set_values(rows)
    t3 = time.perf_counter()
    cm.check_initial_logprobs = True
    cm._initialize_model_checks(dm)
    t4 = time.perf_counter()
    dm.run(25, 5, 1, True, 2024)
    t5 = time.perf_counter()
    s = dm.samples()
    t6 = time.perf_counter()
    print("rep %d: inits %.1f ms, model_create (pack + upload) %.1f ms, set_values %.1f ms, initializeModel checks %.1f ms, run %.1f ms, samples %.1f ms; total %.1f ms"
          % (rep, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t5 - t4) * 1e3, (t6 - t5) * 1e3, (t6 - t0) * 1e3), flush=True)
    dm.close()
