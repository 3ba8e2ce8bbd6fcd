import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import configs, _capi
N = int(os.environ.get("PN", 1_000_000)); C = int(os.environ.get("PC", 1024))
m, conf = configs.logistic(N=N, p=50)
cm = nb.compileNimble(nb.buildMCMC(conf), backend=_capi.default_backend())
dm = cm._ensure(C)
rng = np.random.default_rng(0)
init = np.tile(m.init_vector(), (C, 1)); bs = m.var_slot["beta"]
init[:, bs:bs + 50] = rng.standard_normal((C, 50))
dm.set_values(init)
print("kernel ms", dm.time_calculate(len(cm.builder.decls) - 1, 3))
