"""Time the block-sum evaluation of the last declaration of a configuration: python scripts/probe_family.py <lmm|glmm> [chains]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nimble_b200 as nb
from nimble_b200 import configs, _capi
name = sys.argv[1]; C = int(sys.argv[2]) if len(sys.argv) > 2 else (8 if name == "lmm" else 8192)
m, conf = configs.lmm() if name == "lmm" else configs.glmm()
cm = nb.compileNimble(nb.buildMCMC(conf), backend=_capi.default_backend())
dm = cm._ensure(C)
dm.set_values(np.tile(m.init_vector(), (C, 1)))
y = len(cm.builder.decls) - 1
n = cm.builder.decls[y].n
ms = dm.time_calculate(y, 20)
bytes_alg = n * (8 + 4) if name == "lmm" else n * (8 + 8 + 8 + 4)
print("%s C=%d: %.4f ms per block-sum evaluation, %.3g evals/s, algorithmic %.1f MB -> %.0f GB/s" % (name, C, ms, n * C / ms * 1e3, bytes_alg / 1e6, bytes_alg / ms / 1e6))
