import numpy as np

import nimble_b200 as nb


def compile_on(backend, conf):
    return nb.compileNimble(nb.buildMCMC(conf), backend=backend)


def run(cm, inits, niter, nburnin=0, thin=1, seed=1, reset=True):
    return cm.run_chains(np.asarray(inits, dtype=np.float64), niter, nburnin, thin, reset, seed)


def pooled(samples):
    return samples.reshape(-1, samples.shape[-1])


def rel_err(a, b):
    a, b = np.asarray(a), np.asarray(b)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    both_nan = np.isnan(a) & np.isnan(b)
    d = np.abs(a - b) / np.maximum(1.0, np.abs(b))
    d[both_inf | both_nan] = 0.0
    return d
