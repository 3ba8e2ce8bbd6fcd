"""GPU fuzz: randomly assembled models (tests/fuzz_models.py), default sampler assignment, device against oracle through the
C ABI -- monitored samples and the complete final state of every chain, across at least one adaptation step."""
import os

import numpy as np
import pytest

from tests.fuzz_models import make_case
from tests.helpers import compile_on, rel_err, run

pytestmark = pytest.mark.gpu


_FIRST = int(os.environ.get("NIMB200_FUZZ_FIRST", "0"))
_CASES = int(os.environ.get("NIMB200_FUZZ_CASES", "50"))


@pytest.mark.parametrize("seed", range(_FIRST, _FIRST + _CASES))
def test_random_model_matches_oracle(oracle, b200, seed):
    m, conf, nchains = make_case(seed)
    rng = np.random.default_rng(1000 + seed)
    base = m.init_vector()
    inits = np.tile(base, (nchains, 1))
    cont = np.array([not float(v).is_integer() for v in base])            # perturb the continuous slots only
    inits[:, cont] *= np.exp(0.05 * rng.standard_normal((nchains, int(cont.sum()))))
    co, cg = compile_on(oracle, conf), compile_on(b200, conf)
    niter = 130
    so, sg = run(co, inits, niter, 0, seed=seed + 1), run(cg, inits, niter, 0, seed=seed + 1)
    assert so.shape == sg.shape
    names = [s.name for s in conf.samplerConfs]
    ok = np.isfinite(so)
    assert ok.mean() > 0.99, names
    err = rel_err(sg, so)
    assert err.max() < 1e-8, (names, err.max(), np.unravel_index(err.argmax(), err.shape))
    assert rel_err(cg.dm.get_values(), co.dm.get_values()).max() < 1e-8, names
