"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle."""
import numpy as np


def compare_backends_smoke():
    from nimble_b200 import _capi
    from oracle.oracle_capi import oracle_backend
    from tests import small_models as sm
    from tests.helpers import compile_on, rel_err, run
    gpu, ref = _capi.default_backend(), oracle_backend()
    m, conf = sm.SMALL["logistic"]()
    inits = sm.random_inits(m, 8, 1, 0.1)
    cg, co = compile_on(gpu, conf), compile_on(ref, conf)
    sg, so = run(cg, inits, 60, 10, seed=1), run(co, inits, 60, 10, seed=1)
    err = rel_err(sg, so).max()
    assert err < 1e-9, err
    y = len(cg.builder.decls) - 1
    assert rel_err(cg.dm.calculate(y), co.dm.calculate(y)).max() < 1e-12
    ms, launches = cg.dm.last_timing()
    print("smoke: 60 iterations x 8 chains, max rel err vs oracle %.2e, %d kernel launches, %.2f ms" % (err, launches, ms))
