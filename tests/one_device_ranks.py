"""Two ranks of a data-sharded group on ONE device of ONE process (python -m tests.one_device_ranks <model>; run by
tests/test_multi_gpu.py with CUDA_MODULE_LOADING=EAGER).  Both ranks must produce the same samples, equal to the unsharded
run's to 1e-9, and the online WAIC summed over the ranks' slices must equal the unsharded WAIC."""
import sys
import threading

import numpy as np


def main(name):
    import nimble_b200 as nb
    from nimble_b200 import _capi
    from nimble_b200.parallel import waic_from_partials
    from tests import small_models as sm
    backend = _capi.default_backend()
    m, conf = sm.SMALL[name]()
    C = 4
    inits = sm.random_inits(m, C, 21, 0.1)
    cs = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=0)
    ds = cs._ensure(C)
    ds.waic_enable(True)
    single = cs.run_chains(inits, 120, 10, 1, True, 17)
    w1 = ds.waic_get()
    cms = [nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=0) for _ in (0, 1)]
    dms = [cm._ensure(C) for cm in cms]
    handles = [dm.peer_export(2) for dm in dms]
    for r, dm in enumerate(dms):
        dm.enable_data_sharding_peer(2, r, handles, min_nodes=256)
        dm.waic_enable(True)
    out, errs = [None, None], []

    def work(r):
        try:
            out[r] = cms[r].run_chains(inits, 120, 10, 1, True, 17)
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)
    ths = [threading.Thread(target=work, args=(r,)) for r in (0, 1)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    assert np.array_equal(out[0], out[1]), "the two ranks disagree"
    err = np.abs(out[0] - single) / np.maximum(1.0, np.abs(single))
    assert err.max() < 1e-9, err.max()
    w2 = waic_from_partials([dm.waic_partial() for dm in dms])
    for k in ("WAIC", "lppd", "pWAIC"):
        assert abs(w2[k] - w1[k]) < 1e-8 * max(1.0, abs(w1[k])), (k, w2[k], w1[k])
    print("ok", name, "max rel diff vs unsharded %.2e" % err.max(), "WAIC", w2["WAIC"], w1["WAIC"], "all-reduces per iteration", dms[0].last_allreduces())


if __name__ == "__main__":
    main(sys.argv[1])
