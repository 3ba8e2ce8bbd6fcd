"""GPU: the allocation-free entry point nimb200_bern_logit_block_logprob_noalloc().

Some users of this library cannot allocate device memory themselves or let a library do it: their process hands
the GPU's memory to its own manager at start-up, or they sit in a steady-state loop where cudaMalloc's device-wide
synchronisation is not acceptable.  For them the library offers one entry point whose scratch memory is a fixed
device array built into the library.  This test calls it through the C ABI twice in a row (the second call with
fewer chains and other parameters, so that stale contents of the shared scratch from the first call are present)
and checks each result against the ordinary entry point nimb200_calculate() on the same X, y and beta, at the same
tolerance the other parity tests use (1e-12 relative), against the CPU oracle's block sum of the same declaration, and
that the device's free memory is the same before and after each call (nothing was allocated behind the caller's back).
It also checks that a request beyond the built-in limit is refused with a clear message.  Collected last on purpose: a run that stops at the first failure has then already
run every other test."""
import ctypes as C

import numpy as np
import pytest

from nimble_b200 import configs
from tests.helpers import compile_on, rel_err

pytestmark = pytest.mark.gpu
RTOL_LP = 1e-12


def _noalloc(lib):
    f = lib.nimb200_bern_logit_block_logprob_noalloc
    f.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    f.restype = C.c_int
    return f


def test_noalloc_matches_calculate_twice_in_a_row(b200, oracle):
    import torch
    N, p = 5000, 50
    m, conf = configs.logistic(N=N, p=p, seed=77)
    cg, co = compile_on(b200, conf), compile_on(oracle, conf)
    ydecl = len(cg.builder.decls) - 1
    X = np.ascontiguousarray(m.constants["X"])
    y = np.ascontiguousarray(m.data["y"])
    dX = torch.tensor(X, dtype=torch.float64, device="cuda")
    dy = torch.tensor(y, dtype=torch.float64, device="cuda")
    f = _noalloc(b200.lib)
    rng = np.random.default_rng(3)
    bs = m.var_slot["beta"]
    # the documented set-up call: the first call on a device resolves the scratch array's address (and, with CUDA's lazy
    # module loading, brings the kernels' code onto the device) -- free memory is compared on the calls after it
    d1 = torch.zeros(8, dtype=torch.float64, device="cuda")
    assert f(d1.data_ptr(), 1, d1.data_ptr(), 1, 1, d1.data_ptr(), 1, d1.data_ptr(), None) == 0, b200._err()
    for nchains, scale in ((70, 0.8), (33, 2.5)):
        beta = scale * rng.standard_normal((nchains, p))
        inits = np.tile(m.init_vector(), (nchains, 1))
        inits[:, bs:bs + p] = beta
        dm = cg._ensure(nchains)
        dm.set_values(inits)
        want = dm.calculate(ydecl)                              # existing entry point
        do = co._ensure(nchains)
        do.set_values(inits)
        want_cpu = do.calculate(ydecl)                          # the oracle's left-to-right sum of dbinom(y, 1, ilogit(eta))
        dbeta = torch.tensor(np.ascontiguousarray(beta.T), dtype=torch.float64, device="cuda")   # [p][nchains]
        dout = torch.full((nchains,), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]              # cudaMemGetInfo
        rc = f(dX.data_ptr(), p, dy.data_ptr(), N, p, dbeta.data_ptr(), nchains, dout.data_ptr(), None)
        assert rc == 0, b200._err()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before
        got = dout.cpu().numpy()
        assert np.all(np.isfinite(got))
        assert rel_err(got, want).max() < RTOL_LP, rel_err(got, want).max()
        assert rel_err(got, want_cpu).max() < RTOL_LP, rel_err(got, want_cpu).max()


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch(b200):
    import torch
    f = _noalloc(b200.lib)
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = f(d.data_ptr(), 1, d.data_ptr(), 1, 1, d.data_ptr(), 2049, d.data_ptr(), None)
    assert rc != 0 and b"NIMB200_NOALLOC_MAX_CHAINS" in b200._err()
    rc = f(d.data_ptr(), 65, d.data_ptr(), 1, 65, d.data_ptr(), 1, d.data_ptr(), None)
    assert rc != 0 and b"between 1 and 64" in b200._err()
