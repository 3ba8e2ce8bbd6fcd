"""TEST INFRASTRUCTURE: binds oracle/libnimble_ref.so (the CPU parity oracle) behind the same Python
`Backend` interface the product uses, so tests can run one lowered table through both.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
import ctypes as C
import os
import subprocess

from nimble_b200 import _capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "libnimble_ref.so")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


_backend = None


def oracle_backend():
    global _backend
    if _backend is None:
        if not os.path.exists(LIB):
            build()
        b = _capi.Backend(LIB, prefix="ref_", has_device=False, name="oracle")
        vp = C.c_void_p
        b.prepare = b.lib.ref_mcmc_prepare
        b.prepare.argtypes = [vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32]
        b.run_range = b.lib.ref_mcmc_run_range
        b.run_range.argtypes = [vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_uint64, C.c_int32, C.c_int32]
        b.finish = b.lib.ref_mcmc_finish
        b.finish.argtypes = [vp, C.c_int64, C.c_int64, C.c_int32]
        b.eval_count = b.lib.ref_eval_count
        b.eval_count.argtypes, b.eval_count.restype = [vp], C.c_uint64
        for name in ("dnorm", "dgamma", "dbinom", "dbeta", "dlogis", "dlnorm"):
            f = getattr(b.lib, "ref_" + name)
            f.argtypes, f.restype = [C.c_double] * 3, C.c_double
        for name in ("dpois", "dexp_nimble", "bd0", "lbeta"):
            f = getattr(b.lib, "ref_" + name)
            f.argtypes, f.restype = [C.c_double] * 2, C.c_double
        b.lib.ref_stirlerr.argtypes, b.lib.ref_stirlerr.restype = [C.c_double], C.c_double
        b.lib.ref_dcat.argtypes, b.lib.ref_dcat.restype = [C.c_double, C.POINTER(C.c_double), C.c_int], C.c_double
        _backend = b
    return _backend
