"""First contact with the tcgen05 int8 path: linear predictors and block sums against numpy."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nimble_b200 import _capi
lib = _capi.default_backend().lib
f = lib.nimb200_debug_ozaki
dp = C.POINTER(C.c_double)
f.argtypes = [dp, dp, C.c_int32, C.c_int32, dp, C.c_int32, dp, dp, dp]
f.restype = C.c_int
rng = np.random.default_rng(0)
for n, p, Cn in ((256, 50, 64), (1000, 50, 64), (5000, 7, 100), (100000, 50, 1024)):
    X = rng.standard_normal((n, p)) / np.sqrt(p); X[:, 0] = 1.0
    beta = np.ascontiguousarray(rng.standard_normal((p, Cn)))
    y = (rng.random(n) < 0.5).astype(np.float64)
    eta = np.empty((n, Cn)); sums = np.empty(Cn); ms = C.c_double()
    rc = f(X.ctypes.data_as(dp), y.ctypes.data_as(dp), n, p, beta.ctypes.data_as(dp), Cn, eta.ctypes.data_as(dp), sums.ctypes.data_as(dp), C.byref(ms))
    if rc:
        lib.nimb200_last_error.restype = C.c_char_p
        print("error:", lib.nimb200_last_error()); break
    want = X @ beta
    lp = -np.logaddexp(0, -want) * y[:, None] - np.logaddexp(0, want) * (1 - y[:, None])
    print("n=%d p=%d C=%d: max |eta - X.beta| = %.3e (|eta| max %.2f), block sums rel err %.3e, %.3f ms"
          % (n, p, Cn, np.abs(eta - want).max(), np.abs(want).max(), np.max(np.abs(sums - lp.sum(0)) / np.abs(lp.sum(0))), ms.value), flush=True)
# timing without the diagnostic eta output, at the benchmark's size
n, p, Cn = 1_000_000, 50, 1024
X = rng.standard_normal((n, p)) / np.sqrt(p); X[:, 0] = 1.0
beta = np.ascontiguousarray(0.3 * rng.standard_normal((p, Cn)))
y = (rng.random(n) < 0.5).astype(np.float64)
sums = np.empty(Cn); ms = C.c_double()
rc = f(X.ctypes.data_as(dp), y.ctypes.data_as(dp), n, p, beta.ctypes.data_as(dp), Cn, None, sums.ctypes.data_as(dp), C.byref(ms))
want = X @ beta[:, :8]
lp = (-np.logaddexp(0, -want) * y[:, None] - np.logaddexp(0, want) * (1 - y[:, None])).sum(0)
print("n=%d p=%d C=%d: %.3f ms per launch; block sums rel err (8 chains) %.3e" % (n, p, Cn, ms.value, np.max(np.abs(sums[:8] - lp) / np.abs(lp))))
