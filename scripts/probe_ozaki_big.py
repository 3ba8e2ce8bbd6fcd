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
n, p, Cn = int(os.environ.get("OZ_N", "1000000")), 50, 1024
X = rng.standard_normal((n, p)) / np.sqrt(p); X[:, 0] = 1.0
beta = np.ascontiguousarray(0.3 * rng.standard_normal((p, Cn)))
y = (rng.random(n) < 0.5).astype(np.float64)
sums = np.empty(Cn); ms = C.c_double()
rc = f(X.ctypes.data_as(dp), y.ctypes.data_as(dp), n, p, beta.ctypes.data_as(dp), Cn, None, sums.ctypes.data_as(dp), C.byref(ms))
print("n=%d: %.3f ms per launch" % (n, ms.value))
