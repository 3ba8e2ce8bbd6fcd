"""Timeline of CTA (0,0) of the two-CTA tcgen05 kernel (nimb200_debug_ozaki_trace): per tile, MMA warp and epilogue warps."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nimble_b200 import _capi
lib = _capi.default_backend().lib
f = lib.nimb200_debug_ozaki
dp = C.POINTER(C.c_double)
f.argtypes = [dp, dp, C.c_int32, C.c_int32, dp, C.c_int32, dp, dp, dp]
f.restype = C.c_int
tr = lib.nimb200_debug_ozaki_trace
tr.argtypes = [C.c_void_p]; tr.restype = C.c_int
rng = np.random.default_rng(0)
n, p, Cn = 1000000, 50, 1024
X = rng.standard_normal((n, p)) / np.sqrt(p); X[:, 0] = 1.0
beta = np.ascontiguousarray(0.3 * rng.standard_normal((p, Cn)))
y = (rng.random(n) < 0.5).astype(np.float64)
sums = np.empty(Cn)
assert tr(None) == 0
rc = f(X.ctypes.data_as(dp), y.ctypes.data_as(dp), n, p, beta.ctypes.data_as(dp), Cn, None, sums.ctypes.data_as(dp), None)
out = np.zeros((18, 64, 2, 4), dtype=np.int64)
assert tr(out.ctypes.data) == 0
t0 = out[1, 20, 0, 0]
print("MMA warp (tile): wait-start, wait-done, issued+commit")
for it in range(20, 26):
    print(it, out[1, it, 0, :3] - t0)
for w in (2, 3, 6, 9):
    print("epilogue warp %d (tile): before t_full wait, after | batch 0 loads done, math done | batch 1 loads done, math done" % w)
    for it in range(20, 24):
        print(it, out[w, it, 0, :2] - t0, out[w, it, 0, 2:] - t0, out[w, it, 1, 2:] - t0)
e = out[2:10, 20:60]
print("mean cycles per tile and warp: wait %.0f, loads0 %.0f, math0 %.0f, loads1 %.0f, math1 %.0f; period %.0f" % (
    (e[:, :, 0, 1] - e[:, :, 0, 0]).mean(), (e[:, :, 0, 2] - e[:, :, 0, 1]).mean(), (e[:, :, 0, 3] - e[:, :, 0, 2]).mean(),
    (e[:, :, 1, 2] - e[:, :, 0, 3]).mean(), (e[:, :, 1, 3] - e[:, :, 1, 2]).mean(), (out[2, 59, 0, 0] - out[2, 20, 0, 0]) / 39.0))
