"""Timeline of CTA (0,0) of the tcgen05 kernel: clock64 stamps per warp, tile and half (nimb200_debug_ozaki_trace)."""
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
print("MMA warp (tile, half): wait-start, wait-done, issued+commit   [cycles relative to tile 20]")
for it in range(20, 25):
    for h in range(2):
        print(it, h, out[1, it, h, :3] - t0)
for w in (2, 9, 17):
    print("epilogue warp %d (tile, half): before t_full wait, after, loads done, math done" % w)
    for it in range(20, 23):
        for h in range(2):
            print(it, h, out[w, it, h] - t0)
print("mean cycles per half and warp: wait %.0f, loads %.0f, math %.0f" % ((out[2:, 20:40, :, 1] - out[2:, 20:40, :, 0]).mean(),
      (out[2:, 20:40, :, 2] - out[2:, 20:40, :, 1]).mean(), (out[2:, 20:40, :, 3] - out[2:, 20:40, :, 2]).mean()))
print("cycles per tile (warp 2, tiles 20..60): %.0f" % ((out[2, 60, 0, 0] - out[2, 20, 0, 0]) / 40.0))
