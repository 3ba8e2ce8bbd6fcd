import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nimble_b200 import _capi
b = _capi.default_backend()
print("dfma, dmma peaks", b.fp64_peaks(0))
f = b.lib.nimb200_measure_fp64_mixed
f.argtypes = [C.c_int, C.POINTER(C.c_double)]
t = C.c_double()
print("rc", f(0, C.byref(t)), "mixed TFLOP/s", t.value)
