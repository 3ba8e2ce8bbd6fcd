"""Pageable vs registered host-to-device copy of a 400 MB numpy array (the C2 design matrix): python scripts/probe_upload.py"""
import ctypes as C, time
import numpy as np
import torch
rt = C.CDLL("libcudart.so.12")
torch.cuda.init(); torch.zeros(1, device="cuda")
n = 50_000_000
x = np.random.default_rng(0).standard_normal(n)
d = C.c_void_p()
assert rt.cudaMalloc(C.byref(d), C.c_size_t(n * 8)) == 0
def sync(): assert rt.cudaDeviceSynchronize() == 0
for rep in range(3):
    sync(); t0 = time.perf_counter()
    assert rt.cudaMemcpy(d, C.c_void_p(x.ctypes.data), C.c_size_t(n * 8), 1) == 0
    sync(); t1 = time.perf_counter()
    print("pageable cudaMemcpy      %.1f ms  (%.1f GB/s)" % ((t1 - t0) * 1e3, n * 8 / (t1 - t0) / 1e9))
for rep in range(3):
    sync(); t0 = time.perf_counter()
    assert rt.cudaHostRegister(C.c_void_p(x.ctypes.data), C.c_size_t(n * 8), 0) == 0
    t1 = time.perf_counter()
    assert rt.cudaMemcpy(d, C.c_void_p(x.ctypes.data), C.c_size_t(n * 8), 1) == 0
    sync(); t2 = time.perf_counter()
    assert rt.cudaHostUnregister(C.c_void_p(x.ctypes.data)) == 0
    t3 = time.perf_counter()
    print("register %.1f ms + copy %.1f ms (%.1f GB/s) + unregister %.1f ms = %.1f ms" %
          ((t1 - t0) * 1e3, (t2 - t1) * 1e3, n * 8 / (t2 - t1) / 1e9, (t3 - t2) * 1e3, (t3 - t0) * 1e3))
