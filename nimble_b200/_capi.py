"""ctypes binding of the C ABI in include/nimb200.h (libnimb200.so, the sm_100a library).

There is no CPU fallback: if the shared library is missing or no CUDA device is present, every call
raises.  (The same `Backend` class can bind a library that exports the identical entry points under
another prefix; the test suite uses that to drive the CPU oracle through the very same host code.)
"""
import ctypes as C
import os

import numpy as np

from . import table as T

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnimb200.so")

_dp = C.POINTER(C.c_double)


def _ptr(a):
    return a.ctypes.data_as(_dp)


class BackendError(RuntimeError):
    pass


class Backend:
    """Thin object wrapper over the C entry points `<prefix>model_create`, `<prefix>mcmc_run`, ..."""

    def __init__(self, path=LIB_PATH, prefix="nimb200_", has_device=True, name="b200"):
        if not os.path.exists(path):
            raise BackendError("%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)" % path)
        self.lib = C.CDLL(path)
        self.prefix, self.has_device, self.name, self.path = prefix, has_device, name, path
        L, P = self.lib, prefix
        vp = C.c_void_p

        def fn(name_, argtypes, restype=C.c_int):
            f = getattr(L, P + name_)
            f.argtypes, f.restype = argtypes, restype
            return f
        self._err = fn("last_error", [], C.c_char_p)
        if has_device:
            self._create = fn("model_create", [C.POINTER(T.Table), C.c_int32, C.c_int32, C.c_int32, C.POINTER(vp)])
        else:
            self._create = fn("model_create", [C.POINTER(T.Table), C.c_int32, C.c_int32, C.POINTER(vp)])
        self._free = fn("model_free", [vp])
        self._set = fn("set_values", [vp, _dp])
        self._get = fn("get_values", [vp, _dp])
        self._calc = fn("calculate", [vp, C.c_int32, C.c_int32, C.c_int32, _dp])
        self._calc_nodes = fn("calculate_nodes", [vp, C.c_int32, C.c_int32, C.c_int32, _dp])
        self._calc_diff = fn("calculateDiff", [vp, C.c_int32, _dp])
        self._get_lp = fn("getLogProb", [vp, C.c_int32, _dp])
        self._run = fn("mcmc_run", [vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_uint64])
        self._samples = fn("mcmc_samples", [vp, _dp, C.POINTER(C.c_int64)])
        self._sstate = fn("sampler_state", [vp, C.c_int32, C.c_int32, _dp, C.c_int64])
        self._waic_enable = fn("waic_enable", [vp, C.c_int32])
        self._waic_get = fn("waic_get", [vp, _dp])
        self._init_model = fn("initialize_model", [vp, C.POINTER(C.c_int32), C.c_int32, C.c_uint64])
        self._draws = fn("draws", [C.c_uint64, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_double,
                                   C.c_int32, _dp])
        if has_device:
            self._timing = fn("mcmc_last_timing", [vp, _dp, C.POINTER(C.c_int64)])
            self._waic_partial = fn("waic_partial", [vp, _dp])
            self._summary = fn("samples_summary", [vp, C.c_int32, C.c_int32, _dp])
            self._hist_enable = fn("sampler_history_enable", [vp, C.c_int32])
            self._hist = fn("sampler_history", [vp, C.c_int32, C.c_int32, _dp, C.c_int64, C.POINTER(C.c_int32)])
            self._set_timing = fn("mcmc_set_timing", [vp, C.c_int32])
            self._sampler_times = fn("mcmc_sampler_times", [vp, _dp, C.c_int32])
            self._time_calc = fn("time_calculate", [vp, C.c_int32, C.c_int32, _dp])
            self._peaks = fn("measure_fp64_peaks", [C.c_int, _dp, _dp])
            self._mixed = fn("measure_fp64_mixed", [C.c_int, _dp])
            self._issue_mix = fn("measure_fp64_issue_mix", [C.c_int, _dp])
            self._nccl_id = fn("nccl_unique_id", [C.c_void_p])
            self._shard = fn("enable_data_sharding", [vp, C.c_int32, C.c_int32, C.c_void_p, C.c_int32])
            self._peer_export = fn("peer_export", [vp, C.c_int32, C.c_void_p])
            self._shard_peer = fn("enable_data_sharding_peer", [vp, C.c_int32, C.c_int32, C.c_void_p, C.c_int32])
            self._allreduces = fn("mcmc_last_allreduces", [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int32)])
            self._time_ar = fn("time_allreduce", [vp, C.c_int32, C.c_int32, _dp])

    def check(self, rc):
        if rc != 0:
            msg = self._err()
            raise BackendError((msg or b"unknown error").decode())

    def create(self, packed_table, nchains, first_chain=0, device=0):
        h = C.c_void_p()
        if self.has_device:
            self.check(self._create(C.byref(packed_table), nchains, first_chain, device, C.byref(h)))
        else:
            self.check(self._create(C.byref(packed_table), nchains, first_chain, C.byref(h)))
        return h

    def nccl_unique_id(self):
        """128-byte ncclUniqueId (bytes); produce on one rank, hand to all ranks of the data-sharded group."""
        buf = C.create_string_buffer(128)
        self.check(self._nccl_id(buf))
        return buf.raw

    def fp64_mixed_peak(self, device=0):
        a = C.c_double()
        self.check(self._mixed(device, C.byref(a)))
        return a.value

    def fp64_issue_mix(self, device=0):
        """DFMA TFLOP/s of a loop with 0, 4, 8, 16 integer instructions per 8 DFMAs."""
        out = np.zeros(4)
        self.check(self._issue_mix(device, _ptr(out)))
        return out

    def fp64_peaks(self, device=0):
        a, b = C.c_double(), C.c_double()
        self.check(self._peaks(device, C.byref(a), C.byref(b)))
        return a.value, b.value

    def draws(self, seed, chain, it, sampler, target, kind, a, n):
        out = np.empty(n)
        self.check(self._draws(seed, chain, it, sampler, target, kind, a, n, _ptr(out)))
        return out


class DeviceModel:
    """One compiled model instance: nchains chains resident on one device (or in the oracle)."""

    def __init__(self, backend, builder, nchains, first_chain=0, device=0):
        self.backend, self.builder, self.C = backend, builder, nchains
        self.packed = builder.pack()
        self.n_slots, self.n_mon = builder.n_slots, len(builder.monitors)
        self.h = backend.create(self.packed, nchains, first_chain, device)

    def close(self):
        if self.h is not None and self.h.value:
            self.backend._free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_values(self, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        assert v.shape == (self.C, self.n_slots), v.shape
        self.backend.check(self.backend._set(self.h, _ptr(v)))

    def get_values(self):
        out = np.empty((self.C, self.n_slots))
        self.backend.check(self.backend._get(self.h, _ptr(out)))
        return out

    def calculate(self, decl, lo=0, hi=None):
        hi = self.builder.decls[decl].n if hi is None else hi
        out = np.empty(self.C)
        self.backend.check(self.backend._calc(self.h, decl, lo, hi, _ptr(out)))
        return out

    def calculate_nodes(self, decl, lo=0, hi=None):
        hi = self.builder.decls[decl].n if hi is None else hi
        out = np.empty((hi - lo, self.C))
        self.backend.check(self.backend._calc_nodes(self.h, decl, lo, hi, _ptr(out)))
        return out

    def calculateDiff(self, decl):
        out = np.empty(self.C)
        self.backend.check(self.backend._calc_diff(self.h, decl, _ptr(out)))
        return out

    def getLogProb(self, decl):
        out = np.empty(self.C)
        self.backend.check(self.backend._get_lp(self.h, decl, _ptr(out)))
        return out

    def set_timing(self, enable=True):
        """run(time = TRUE): per-sampler CUDA-event timing (profiling mode, device back end only)."""
        if not self.backend.has_device:
            raise BackendError("per-sampler timing is a device feature")
        self.backend.check(self.backend._set_timing(self.h, int(bool(enable))))

    def sampler_times(self):
        """Milliseconds per table sampler since timing was enabled."""
        out = np.zeros(len(self.builder.samplers))
        self.backend.check(self.backend._sampler_times(self.h, _ptr(out), int(out.size)))
        return out

    def waic_enable(self, enable=True):
        self.backend.check(self.backend._waic_enable(self.h, int(bool(enable))))

    def waic_partial(self):
        """(log-sum part, sum of squares, nodes, count) over the data nodes this rank accounts for (data-sharded models)."""
        out = np.zeros(4)
        self.backend.check(self.backend._waic_partial(self.h, _ptr(out)))
        return out

    def waic_get(self):
        """{'WAIC', 'lppd', 'pWAIC', 'n'} of the online conditional WAIC accumulated so far."""
        out = np.empty(4)
        self.backend.check(self.backend._waic_get(self.h, _ptr(out)))
        return {"WAIC": float(out[0]), "lppd": float(out[1]), "pWAIC": float(out[2]), "n": int(out[3])}

    def initialize_model(self, decl_order, seed=0):
        """Simulate the non-data nodes whose value is NaN, declaration by declaration in `decl_order` (parents first)."""
        order = np.ascontiguousarray(decl_order, dtype=np.int32)
        self.backend.check(self.backend._init_model(self.h, order.ctypes.data_as(C.POINTER(C.c_int32)), int(order.size),
                                                    int(seed)))

    def run(self, niter, nburnin=0, thin=1, reset=True, seed=0):
        self.backend.check(self.backend._run(self.h, niter, nburnin, thin, 1 if reset else 0, seed))

    def samples(self):
        n = C.c_int64()
        self.backend.check(self.backend._samples(self.h, None, C.byref(n)))
        out = np.empty((self.C, n.value, self.n_mon))
        if out.size:
            self.backend.check(self.backend._samples(self.h, _ptr(out), C.byref(n)))
        return out

    def history_enable(self, max_adaptations=1024):
        self.backend.check(self.backend._hist_enable(self.h, int(max_adaptations)))

    def sampler_history(self, sampler, what):
        """what 0 scaleHistory / 1 acceptanceHistory: [adaptations][T][C]; 2 propCovHistory: [adaptations][d*d][C]."""
        n = C.c_int32()
        self.backend.check(self.backend._hist(self.h, sampler, what, None, 0, C.byref(n)))
        s = self.builder.samplers[sampler]
        Tn = 1 if s.type == T.SAMPLER_RW_BLOCK else s.n_targets
        inner = s.d * s.d if what == 2 else Tn
        out = np.empty((n.value, inner, self.C))
        if out.size:
            self.backend.check(self.backend._hist(self.h, sampler, what, _ptr(out), out.size, C.byref(n)))
        return out

    def samples_summary(self, chain_lo=0, chain_hi=None):
        """samplesSummary on the device: array [5][n_monitors] = Mean, Median, St.Dev., 95%CI_low, 95%CI_upp (chains pooled)."""
        out = np.empty((5, self.n_mon))
        self.backend.check(self.backend._summary(self.h, chain_lo, self.C if chain_hi is None else chain_hi, _ptr(out)))
        return out

    def sampler_state(self, sampler, what):
        s = self.builder.samplers[sampler]
        Tn = 1 if s.type == T.SAMPLER_RW_BLOCK else s.n_targets
        n = (s.d * s.d if what == 2 else Tn) * self.C
        out = np.empty(n)
        self.backend.check(self.backend._sstate(self.h, sampler, what, _ptr(out), n))
        return out.reshape(-1, self.C)

    def enable_data_sharding(self, nranks, rank, unique_id=None, min_nodes=65536):
        """Split large data declarations across the ranks of a group; unique_id=None keeps block sums as partial sums."""
        buf = C.create_string_buffer(unique_id, 128) if unique_id is not None else None
        self.backend.check(self.backend._shard(self.h, nranks, rank, buf, min_nodes))

    def peer_export(self, nranks):
        """Allocate this rank's exchange buffer for the device-side all-reduce; returns the 128-byte handle to gather."""
        buf = C.create_string_buffer(128)
        self.backend.check(self.backend._peer_export(self.h, nranks, buf))
        return buf.raw

    def enable_data_sharding_peer(self, nranks, rank, handles, min_nodes=65536):
        """handles: the nranks exported handles in rank order (this rank's own included)."""
        blob = b"".join(handles)
        assert len(blob) == 128 * nranks
        buf = C.create_string_buffer(blob, len(blob))
        self.backend.check(self.backend._shard_peer(self.h, nranks, rank, buf, min_nodes))

    def last_allreduces(self):
        n, g = C.c_int64(), C.c_int32()
        self.backend.check(self.backend._allreduces(self.h, C.byref(n), C.byref(g)))
        return n.value, bool(g.value)

    def time_allreduce(self, count, reps=100):
        ms = C.c_double()
        self.backend.check(self.backend._time_ar(self.h, count, reps, C.byref(ms)))
        return ms.value

    def time_calculate(self, decl, reps=5):
        ms = C.c_double()
        self.backend.check(self.backend._time_calc(self.h, decl, reps, C.byref(ms)))
        return ms.value

    def last_timing(self):
        ms, k = C.c_double(), C.c_int64()
        self.backend.check(self.backend._timing(self.h, C.byref(ms), C.byref(k)))
        return ms.value, k.value


_default = None


def default_backend():
    """The CUDA library.  Fails loudly when it is missing; nothing else is ever substituted."""
    global _default
    if _default is None:
        _default = Backend()
    return _default
