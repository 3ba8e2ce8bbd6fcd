"""ctypes mirror of include/nimb200.h: the node table that crosses the C ABI.

The table is the flat, chain-invariant description of a lowered model; it plays the role of the
objects `compileNimble` populates in the reference (nodeFunctionVector -> NodeInstruction lists,
P/R/types_nodeFxnVector.R:70-96 and P/inst/CppCode/accessorClasses.cpp:961-1058).
"""
import ctypes as C

import numpy as np

TABLE_VERSION = 1
MAX_PARAM = 3
MAX_BLOCK_D = 64

# distributions
DNORM, DGAMMA, DPOIS, DBINOM, DBETA, DLOGIS, DCAT, DEXP, DLNORM = 1, 2, 3, 4, 5, 6, 7, 8, 9
DIST_NAMES = {DNORM: "dnorm", DGAMMA: "dgamma", DPOIS: "dpois", DBINOM: "dbinom", DBETA: "dbeta",
              DLOGIS: "dlogis", DCAT: "dcat", DEXP: "dexp", DLNORM: "dlnorm"}
# ops
OP_CONST, OP_DATA, OP_CVAR, OP_CVARN, OP_INPROD, OP_DATAJ, OP_CVARJ = 1, 2, 3, 4, 5, 6, 7
OP_CVAR_IDX, OP_DATA_IDX = 48, 49
OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_POW = 16, 17, 18, 19, 20
OP_NEG, OP_EXP, OP_LOG, OP_SQRT, OP_ILOGIT, OP_LOGIT, OP_INV, OP_SQ, OP_ABS, OP_LOG1P, OP_ICLOGLOG, OP_STEP = (
    32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43)
IDX_I, IDX_DIV, IDX_MOD, IDX_IARR = 0, 1, 2, 3
FAM_GENERIC, FAM_BERN_LOGIT_INPROD, FAM_POIS_LOGLIN, FAM_NORM_GATHER, FAM_BERN_LOGIT_LIN = 0, 1, 2, 3, 4
SEL_ALL, SEL_SAME, SEL_STRIDE, SEL_SEG, SEL_CSR, SEL_RANGE = 0, 1, 2, 3, 4, 5
PARAM_RATE, PARAM_TAU = 16, 17
LINK_IDENTITY, LINK_ADDITIVE, LINK_MULTIPLICATIVE, LINK_LINEAR = 0, 1, 2, 3
SAMPLER_RW, SAMPLER_RW_BLOCK, SAMPLER_BINARY, SAMPLER_CONJ_GAMMA, SAMPLER_CONJ_NORMAL, SAMPLER_CONJ_BETA = 1, 2, 3, 4, 5, 6
SAMPLER_POSTERIOR_PREDICTIVE, SAMPLER_SLICE, SAMPLER_CATEGORICAL = 7, 8, 9
MAX_CATEGORIES = 64
SAMPLER_NAMES = {SAMPLER_RW: "RW", SAMPLER_RW_BLOCK: "RW_block", SAMPLER_BINARY: "binary",
                 SAMPLER_CONJ_GAMMA: "conjugate_dgamma", SAMPLER_CONJ_NORMAL: "conjugate_dnorm",
                 SAMPLER_CONJ_BETA: "conjugate_dbeta", SAMPLER_POSTERIOR_PREDICTIVE: "posterior_predictive",
                 SAMPLER_SLICE: "slice", SAMPLER_CATEGORICAL: "categorical"}


class Op(C.Structure):
    _fields_ = [("op", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("c", C.c_int32), ("k", C.c_double)]


class Decl(C.Structure):
    _fields_ = [("dist", C.c_int32), ("n", C.c_int32), ("value_kind", C.c_int32), ("value_ref", C.c_int32),
                ("nparam", C.c_int32), ("prog_off", C.c_int32 * MAX_PARAM), ("prog_len", C.c_int32 * MAX_PARAM),
                ("family", C.c_int32), ("fam_arg", C.c_int32 * 6)]


class Dep(C.Structure):
    _fields_ = [("decl", C.c_int32), ("sel", C.c_int32), ("a", C.c_int32), ("b", C.c_int32),
                ("stage", C.c_int32), ("link", C.c_int32), ("param", C.c_int32), ("pad", C.c_int32)]


class Sampler(C.Structure):
    _fields_ = [("type", C.c_int32), ("n_targets", C.c_int32), ("target_slot", C.c_int32), ("d", C.c_int32),
                ("dep_off", C.c_int32), ("n_dep", C.c_int32), ("log_scale", C.c_int32), ("reflective", C.c_int32),
                ("adaptive", C.c_int32), ("adapt_interval", C.c_int32), ("adapt_scale_only", C.c_int32),
                ("tries", C.c_int32), ("propcov_darr", C.c_int32), ("pad", C.c_int32),
                ("scale", C.c_double), ("adapt_factor_exponent", C.c_double),
                ("lower", C.c_double), ("upper", C.c_double)]


class Table(C.Structure):
    _fields_ = [("version", C.c_int32), ("n_slots", C.c_int32), ("n_darr", C.c_int32), ("n_iarr", C.c_int32),
                ("darr", C.POINTER(C.POINTER(C.c_double))), ("darr_len", C.POINTER(C.c_int64)),
                ("iarr", C.POINTER(C.POINTER(C.c_int32))), ("iarr_len", C.POINTER(C.c_int64)),
                ("n_ops", C.c_int32), ("n_decl", C.c_int32),
                ("ops", C.POINTER(Op)), ("decl", C.POINTER(Decl)),
                ("n_samplers", C.c_int32), ("n_deps", C.c_int32),
                ("samplers", C.POINTER(Sampler)), ("deps", C.POINTER(Dep)),
                ("n_monitors", C.c_int32), ("pad", C.c_int32),
                ("monitor_slots", C.POINTER(C.c_int32))]


class TableBuilder:
    """Accumulates arrays / programs / declarations / samplers and packs them into a `Table`.

    Keeps every numpy buffer alive for as long as the packed table is referenced."""

    def __init__(self):
        self.darr, self.iarr = [], []
        self.ops, self.decls, self.samplers, self.deps = [], [], [], []
        self.monitors = []
        self.n_slots = 0
        self._keep = []

    def add_darr(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
        self.darr.append(a)
        return len(self.darr) - 1

    def add_iarr(self, a):
        a = np.ascontiguousarray(a, dtype=np.int32).reshape(-1)
        self.iarr.append(a)
        return len(self.iarr) - 1

    def add_prog(self, ops):
        off = len(self.ops)
        self.ops.extend(ops)
        return off, len(ops)

    def pack(self):
        t = Table()
        t.version = TABLE_VERSION
        t.n_slots = self.n_slots
        nd, ni = len(self.darr), len(self.iarr)
        t.n_darr, t.n_iarr = nd, ni
        dptr = (C.POINTER(C.c_double) * max(nd, 1))()
        dlen = (C.c_int64 * max(nd, 1))()
        for i, a in enumerate(self.darr):
            dptr[i] = a.ctypes.data_as(C.POINTER(C.c_double))
            dlen[i] = a.size
        iptr = (C.POINTER(C.c_int32) * max(ni, 1))()
        ilen = (C.c_int64 * max(ni, 1))()
        for i, a in enumerate(self.iarr):
            iptr[i] = a.ctypes.data_as(C.POINTER(C.c_int32))
            ilen[i] = a.size
        t.darr, t.darr_len, t.iarr, t.iarr_len = dptr, dlen, iptr, ilen
        ops = (Op * max(len(self.ops), 1))()
        for i, (op, a, b, c, k) in enumerate(self.ops):
            ops[i] = Op(op, a, b, c, k)
        decl = (Decl * max(len(self.decls), 1))(*self.decls)
        samp = (Sampler * max(len(self.samplers), 1))(*self.samplers)
        deps = (Dep * max(len(self.deps), 1))(*self.deps)
        mon = (C.c_int32 * max(len(self.monitors), 1))(*self.monitors)
        t.n_ops, t.n_decl, t.ops, t.decl = len(self.ops), len(self.decls), ops, decl
        t.n_samplers, t.n_deps, t.samplers, t.deps = len(self.samplers), len(self.deps), samp, deps
        t.n_monitors, t.monitor_slots = len(self.monitors), mon
        self._keep = [dptr, dlen, iptr, ilen, ops, decl, samp, deps, mon]
        return t
