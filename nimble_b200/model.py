"""nimbleModel for the supported sampling path: BUGS code -> lowered declarations.

Host-side mirror of what the reference's model-definition layer hands to the compiled MCMC:
declarations with their unrolled index ranges (declInfo, P/R/BUGS_BUGSdecl.R:111-170,
P/R/BUGS_modelDef.R:150-226), links turned into inverse-link expressions (BUGS_modelDef.R:197,
P/R/BUGS_mathCompatibility.R:3-12), reparameterisation to the canonical ("Rdist") parameters
(P/R/distributions_inputList.R), and the dependency structure getDependencies answers
(P/R/BUGS_getDependencies.R).  Deterministic nodes are inlined into the parameter expressions of the
stochastic declarations that read them, so only stochastic nodes exist in the lowered model.

Lowered expression trees are tuples:
  ('const', v) ('data', values[n]) ('cvar', var, pos) with pos an int or an int array[n] of element positions
  ('inprod', X[n,p], var, lo, p) ('un', op, a) ('bin', op, a, b) ('dataj', M[n,K]) ('cvarj', var, lo)
"""
import numpy as np

from . import bugs
from . import table as T

_UNARY = {"exp": T.OP_EXP, "log": T.OP_LOG, "sqrt": T.OP_SQRT, "ilogit": T.OP_ILOGIT, "expit": T.OP_ILOGIT,
          "logit": T.OP_LOGIT, "abs": T.OP_ABS, "log1p": T.OP_LOG1P, "icloglog": T.OP_ICLOGLOG, "step": T.OP_STEP}
_LINK_INVERSE = {"logit": T.OP_ILOGIT, "log": T.OP_EXP, "cloglog": T.OP_ICLOGLOG}  # BUGS_mathCompatibility.R:3-6
_BIN = {"+": T.OP_ADD, "-": T.OP_SUB, "*": T.OP_MUL, "/": T.OP_DIV, "^": T.OP_POW}

# BUGS distribution signatures -> canonical parameter trees.  distributions_inputList.R:20-149
_DIST_ARGS = {
    "dnorm": ["mean", "tau"], "dgamma": ["shape", "rate"], "dpois": ["lambda"], "dbern": ["prob"],
    "dbin": ["prob", "size"], "dbinom": ["prob", "size"], "dbeta": ["shape1", "shape2"],
    "dlogis": ["location", "rate"], "dexp": ["rate"], "dcat": ["prob"], "dlnorm": ["meanlog", "taulog"],
}


class ModelError(ValueError):
    pass


class Declaration:
    """One stochastic BUGS declaration over its flattened loop grid (n node instances)."""

    def __init__(self):
        self.var = None          # LHS variable name
        self.pos = None          # int array[n]: element position (column-major) of node i inside var
        self.n = 0
        self.dist = None         # T.D*
        self.dist_name = None
        self.params = []         # canonical parameter trees
        self.bugs_params = {}    # as-declared parameter trees by BUGS name (conjugacy checks)
        self.is_data = False
        self.K = 0               # dcat categories
        self.loop_extents = []
        self.index = -1          # position in Model.decls

    def reads(self):
        out = []
        for p in self.params:
            _collect_reads(p, out)
        return out


def _collect_reads(tree, out):
    k = tree[0]
    if k == "cvar":
        out.append((tree[1], tree[2]))
    elif k == "inprod":
        out.append((tree[2], ("block", tree[3], tree[4])))
    elif k == "cvarj":
        out.append((tree[1], ("block", tree[2], tree[3])))
    elif k == "dyn":  # ("dyn", latent?, name | array, base, K, index tree): any element may be read, plus whatever the index reads
        if tree[1]:
            out.append((tree[2], ("block", tree[3], tree[4])))
        _collect_reads(tree[5], out)
    elif k == "un":
        _collect_reads(tree[2], out)
    elif k == "bin":
        _collect_reads(tree[2], out)
        _collect_reads(tree[3], out)


def tree_reads_var(tree, var):
    out = []
    _collect_reads(tree, out)
    return any(v == var for v, _ in out)


class Model:
    """`nimbleModel(code, constants, data, inits)` restricted to the supported distribution set."""

    def __init__(self, code, constants=None, data=None, inits=None, name="model"):
        self.name = name
        self.constants = {k: np.asarray(v, dtype=np.float64) for k, v in (constants or {}).items()}
        self.data = {k: np.asarray(v, dtype=np.float64) for k, v in (data or {}).items()}
        self.inits = dict(inits or {})
        stmts = bugs.parse(code) if isinstance(code, str) else code
        self._raw = []  # (kind, loops, stmt)
        self._flatten(stmts, [])
        self._determ = {}
        for kind, loops, st in self._raw:
            if kind == "determ":
                lhs = st[2]
                if lhs[0] != "var":
                    raise ModelError("unsupported left-hand side")
                self._determ.setdefault(lhs[1], []).append((loops, st))
        # latent variable shapes: from the LHS index ranges of their stochastic declarations
        self.var_shape, self.var_slot = {}, {}
        stoch = [(loops, st) for kind, loops, st in self._raw if kind == "stoch"]
        lhs_info = []
        for loops, st in stoch:
            env, n = self._grid(loops)
            lhs = st[1]
            idx = [self._index_array(e, env, n) for e in lhs[2]]
            lhs_info.append((env, n, idx))
            name_ = lhs[1]
            if name_ in self.data:
                continue
            shape = [int(np.max(a)) for a in idx] if idx else []
            old = self.var_shape.get(name_)
            if old is not None and len(old) != len(shape):
                raise ModelError("inconsistent dimensions for %s" % name_)
            self.var_shape[name_] = [max(a, b) for a, b in zip(old, shape)] if old else shape
        slot = 0
        for name_, shape in self.var_shape.items():
            self.var_slot[name_] = slot
            slot += int(np.prod(shape)) if shape else 1
        self.n_slots = slot
        self.decls = []
        for (loops, st), (env, n, idx) in zip(stoch, lhs_info):
            self.decls.append(self._lower_stoch(loops, st, env, n, idx))
        for i, d in enumerate(self.decls):
            d.index = i
        self._check_coverage()

    # ------------------------------------------------------------------ structure
    def _flatten(self, stmts, loops):
        for st in stmts:
            if st[0] == "for":
                lo = int(round(self._const_eval(st[2])))
                hi = int(round(self._const_eval(st[3])))
                self._flatten(st[4], loops + [(st[1], lo, hi)])
            else:
                self._raw.append((st[0], loops, st))

    def _const_eval(self, e):
        k = e[0]
        if k == "num":
            return e[1]
        if k == "neg":
            return -self._const_eval(e[1])
        if k == "bin":
            a, b = self._const_eval(e[2]), self._const_eval(e[3])
            return {"+": a + b, "-": a - b, "*": a * b, "/": a / b, "^": a ** b}[e[1]]
        if k == "var" and not e[2] and e[1] in self.constants:
            return float(self.constants[e[1]])
        if k == "call" and e[1] == "length":
            v = e[2][0]
            return float(np.size(self.constants.get(v[1], self.data.get(v[1]))))
        raise ModelError("loop bound is not a constant expression: %r" % (e,))

    def _grid(self, loops):
        """loop variable -> value per flattened node instance (outer loop slowest)."""
        ext = [hi - lo + 1 for _, lo, hi in loops]
        n = int(np.prod(ext)) if ext else 1
        env = {}
        rep_inner = n
        for (v, lo, hi), e in zip(loops, ext):
            rep_inner //= e
            env[v] = (np.arange(n) // rep_inner) % e + lo
        env["__ext__"] = ext
        return env, n

    def _index_array(self, e, env, n):
        """Evaluate an index expression to an int array[n] (1-based, as written in the BUGS code)."""
        k = e[0]
        if k == "num":
            return np.full(n, int(round(e[1])), dtype=np.int64)
        if k == "neg":
            return -self._index_array(e[1], env, n)
        if k == "bin":
            a, b = self._index_array(e[2], env, n), self._index_array(e[3], env, n)
            return {"+": a + b, "-": a - b, "*": a * b, "/": a // b}[e[1]]
        if k == "var":
            name_ = e[1]
            if not e[2]:
                if name_ in env:
                    return env[name_].astype(np.int64)
                if name_ in self.constants:
                    return np.full(n, int(round(float(self.constants[name_]))), dtype=np.int64)
            src = self.constants.get(name_, self.data.get(name_))
            if src is not None and e[2]:
                sub = [self._index_array(x, env, n) for x in e[2]]
                return np.rint(self._gather(src, sub)).astype(np.int64)
        raise ModelError("index expression must be built from loop variables and constants: %r" % (e,))

    @staticmethod
    def _gather(src, sub):
        src = np.asarray(src)
        if src.ndim != len(sub):
            raise ModelError("wrong number of indices")
        return src[tuple(s - 1 for s in sub)]

    def _positions(self, name_, idx, n):
        shape = self.var_shape[name_]
        if not shape:
            return 0
        if len(idx) != len(shape):
            raise ModelError("wrong number of indices for %s" % name_)
        pos, mult = np.zeros(n, dtype=np.int64), 1
        for a, s in zip(idx, shape):  # column-major, like R
            if np.any(a < 1) or np.any(a > s):
                raise ModelError("index out of range for %s" % name_)
            pos += (a - 1) * mult
            mult *= s
        return pos

    # ------------------------------------------------------------------ expression lowering
    def _lower(self, e, env, n, depth=0):
        if depth > 50:
            raise ModelError("deterministic nodes nest too deeply (cycle?)")
        k = e[0]
        if k == "num":
            return ("const", float(e[1]))
        if k == "neg":
            return ("un", T.OP_NEG, self._lower(e[1], env, n, depth))
        if k == "sub":
            # (X[i, 1:p] %*% beta[1:p])[1, 1]: the single element of a row-by-vector product (test-mcmc.R:1793-1803)
            inner = e[1]
            if inner[0] == "bin" and inner[1] == "%*%" and all(x == ("num", 1.0) for x in e[2]):
                return self._lower_inprod(inner[2], inner[3], env, n)
            raise ModelError("indexing an expression is supported only as (row %*% vector)[1, 1]")
        if k == "bin" and e[1] == "%*%":
            return self._lower_inprod(e[2], e[3], env, n)
        if k == "bin":
            a, b = self._lower(e[2], env, n, depth), self._lower(e[3], env, n, depth)
            if e[1] == "^" and b == ("const", 2.0):
                return ("un", T.OP_SQ, a)
            return ("bin", _BIN[e[1]], a, b)
        if k == "call":
            f, args = e[1], e[2]
            if f in _UNARY and len(args) == 1:
                return ("un", _UNARY[f], self._lower(args[0], env, n, depth))
            if f == "pow" and len(args) == 2:
                return ("bin", T.OP_POW, self._lower(args[0], env, n, depth), self._lower(args[1], env, n, depth))
            if f == "inprod" and len(args) == 2:
                return self._lower_inprod(args[0], args[1], env, n)
            if f == "sum" and len(args) == 1 and args[0][0] == "bin" and args[0][1] == "*":
                # sum(beta[1:p] * X[i, 1:p]) is the same node function as inprod() (MCMC_conjugacy.R:1478-1486)
                return self._lower_inprod(args[0][2], args[0][3], env, n)
            raise ModelError("unsupported function in model code: %s()" % f)
        if k == "var":
            name_, idx = e[1], e[2]
            if name_ in env and not idx:
                return ("data", env[name_].astype(np.float64))
            dyn = self._dynamic_index(name_, idx, env, n, depth)
            if dyn is not None:
                return dyn
            if name_ in self.var_slot:
                ia = [self._index_array(x, env, n) for x in idx]
                pos = self._positions(name_, ia, n)
                if not np.isscalar(pos) and pos.size and np.all(pos == pos[0]):
                    pos = int(pos[0])
                return ("cvar", name_, pos)
            if name_ in self._determ:
                return self._inline_determ(name_, idx, env, n, depth)
            src = self.constants.get(name_, self.data.get(name_))
            if src is not None:
                if not idx:
                    if src.ndim != 0 and src.size != 1:
                        raise ModelError("%s used without an index" % name_)
                    return ("const", float(src.reshape(-1)[0]))
                vals = self._gather(src, [self._index_array(x, env, n) for x in idx]).astype(np.float64)
                if vals.size and np.all(vals == vals[0]):
                    return ("const", float(vals[0]))
                return ("data", vals)
            raise ModelError("variable %s is neither data, constant, nor defined in the model" % name_)
        raise ModelError("cannot lower expression %r" % (e,))

    def _reads_latent(self, e):
        """Does an index expression read a latent variable or a deterministic node (i.e. is it not static)?"""
        k = e[0]
        if k == "var":
            if e[1] in self.var_slot or e[1] in self._determ:
                return True
            return any(self._reads_latent(x) for x in e[2] if x[0] != "range")
        if k == "neg":
            return self._reads_latent(e[1])
        if k == "bin":
            return self._reads_latent(e[2]) or self._reads_latent(e[3])
        if k == "call":
            return any(self._reads_latent(x) for x in e[2])
        return False

    def _dynamic_index(self, name_, idx, env, n, depth):
        """mu[z[i]]: a vector (latent, constant or data) indexed by an expression that reads the chain state.  Returns
        ("dyn", latent?, name | values, base position, K, index tree) or None when the index is static."""
        if len(idx) != 1 or idx[0][0] == "range" or not self._reads_latent(idx[0]):
            if any(x[0] != "range" and self._reads_latent(x) for x in idx):
                raise ModelError("dynamic indexing is supported for vectors only: %s" % name_)
            return None
        itree = self._lower(idx[0], env, n, depth + 1)
        if name_ in self.var_slot:
            shape = self.var_shape[name_]
            if len(shape) != 1:
                raise ModelError("dynamic indexing is supported for vectors only: %s" % name_)
            base = self._positions(name_, [np.full(1, 1, dtype=np.int64)], 1)
            return ("dyn", True, name_, int(base[0]), int(shape[0]), itree)
        src = self.constants.get(name_, self.data.get(name_))
        if src is None or src.ndim != 1:
            raise ModelError("dynamic indexing needs a latent, constant or data vector: %s" % name_)
        return ("dyn", False, np.ascontiguousarray(src, dtype=np.float64), 0, int(src.size), itree)

    def _inline_determ(self, name_, idx, env, n, depth):
        """Replace a read of a deterministic node by its defining expression (links inverted)."""
        use = [self._index_array(x, env, n) for x in idx]
        for loops, st in self._determ[name_]:
            link, lhs, rhs = st[1], st[2], st[3]
            if len(lhs[2]) != len(use):
                continue
            new_env, ok = {"__ext__": env.get("__ext__", [])}, True
            loopvars = {v: (lo, hi) for v, lo, hi in loops}
            for li, u in zip(lhs[2], use):
                if li[0] == "var" and not li[2] and li[1] in loopvars:
                    lo, hi = loopvars[li[1]]
                    if np.any(u < lo) or np.any(u > hi):
                        ok = False
                        break
                    new_env[li[1]] = u
                else:
                    try:
                        c = self._index_array(li, {}, n)
                    except ModelError:
                        ok = False
                        break
                    if np.any(c != u):
                        ok = False
                        break
            if not ok or any(v not in new_env for v in loopvars):
                continue
            tree = self._lower(rhs, new_env, n, depth + 1)
            if link is not None:
                if link not in _LINK_INVERSE:
                    raise ModelError("unsupported link function %s()" % link)
                tree = ("un", _LINK_INVERSE[link], tree)
            return tree
        raise ModelError("no deterministic declaration covers this use of %s" % name_)

    def _lower_inprod(self, a, b, env, n):
        def is_latent_slice(x):
            return x[0] == "var" and x[1] in self.var_slot and len(x[2]) == 1 and x[2][0][0] == "range"
        if is_latent_slice(a) and not is_latent_slice(b):
            a, b = b, a
        if not is_latent_slice(b) or a[0] != "var":
            raise ModelError("inprod() is supported as inprod(<constant/data row slice>, <latent vector slice>)")
        lo = int(round(self._const_eval(b[2][0][1])))
        hi = int(round(self._const_eval(b[2][0][2])))
        p = hi - lo + 1
        src = self.constants.get(a[1], self.data.get(a[1]))
        if src is None:
            raise ModelError("inprod(): %s must be data or a constant" % a[1])
        rows = None
        cols = None
        for ax, ie in enumerate(a[2]):
            if ie[0] == "range":
                c0 = int(round(self._const_eval(ie[1])))
                c1 = int(round(self._const_eval(ie[2])))
                cols = (ax, c0, c1)
            else:
                rows = (ax, self._index_array(ie, env, n))
        if cols is None or cols[2] - cols[1] + 1 != p:
            raise ModelError("inprod(): slice lengths differ")
        if src.ndim == 1:
            X = np.broadcast_to(src[cols[1] - 1:cols[2]], (n, p))
        elif cols[0] == 1:
            X = src[rows[1] - 1, cols[1] - 1:cols[2]]
        else:
            X = src[cols[1] - 1:cols[2], rows[1] - 1].T
        base = self._positions(b[1], [np.full(1, lo, dtype=np.int64)], 1)
        return ("inprod", np.ascontiguousarray(X, dtype=np.float64), b[1], int(base[0]), p)

    # ------------------------------------------------------------------ stochastic declarations
    def _lower_stoch(self, loops, st, env, n, idx):
        lhs, dist, args, kw = st[1], st[2], st[3], st[4]
        d = Declaration()
        d.var, d.n, d.dist_name = lhs[1], n, dist
        d.loop_extents = env["__ext__"]
        d.is_data = d.var in self.data
        if d.is_data:
            d.values = self._gather(self.data[d.var], idx).astype(np.float64) if idx else \
                np.full(n, float(self.data[d.var]))
            d.pos = None
        else:
            d.pos = self._positions(d.var, idx, n)
            if np.isscalar(d.pos):
                d.pos = np.zeros(1, dtype=np.int64)
        if dist not in _DIST_ARGS:
            raise ModelError("distribution %s is outside the supported set %s" % (dist, sorted(_DIST_ARGS)))
        names = _DIST_ARGS[dist]
        given = {}
        for name_, a in zip(names, args):
            given[name_] = a
        given.update(kw)
        if dist == "dcat":
            pe = given["prob"]
            if pe[0] != "var" or not pe[2] or pe[2][-1][0] != "range":
                raise ModelError("dcat(prob): prob must be a vector slice such as p[1:K] or P[i, 1:K]")
            k0 = int(round(self._const_eval(pe[2][-1][1])))
            k1 = int(round(self._const_eval(pe[2][-1][2])))
            d.K = k1 - k0 + 1
            if pe[1] in self.var_slot:
                base = self._positions(pe[1], [np.full(1, k0, dtype=np.int64)], 1)
                tree = ("cvarj", pe[1], int(base[0]), d.K)
            else:
                src = self.constants.get(pe[1], self.data.get(pe[1]))
                if src is None:
                    raise ModelError("dcat(prob): unknown variable %s" % pe[1])
                if src.ndim == 1:
                    M = np.broadcast_to(src[k0 - 1:k1], (n, d.K))
                else:
                    rows = self._index_array(pe[2][0], env, n)
                    M = src[rows - 1, k0 - 1:k1]
                tree = ("dataj", np.ascontiguousarray(M, dtype=np.float64))
            d.dist, d.params, d.bugs_params = T.DCAT, [tree], {"prob": tree}
            return d
        low = {k: self._lower(v, env, n) for k, v in given.items()}
        d.bugs_params = low
        inv = lambda t: ("un", T.OP_INV, t)
        if dist == "dnorm":
            if "sd" in low:
                sd = low["sd"]
            elif "var" in low:
                sd = ("un", T.OP_SQRT, low["var"])
            else:
                sd = inv(("un", T.OP_SQRT, low["tau"]))  # sd = 1/sqrt(tau), inputList.R:145
            d.dist, d.params = T.DNORM, [low["mean"], sd]
        elif dist == "dgamma":
            scale = low["scale"] if "scale" in low else inv(low["rate"])  # scale = 1/rate, inputList.R:104
            d.dist, d.params = T.DGAMMA, [low["shape"], scale]
        elif dist == "dpois":
            d.dist, d.params = T.DPOIS, [low["lambda"]]
        elif dist == "dbern":
            d.dist, d.params = T.DBINOM, [("const", 1.0), low["prob"]]  # dbinom(size = 1, prob), inputList.R:21
        elif dist in ("dbin", "dbinom"):
            d.dist, d.params = T.DBINOM, [low["size"], low["prob"]]
        elif dist == "dbeta":
            d.dist, d.params = T.DBETA, [low["shape1"], low["shape2"]]
        elif dist == "dlogis":
            scale = low["scale"] if "scale" in low else inv(low["rate"])
            d.dist, d.params = T.DLOGIS, [low["location"], scale]
        elif dist == "dexp":
            rate = low["rate"] if "rate" in low else inv(low["scale"])
            d.dist, d.params = T.DEXP, [rate]
        elif dist == "dlnorm":   # dlnorm(meanlog, sdlog = 1/sqrt(taulog)) | sqrt(varlog), inputList.R:131-137
            if "sdlog" in low:
                sdl = low["sdlog"]
            elif "varlog" in low:
                sdl = ("un", T.OP_SQRT, low["varlog"])
            else:
                sdl = inv(("un", T.OP_SQRT, low["taulog"]))
            d.dist, d.params = T.DLNORM, [low["meanlog"], sdl]
        return d

    def _check_coverage(self):
        for name_, shape in self.var_shape.items():
            size = int(np.prod(shape)) if shape else 1
            seen = np.zeros(size, dtype=np.int64)
            for d in self.decls:
                if d.var == name_ and not d.is_data:
                    np.add.at(seen, d.pos, 1)
            if np.any(seen != 1):
                raise ModelError("every element of %s must be declared exactly once" % name_)

    # ------------------------------------------------------------------ queries (model$... in the reference)
    def getNodeNames(self, stochOnly=True, includeData=False):
        out = []
        for d in self.decls:
            if d.is_data and not includeData:
                continue
            out.extend(self.element_names(d.var, d.pos if not d.is_data else np.arange(d.n)))
        return out

    def element_names(self, var, pos):
        """expandMVNames-style names, column-major (P/R/types_util.R:155-179)."""
        shape = self.var_shape.get(var, [])
        if not shape:
            return [var]
        names = []
        for p in np.asarray(pos).reshape(-1):
            sub, r = [], int(p)
            for s in shape:
                sub.append(r % s + 1)
                r //= s
            names.append("%s[%s]" % (var, ", ".join(str(x) for x in sub)))
        return names

    def decl_of(self, var):
        return [d for d in self.decls if d.var == var and not d.is_data]

    def dependents(self, var):
        """Declarations (other than var's own) whose parameters read `var`."""
        return [d for d in self.decls if any(v == var for v, _ in d.reads())]

    def top_level_vars(self):
        """Variables with no stochastic parents: default monitors (MCMC_configuration.R `monitors`)."""
        out = []
        for name_ in self.var_slot:
            if all(not d.reads() for d in self.decl_of(name_)):
                out.append(name_)
        return out

    def topo_vars(self):
        """Latent variables ordered parents-first (ties: declaration order)."""
        order, done = [], set()
        names = list(self.var_slot)
        while len(order) < len(names):
            progressed = False
            for name_ in names:
                if name_ in done:
                    continue
                parents = {v for d in self.decl_of(name_) for v, _ in d.reads()}
                if all(p in done or p == name_ for p in parents):
                    order.append(name_)
                    done.add(name_)
                    progressed = True
            if not progressed:
                raise ModelError("cycle in the model graph")
        return order

    def init_vector(self, inits=None):
        """Flatten an inits list into the per-chain state vector (model$setInits)."""
        vals = dict(self.inits)
        vals.update(inits or {})
        out = np.full(self.n_slots, np.nan)
        for name_, slot in self.var_slot.items():
            shape = self.var_shape[name_]
            size = int(np.prod(shape)) if shape else 1
            if name_ in vals:
                v = np.asarray(vals[name_], dtype=np.float64)
                out[slot:slot + size] = v.reshape(-1, order="F") if v.size == size else np.broadcast_to(v, size)
        # missing values stay NaN (NA): initializeModel simulates them from their distributions before sampling starts
        # (P/R/initializeModel.R:78-139; CompiledMCMC.run_chains -> nimb200_initialize_model)
        return out

    def init_matrix(self, inits_list):
        """init_vector for a list of per-chain inits at once ([chains][slots]): one assignment per variable when every
        chain names the same variables (the usual case), instead of one Python pass over the variables per chain."""
        n = len(inits_list)
        dicts = [d or {} for d in inits_list]
        keys = set(dicts[0])
        if any(set(d) != keys for d in dicts):
            return np.stack([self.init_vector(d) for d in dicts])
        out = np.tile(self.init_vector(None), (n, 1))
        for name_ in keys:
            if name_ not in self.var_slot:
                continue
            slot, shape = self.var_slot[name_], self.var_shape[name_]
            size = int(np.prod(shape)) if shape else 1
            try:
                vals = np.asarray([np.asarray(d[name_], dtype=np.float64) for d in dicts])
            except ValueError:   # ragged: chain by chain
                return np.stack([self.init_vector(d) for d in dicts])
            if vals.size == n * size:
                if vals.ndim > 2:   # arrays are stored column-major, as init_vector does
                    vals = np.stack([v.reshape(-1, order="F") for v in vals])
                out[:, slot:slot + size] = vals.reshape(n, size)
            elif vals.size == n:    # a scalar broadcast over the variable
                out[:, slot:slot + size] = vals.reshape(n, 1)
            else:
                return np.stack([self.init_vector(d) for d in dicts])
        return out

    def simulable(self, var):
        """Can missing values of `var` be simulated on the device (a variate generator exists for its distribution)?"""
        return all(d.dist in (T.DNORM, T.DGAMMA, T.DBETA, T.DEXP, T.DLOGIS, T.DPOIS, T.DBINOM, T.DLNORM) for d in self.decl_of(var))


def nimbleCode(text):
    return text


def nimbleModel(code, constants=None, data=None, inits=None, name="model", **_ignored):
    return Model(code, constants, data, inits, name)
