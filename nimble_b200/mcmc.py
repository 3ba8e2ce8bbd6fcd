"""configureMCMC / buildMCMC / compileNimble / runMCMC for the B200 back end.

Same names, argument meaning and error behaviour as the reference's R functions:
  configureMCMC   P/R/MCMC_configuration.R:1362-1396 (default sampler assignment :340-446)
  buildMCMC       P/R/MCMC_build.R:101-215 (setup) / :217-321 (run)
  compileNimble   P/R/nimbleProject.R:1115-1215 (here: lower to the node table of include/nimb200.h)
  runMCMC         P/R/MCMC_run.R:82-213
The difference the back end exists for: runMCMC materialises the initial values of all chains first and
advances every chain in ONE batched device run, instead of the sequential chain loop at MCMC_run.R:127-153.
"""
import numpy as np

from . import _capi
from . import table as T
from .model import Model, ModelError, tree_reads_var


# nimbleOptions(): the few system options of the reference that reach this path (P/R/options.R)
_OPTIONS = {"MCMCsaveHistory": False}


def nimbleOptions(**kw):
    """nimbleOptions(MCMCsaveHistory = TRUE): RW / RW_block samplers keep their scale, acceptance-rate and proposal-covariance
    histories (P/R/MCMC_samplers.R:394-399, 726-739).  Read when the compiled MCMC allocates its device model."""
    for k, v in kw.items():
        if k not in _OPTIONS:
            raise KeyError("unknown option %s (supported: %s)" % (k, ", ".join(_OPTIONS)))
        _OPTIONS[k] = v
    return dict(_OPTIONS)


# ----------------------------------------------------------------------------- conjugacy (MCMC_conjugacy.R)
def _read_key(pos):
    """Which element of the variable a `cvar` read touches at each node: one element for all nodes, or one per node."""
    if isinstance(pos, (int, np.integer)):
        return ("el", int(pos))
    return ("els", np.asarray(pos, dtype=np.int64).tobytes())


def _linear_form(tree, var):
    """Is `tree` a + b*target for every scalar element `target` of `var` it reads?  Returns None (not linear), "free" (does
    not read var) or (has_offset, has_coeff, key).  Counterpart of cc_checkLinearity, P/R/MCMC_conjugacy.R:1397-1524, which
    works on ONE scalar target node: the other elements of the same variable are ordinary terms of the offset.  `key` says
    which element a node reads; when the expression reads several elements at one node (inprod over a slice, b[1] + b[2]*x)
    the key is "mixed" and both an offset and a coefficient are present for each of them -- the `linear` link, whose offset
    and coefficient the sampler finds per node by the set-to-0 / set-to-1 probing (MCMC_conjugacy.R:766-820)."""
    if not tree_reads_var(tree, var):
        return "free"
    k = tree[0]
    if k == "cvar":
        return (False, False, _read_key(tree[2]))
    if k == "inprod":
        # inprod(X[i, 1:p], beta[1:p]) seen from beta[j]: coefficient X[i, j], offset = the other p - 1 terms
        # (test-mcmc.R:1684-1700: conjugate_dnorm_dnorm_linear)
        if tree[4] == 1:
            return (False, True, ("el", int(tree[3])))
        return (True, True, "mixed")
    if k == "dyn":
        # mu[z[i]] as a function of one element of mu: that element (index equal) or another one -- offset and coefficient both
        # depend on the node, found by the numerical probing; the index itself must not read the target
        if tree_reads_var(tree[5], var):
            return None
        return (True, True, "mixed")
    if k == "un" and tree[1] == T.OP_NEG:
        f = _linear_form(tree[2], var)
        return None if f in (None, "free") else (f[0], True, f[2])
    if k == "bin":
        op, l, r = tree[1], tree[2], tree[3]
        fl, fr = _linear_form(l, var), _linear_form(r, var)
        if fl is None or fr is None:
            return None
        if op in (T.OP_ADD, T.OP_SUB):
            if fl == "free":
                return (True, fr[1] or op == T.OP_SUB, fr[2])
            if fr == "free":
                return (True, fl[1], fl[2])
            if fl[2] == fr[2] and fl[2] != "mixed":   # the same element on both sides: b + b, b - (a - b)
                return (fl[0] or fr[0], True, fl[2])
            return (True, True, "mixed")              # different elements: each is part of the other's offset
        if op == T.OP_MUL:
            if fl == "free":
                return (fr[0], True, fr[2])
            if fr == "free":
                return (fl[0], True, fl[2])
            # target on both sides is not linear (cc_checkLinearity returns NULL); two provably different elements are
            if fl[2] != "mixed" and fr[2] != "mixed" and fl[2][0] == "el" and fr[2][0] == "el" and fl[2] != fr[2]:
                return (True, True, "mixed")
            return None
        if op == T.OP_DIV and fr == "free":
            return (fl[0], True, fl[2])
    return None


def _link_code(f):
    off, co = f[0], f[1]
    if off and co:
        return T.LINK_LINEAR
    if off:
        return T.LINK_ADDITIVE
    if co:
        return T.LINK_MULTIPLICATIVE
    return T.LINK_IDENTITY


def check_conjugacy(model, decl):
    """Returns (sampler type, {dependent decl index: (link, param code)}) or None.
    Relationship table: P/R/MCMC_conjugacy.R:4-15 (dbeta prior: dbern / dbin dependents, identity link),
    :80-93 (dgamma prior), :109-115 (dnorm prior)."""
    var = decl.var
    if decl.dist not in (T.DGAMMA, T.DNORM, T.DBETA):
        return None
    if any(v == var for v, _ in decl.reads()):
        return None
    info = {}
    for dep in model.dependents(var):
        if dep is decl:
            continue
        bp = dep.bugs_params

        def only_in(name_):
            return name_ in bp and all(not tree_reads_var(t, var) for k, t in bp.items() if k != name_)
        if decl.dist == T.DBETA:
            if dep.dist == T.DBINOM and only_in("prob"):
                pname, code = "prob", 1
            else:
                return None
            f = _linear_form(bp[pname], var)
            if f in (None, "free") or f[0] or f[1]:  # beta prior: identity link only
                return None
        elif decl.dist == T.DGAMMA:
            if dep.dist == T.DPOIS and only_in("lambda"):
                pname, code = "lambda", 0
            elif dep.dist == T.DGAMMA and only_in("rate"):
                pname, code = "rate", T.PARAM_RATE
            elif dep.dist == T.DNORM and only_in("tau"):
                pname, code = "tau", T.PARAM_TAU
            elif dep.dist == T.DEXP and only_in("rate"):
                pname, code = "rate", 0
            elif dep.dist == T.DLNORM and only_in("taulog"):   # MCMC_conjugacy.R:86-92 (dgamma prior on taulog)
                pname, code = "taulog", T.PARAM_TAU
            else:
                return None
            f = _linear_form(bp[pname], var)
            if f in (None, "free") or f[0]:  # gamma prior: identity or multiplicative links only
                return None
        else:
            if dep.dist == T.DNORM and only_in("mean"):
                pname, code = "mean", 0
            elif dep.dist == T.DLNORM and only_in("meanlog"):   # MCMC_conjugacy.R:113: through log(value)
                pname, code = "meanlog", 0
            else:
                return None
            f = _linear_form(bp[pname], var)
            if f in (None, "free"):
                return None
        info[dep.index] = (_link_code(f), code)
    if not info:
        return None
    kind = {T.DGAMMA: T.SAMPLER_CONJ_GAMMA, T.DNORM: T.SAMPLER_CONJ_NORMAL, T.DBETA: T.SAMPLER_CONJ_BETA}[decl.dist]
    return kind, info


# ----------------------------------------------------------------------------- configuration
class SamplerConf:
    """samplerConf: name, target, control (P/R/MCMC_configuration.R:2-42)."""

    def __init__(self, type_, var, positions, control=None, conj=None):
        self.type, self.var, self.positions = type_, var, np.asarray(positions, dtype=np.int64)
        self.control, self.conj = dict(control or {}), conj
        self.name = T.SAMPLER_NAMES[type_]

    def target_names(self, model):
        return model.element_names(self.var, self.positions)


class MCMCconf:
    """MCMCconf: the ordered sampler list and the monitors (P/R/MCMC_configuration.R:73)."""

    def __init__(self, model, monitors=None, thin=1, useConjugacy=True, nodes=None, control=None,
                 multivariateNodesAsScalars=False, autoBlock=False, monitors2=None, thin2=1, enableWAIC=False):
        if autoBlock:
            raise NotImplementedError("autoBlock is outside the B200 sampling path")
        self.model, self.thin, self.samplerConfs = model, thin, []
        self.monitors = list(monitors) if monitors is not None else model.top_level_vars()
        # second monitor set with its own thinning interval -> mvSamples2 (MCMC_configuration.R `monitors2`, `thin2`)
        self.monitors2 = [monitors2] if isinstance(monitors2, str) else list(monitors2 or [])
        self.thin2 = thin2
        self.enableWAIC = bool(enableWAIC)   # online conditional WAIC, one data node per group (MCMC_configuration.R `enableWAIC`)
        self.control_defaults = dict(control or {})
        use = set(nodes) if nodes is not None else None
        self._classify_predictive(use)
        for var in model.topo_vars():
            if use is not None and var not in use:
                continue
            for decl in model.decl_of(var):
                self._add_default(decl, useConjugacy)
        # posterior predictive samplers run last (MCMC_configuration.R:236)
        pp = [sc for sc in self.samplerConfs if sc.type == T.SAMPLER_POSTERIOR_PREDICTIVE]
        self.samplerConfs = [sc for sc in self.samplerConfs if sc.type != T.SAMPLER_POSTERIOR_PREDICTIVE] + pp

    # default assignment rules, MCMC_configuration.R:340-446: conjugate -> binary -> (discrete: unsupported) -> RW
    _PP_DISTS = (T.DNORM, T.DGAMMA, T.DBETA, T.DEXP, T.DLOGIS, T.DPOIS, T.DBINOM, T.DLNORM)

    def _is_terminal(self, decl):
        return not [d for d in self.model.dependents(decl.var) if d is not decl]

    def _children(self, decl):
        return [d for d in self.model.dependents(decl.var) if d is not decl]

    def _downstream(self, decl):
        """Declarations downstream of `decl` (not itself), parents first."""
        seen, order = set(), []

        def visit(d):
            for c in self._children(d):
                if c.index not in seen:
                    seen.add(c.index)
                    visit(c)
                    order.append(c)
        visit(decl)
        rank = {v: i for i, v in enumerate(self.model.topo_vars())}
        return sorted(order, key=lambda d: (rank.get(d.var, len(rank)), d.index))

    def _classify_predictive(self, use):
        """Predictive declarations: stochastic, not data, no data anywhere downstream (BUGS_model.R:442).  Among them, the
        roots of simulable networks get ONE posterior_predictive sampler that simulates the whole downstream network, whose
        members get no sampler of their own (MCMC_configuration.R:247-287; sampler MCMC_samplers.R:86-105).  Granularity is
        the declaration: one whose nodes are only partly predictive is treated as non-predictive."""
        model = self.model
        pred = {d.index for d in model.decls if not d.is_data}
        changed = True
        while changed:
            changed = False
            for d in model.decls:
                if d.index in pred and any(c.index not in pred for c in self._children(d)):
                    pred.discard(d.index)
                    changed = True
        self.predictive_decls = pred
        # a network can be simulated on the device when every member has a variate generator and does not read itself
        ok = {i for i in pred if model.decls[i].dist in self._PP_DISTS
              and not any(v == model.decls[i].var for v, _ in model.decls[i].reads())
              and (use is None or model.decls[i].var in use)}
        changed = True
        while changed:
            changed = False
            for i in list(ok):
                if any(c.index not in ok for c in self._children(model.decls[i])):
                    ok.discard(i)
                    changed = True
        parents_ok = lambda d: any(model.decls[i].var == v for v, _ in d.reads() for i in ok)
        self.pp_roots = {i for i in ok if not parents_ok(model.decls[i])}
        self.pp_covered = ok - self.pp_roots

    def _add_default(self, decl, useConjugacy):
        model = self.model
        # the root of an entirely non-data network (an end node included) gets a posterior_predictive sampler which simulates the
        # root and everything downstream of it; the downstream members get nothing else (MCMC_configuration.R:283-287, 368)
        if decl.index in self.pp_covered:
            return
        if decl.index in self.pp_roots:
            sc = SamplerConf(T.SAMPLER_POSTERIOR_PREDICTIVE, decl.var, decl.pos, self.control_defaults)
            sc.network = [d.index for d in self._downstream(decl)]
            self.samplerConfs.append(sc)
            return
        conj = check_conjugacy(model, decl) if useConjugacy else None
        if conj is not None:
            self.samplerConfs.append(SamplerConf(conj[0], decl.var, decl.pos, self.control_defaults, conj=conj[1]))
            return
        if decl.dist == T.DBINOM and decl.params[0] == ("const", 1.0):
            self.samplerConfs.append(SamplerConf(T.SAMPLER_BINARY, decl.var, decl.pos, self.control_defaults))
            return
        if decl.dist == T.DCAT:  # MCMC_configuration.R:425
            if decl.K > T.MAX_CATEGORIES:
                raise NotImplementedError("node %s: categorical sampler supports at most %d categories" % (decl.var, T.MAX_CATEGORIES))
            self.samplerConfs.append(SamplerConf(T.SAMPLER_CATEGORICAL, decl.var, decl.pos, self.control_defaults))
            return
        if decl.dist in (T.DPOIS, T.DBINOM):  # other discrete nodes: slice, MCMC_configuration.R:428
            self.samplerConfs.append(SamplerConf(T.SAMPLER_SLICE, decl.var, decl.pos, self.control_defaults))
            return
        self.samplerConfs.append(SamplerConf(T.SAMPLER_RW, decl.var, decl.pos, self.control_defaults))

    def _resolve_target(self, target):
        """'beta', 'beta[3]', 'beta[1:50]' or a list of those -> (var, positions)."""
        import re
        targets = [target] if isinstance(target, str) else list(target)
        var, pos = None, []
        for tname in targets:
            m = re.match(r"^\s*([A-Za-z_.][\w.]*)\s*(?:\[(.*)\])?\s*$", tname)
            if not m or m.group(1) not in self.model.var_slot:
                raise ModelError("unknown target node %r" % tname)
            if var not in (None, m.group(1)):
                raise NotImplementedError("a sampler over several variables is not supported")
            var = m.group(1)
            shape = self.model.var_shape[var]
            if m.group(2) is None:
                pos.extend(range(int(np.prod(shape)) if shape else 1))
                continue
            subs = []
            for part, s in zip(m.group(2).split(","), shape):
                part = part.strip()
                if part == "":
                    subs.append(np.arange(1, s + 1))
                elif ":" in part:
                    a, b = part.split(":")
                    subs.append(np.arange(int(a), int(b) + 1))
                else:
                    subs.append(np.array([int(part)]))
            grid = np.meshgrid(*subs, indexing="ij")
            flat, mult = np.zeros(grid[0].size, dtype=np.int64), 1
            for g, s in zip(grid, shape):
                flat += (g.reshape(-1, order="F") - 1) * mult
                mult *= s
            pos.extend(flat.tolist())
        return var, np.array(pos, dtype=np.int64)

    def addSampler(self, target, type="RW", control=None, **kw):
        """conf$addSampler(target, type, control): P/R/MCMC_configuration.R:554-714."""
        ctrl = dict(self.control_defaults)
        ctrl.update(control or {})
        ctrl.update(kw)
        var, pos = self._resolve_target(target)
        tname = type.replace("sampler_", "")
        if tname == "RW":
            if len(pos) > 1:
                raise ValueError("cannot use RW sampler on more than one target; try RW_block sampler")
            self.samplerConfs.append(SamplerConf(T.SAMPLER_RW, var, pos, ctrl))
        elif tname == "RW_block":
            self.samplerConfs.append(SamplerConf(T.SAMPLER_RW_BLOCK, var, pos, ctrl))
        elif tname == "binary":
            if len(pos) > 1:
                raise ValueError("cannot use binary sampler on more than one target node")
            self.samplerConfs.append(SamplerConf(T.SAMPLER_BINARY, var, pos, ctrl))
        elif tname == "conjugate":
            for decl in self.model.decl_of(var):
                conj = check_conjugacy(self.model, decl)
                if conj is None:
                    raise ValueError("no conjugate relationship found for %s" % var)
                self.samplerConfs.append(SamplerConf(conj[0], var, pos, ctrl, conj=conj[1]))
        elif tname == "categorical":
            if len(pos) > 1:
                raise ValueError("cannot use categorical sampler on more than one target node")  # MCMC_samplers.R:182
            if any(d.dist != T.DCAT for d in self.model.decl_of(var)):
                raise ValueError("Can only use categorical sampler on node with dcat distribution.")  # :183
            self.samplerConfs.append(SamplerConf(T.SAMPLER_CATEGORICAL, var, pos, ctrl))
        elif tname == "slice":
            if len(pos) > 1:
                raise ValueError("cannot use slice sampler on more than one target node")  # MCMC_samplers.R:891
            self.samplerConfs.append(SamplerConf(T.SAMPLER_SLICE, var, pos, ctrl))
        elif tname == "posterior_predictive":
            net = []
            for decl in self.model.decl_of(var):
                if decl.index not in self.predictive_decls:
                    # MCMC_samplers.R:93-94
                    raise ValueError("posterior_predictive sampler should not be assigned to non-posterior-predictive nodes")
                for d in [decl] + self._downstream(decl):
                    if d.dist not in self._PP_DISTS:
                        raise NotImplementedError("posterior_predictive: no device variate generator for the distribution of %s" % d.var)
                net.extend(d.index for d in self._downstream(decl) if d.index not in net)
            sc = SamplerConf(T.SAMPLER_POSTERIOR_PREDICTIVE, var, pos, ctrl)
            sc.network = net
            self.samplerConfs.append(sc)
        else:
            raise NotImplementedError("sampler type %r is not part of the B200 sampling path (supported: RW, "
                                      "RW_block, binary, conjugate, posterior_predictive, slice, categorical); there is no CPU fallback" % type)
        return self

    def removeSamplers(self, target=None):
        if target is None:
            self.samplerConfs = []
            return self
        var, pos = self._resolve_target(target)
        keep = []
        for s in self.samplerConfs:
            if s.var != var:
                keep.append(s)
                continue
            left = np.setdiff1d(s.positions, pos)
            if left.size and s.type != T.SAMPLER_RW_BLOCK and left.size < s.positions.size:
                s.positions = left
                keep.append(s)
            elif left.size == s.positions.size:
                keep.append(s)
        self.samplerConfs = keep
        return self

    removeSampler = removeSamplers

    def addMonitors(self, *names):
        for n_ in names:
            for x in ([n_] if isinstance(n_, str) else n_):
                if x not in self.monitors:
                    self.monitors.append(x)
        return self

    def addMonitors2(self, *names):
        for n_ in names:
            for x in ([n_] if isinstance(n_, str) else n_):
                if x not in self.monitors2:
                    self.monitors2.append(x)
        return self

    def setThin(self, thin):
        self.thin = thin
        return self

    def setThin2(self, thin2):
        self.thin2 = thin2
        return self

    def setMonitors(self, *names):
        self.monitors = []
        return self.addMonitors(*names)

    def resetMonitors(self):
        self.monitors = []
        return self

    def getSamplers(self):
        return list(self.samplerConfs)

    def printSamplers(self):
        lines = []
        for i, s in enumerate(self.samplerConfs):
            names = s.target_names(self.model)
            tgt = names[0] if len(names) == 1 else "%s .. %s (%d nodes)" % (names[0], names[-1], len(names))
            lines.append("[%d] %s sampler: %s" % (i + 1, s.name, tgt))
        print("\n".join(lines))
        return lines


def configureMCMC(model, monitors=None, thin=1, useConjugacy=True, nodes=None, control=None, **kw):
    if not isinstance(model, Model):
        raise TypeError("configureMCMC: first argument must be a model")
    return MCMCconf(model, monitors, thin, useConjugacy, nodes, control, **kw)


# ----------------------------------------------------------------------------- lowering to the node table
class _Lowering:
    def __init__(self, model, conf):
        self.model, self.conf = model, conf
        self.b = T.TableBuilder()
        self.b.n_slots = model.n_slots
        self._dcache = {}
        self.decl_ids = {}
        self.colnames = []
        self.sampler_map = []  # table sampler index -> (SamplerConf, element positions)

    # -- shared arrays (deduplicated)
    def _darr(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
        key = (a.size, float(a[0]) if a.size else 0.0, float(a[-1]) if a.size else 0.0, float(a[::max(1, a.size // 64)].sum()))
        for cand in self._dcache.get(key, []):
            if np.array_equal(self.b.darr[cand], a):
                return cand
        i = self.b.add_darr(a)
        self._dcache.setdefault(key, []).append(i)
        return i

    def _flatten(self, tree, decl, ops):
        k = tree[0]
        if k == "const":
            ops.append((T.OP_CONST, 0, 0, 0, tree[1]))
        elif k == "data":
            v, n = tree[1], decl.n
            J = decl.loop_extents[-1] if len(decl.loop_extents) > 1 else 0
            if J > 1 and n % J == 0 and np.all(v.reshape(-1, J) == v.reshape(-1, J)[:, :1]):
                ops.append((T.OP_DATA, self._darr(v.reshape(-1, J)[:, 0]), T.IDX_DIV, J, 0.0))
            else:
                ops.append((T.OP_DATA, self._darr(v), T.IDX_I, 0, 0.0))
        elif k == "cvar":
            slot, pos = self.model.var_slot[tree[1]], tree[2]
            if np.isscalar(pos) or isinstance(pos, (int, np.integer)):
                ops.append((T.OP_CVAR, slot + int(pos), 0, 0, 0.0))
            else:
                n = decl.n
                ar = np.arange(n)
                J = decl.loop_extents[-1] if len(decl.loop_extents) > 1 else 0
                if np.all(pos - pos[0] == ar):
                    ops.append((T.OP_CVARN, slot + int(pos[0]), T.IDX_I, 0, 0.0))
                elif J > 1 and np.all(pos - pos[0] == ar // J):
                    ops.append((T.OP_CVARN, slot + int(pos[0]), T.IDX_DIV, J, 0.0))
                else:
                    ops.append((T.OP_CVARN, slot, T.IDX_IARR, self.b.add_iarr(pos), 0.0))
        elif k == "inprod":
            X, var, lo, p = tree[1], tree[2], tree[3], tree[4]
            ops.append((T.OP_INPROD, self._darr(X), self.model.var_slot[var] + lo, p, float(p)))
        elif k == "dataj":
            ops.append((T.OP_DATAJ, self._darr(tree[1]), 0, tree[1].shape[1], 0.0))
        elif k == "cvarj":
            ops.append((T.OP_CVARJ, self.model.var_slot[tree[1]] + tree[2], 0, 0, 0.0))
        elif k == "dyn":
            self._flatten(tree[5], decl, ops)
            if tree[1]:
                ops.append((T.OP_CVAR_IDX, self.model.var_slot[tree[2]] + tree[3], tree[4], 0, 0.0))
            else:
                ops.append((T.OP_DATA_IDX, self._darr(tree[2]), tree[4], 0, 0.0))
        elif k == "un":
            self._flatten(tree[2], decl, ops)
            ops.append((tree[1], 0, 0, 0, 0.0))
        elif k == "bin":
            self._flatten(tree[2], decl, ops)
            self._flatten(tree[3], decl, ops)
            ops.append((tree[1], 0, 0, 0, 0.0))
        else:
            raise ModelError("cannot flatten %r" % (k,))

    def _add_decl(self, decl):
        d = T.Decl()
        d.dist, d.n, d.nparam = decl.dist, decl.n, len(decl.params)
        if decl.is_data:
            d.value_kind, d.value_ref = 0, self._darr(decl.values)
        else:
            base = self.model.var_slot[decl.var]
            if not np.all(decl.pos - decl.pos[0] == np.arange(decl.n)):
                raise ModelError("%s: elements of a latent variable must be declared in storage order" % decl.var)
            d.value_kind, d.value_ref = 1, base + int(decl.pos[0])
        for i, tree in enumerate(decl.params):
            ops = []
            self._flatten(tree, decl, ops)
            d.prog_off[i], d.prog_len[i] = self.b.add_prog(ops)
        d.family = T.FAM_GENERIC
        if decl.dist == T.DCAT:
            d.fam_arg[0] = decl.K
        # fused family: y[i] ~ dbern(ilogit(inprod(X[i,], beta)))
        if (decl.dist == T.DBINOM and decl.is_data and decl.params[0] == ("const", 1.0)
                and decl.params[1][0] == "un" and decl.params[1][1] == T.OP_ILOGIT and decl.params[1][2][0] == "inprod"):
            ip = decl.params[1][2]
            d.family = T.FAM_BERN_LOGIT_INPROD
            d.fam_arg[0] = self._darr(ip[1])
            d.fam_arg[1] = self.model.var_slot[ip[2]] + ip[3]
            d.fam_arg[2] = ip[4]
            d.fam_arg[3] = ip[4]
        self._match_pois_loglin(decl, d)
        self._match_norm_gather(decl, d)
        self._match_bern_logit_lin(decl, d)
        self.b.decls.append(d)
        return len(self.b.decls) - 1

    @staticmethod
    def _chain_level(tree):
        """True if the expression reads nothing that depends on the node index."""
        k = tree[0]
        if k == "const":
            return True
        if k == "cvar":
            return isinstance(tree[2], (int, np.integer))
        if k == "un":
            return _Lowering._chain_level(tree[2])
        if k == "bin":
            return _Lowering._chain_level(tree[2]) and _Lowering._chain_level(tree[3])
        return False

    def _match_norm_gather(self, decl, d):
        """y[i] ~ dnorm(A + v[G(i)], sd = S) with A, S chain-level -> streaming normal family."""
        if decl.dist != T.DNORM or not decl.is_data or not self._chain_level(decl.params[1]):
            return
        mean = decl.params[0]
        A, v = None, None
        if self._chain_level(mean):
            A = mean
        elif mean[0] == "cvar":
            A, v = ("const", 0.0), mean
        elif mean[0] == "bin" and mean[1] == T.OP_ADD:
            for x, y in ((mean[2], mean[3]), (mean[3], mean[2])):
                if self._chain_level(x) and y[0] == "cvar" and not isinstance(y[2], (int, np.integer)):
                    A, v = x, y
                    break
        if A is None:
            return
        aops = []
        self._flatten(A, decl, aops)
        d.fam_arg[0], d.fam_arg[1] = self.b.add_prog(aops)
        d.fam_arg[3] = -1
        if v is not None:
            vops = []
            self._flatten(v, decl, vops)
            op, a, bmode, carg, _ = vops[0]
            d.fam_arg[2], d.fam_arg[3], d.fam_arg[4] = a, bmode, carg
        d.family = T.FAM_NORM_GATHER

    def _match_bern_logit_lin(self, decl, d):
        """v[i] ~ dbern(gate[G(i)] * ilogit(c0 + c1 * x[i])) (gate optional) -> Bernoulli-logit-linear family."""
        if d.family != T.FAM_GENERIC or decl.dist != T.DBINOM or decl.params[0] != ("const", 1.0):
            return
        prob, gate = decl.params[1], None
        if prob[0] == "bin" and prob[1] == T.OP_MUL:
            for g, q in ((prob[2], prob[3]), (prob[3], prob[2])):
                if g[0] == "cvar" and q[0] == "un" and q[1] == T.OP_ILOGIT:
                    gate, prob = g, q
                    break
        if prob[0] != "un" or prob[1] != T.OP_ILOGIT:
            return
        lin = prob[2]
        if lin[0] != "bin" or lin[1] != T.OP_ADD:
            return
        for c0, cx in ((lin[2], lin[3]), (lin[3], lin[2])):
            if c0[0] != "cvar" or not isinstance(c0[2], (int, np.integer)) or cx[0] != "bin" or cx[1] != T.OP_MUL:
                continue
            for c1, x in ((cx[2], cx[3]), (cx[3], cx[2])):
                if c1[0] != "cvar" or not isinstance(c1[2], (int, np.integer)) or x[0] != "data":
                    continue
                xops = []
                self._flatten(x, decl, xops)
                if xops[0][2] != T.IDX_I:
                    continue
                d.fam_arg[0] = xops[0][1]
                d.fam_arg[1] = self.model.var_slot[c0[1]] + int(c0[2])
                d.fam_arg[2] = self.model.var_slot[c1[1]] + int(c1[2])
                d.fam_arg[3] = -1
                if gate is not None:
                    gops = []
                    self._flatten(gate, decl, gops)
                    op, a, bmode, carg, _ = gops[0]
                    if op == T.OP_CVAR:
                        a, bmode, carg = a, T.IDX_DIV, max(decl.n, 1)
                    d.fam_arg[3], d.fam_arg[4], d.fam_arg[5] = a, bmode, carg
                d.family = T.FAM_BERN_LOGIT_LIN
                return

    def _match_pois_loglin(self, decl, d):
        """y[i] ~ dpois(lam[G(i)] * exp(b0 + b1 * x[i])) (any commutative arrangement) -> fused Poisson family."""
        if decl.dist != T.DPOIS or not decl.is_data:
            return
        t = decl.params[0]
        if t[0] != "bin" or t[1] != T.OP_MUL:
            return
        for lam, ex in ((t[2], t[3]), (t[3], t[2])):
            if lam[0] != "cvar" or ex[0] != "un" or ex[1] != T.OP_EXP:
                continue
            lin = ex[2]
            if lin[0] != "bin" or lin[1] != T.OP_ADD:
                continue
            for b0, bx in ((lin[2], lin[3]), (lin[3], lin[2])):
                if b0[0] != "cvar" or not isinstance(b0[2], (int, np.integer)) or bx[0] != "bin" or bx[1] != T.OP_MUL:
                    continue
                for b1, x in ((bx[2], bx[3]), (bx[3], bx[2])):
                    if b1[0] != "cvar" or not isinstance(b1[2], (int, np.integer)) or x[0] != "data":
                        continue
                    ops = []
                    self._flatten(lam, decl, ops)
                    op, a, bmode, carg, _ = ops[0]
                    if op == T.OP_CVAR:
                        base, mode, arg = a, T.IDX_DIV, max(decl.n, 1)      # one lam for every node
                    else:
                        base, mode, arg = a, bmode, carg
                    xops = []
                    self._flatten(x, decl, xops)
                    if xops[0][2] != T.IDX_I:
                        continue
                    d.family = T.FAM_POIS_LOGLIN
                    d.fam_arg[0], d.fam_arg[1], d.fam_arg[2], d.fam_arg[3] = xops[0][1], base, mode, arg
                    d.fam_arg[4] = self.model.var_slot[b0[1]] + int(b0[2])
                    d.fam_arg[5] = self.model.var_slot[b1[1]] + int(b1[2])
                    return

    # -- which nodes of `dep` read which target (relative position) of var
    def _target_of_nodes(self, dep, var, pos0, Tn):
        reads = [p for v, p in dep.reads() if v == var]
        if not reads:
            return None
        if any(isinstance(p, tuple) for p in reads):
            return "multi"
        arrs = [np.full(dep.n, int(p), dtype=np.int64) if np.isscalar(p) or isinstance(p, (int, np.integer)) else p
                for p in reads]
        for a in arrs[1:]:
            if not np.array_equal(a, arrs[0]):
                return "multi"
        return arrs[0] - pos0

    def _selection(self, tgt, n, Tn):
        """tgt[node] = target index (or out of [0,Tn)): pick the cheapest selection form."""
        ar = np.arange(n)
        inside = (tgt >= 0) & (tgt < Tn)
        if Tn == 1:
            if np.all(inside):
                return (T.SEL_ALL, 0, 0)
            nodes = np.nonzero(inside)[0]
            if nodes.size and np.all(np.diff(nodes) == 1):
                return (T.SEL_RANGE, int(nodes[0]), int(nodes[-1]) + 1)
            return (T.SEL_CSR, self.b.add_iarr(np.array([0, nodes.size])), self.b.add_iarr(nodes))
        if np.all(inside):
            if n == Tn and np.all(tgt == ar):
                return (T.SEL_SAME, 0, 0)
            if n % Tn == 0 and np.all(tgt == ar // (n // Tn)):
                return (T.SEL_STRIDE, n // Tn, 0)
            if np.all(np.diff(tgt) >= 0):
                off = np.searchsorted(tgt, np.arange(Tn + 1), side="left")
                return (T.SEL_SEG, self.b.add_iarr(off), 0)
        nodes = np.nonzero(inside)[0]
        order = nodes[np.argsort(tgt[nodes], kind="stable")]
        off = np.searchsorted(tgt[order], np.arange(Tn + 1), side="left")
        return (T.SEL_CSR, self.b.add_iarr(off), self.b.add_iarr(order))

    def _emit_sampler(self, sc, positions, vectorised, owner=None):
        model, b = self.model, self.b
        var = sc.var
        decl = [d for d in model.decl_of(var) if np.isin(positions[0], d.pos)][0]
        pos0, Tn = int(positions[0]), (len(positions) if vectorised else 1)
        block = sc.type == T.SAMPLER_RW_BLOCK
        d = len(positions) if block else 1
        s = T.Sampler()
        s.type, s.n_targets, s.d = sc.type, (1 if block else Tn), d
        s.target_slot = model.var_slot[var] + pos0
        ctl = sc.control
        s.log_scale = int(bool(ctl.get("log", False)))
        s.reflective = int(bool(ctl.get("reflective", False)))
        s.adaptive = int(bool(ctl.get("adaptive", True)))
        s.adapt_interval = int(ctl.get("adaptInterval", 200))
        s.adapt_scale_only = int(bool(ctl.get("adaptScaleOnly", False)))
        s.tries = int(ctl.get("tries", 1))
        s.scale = float(ctl.get("scale", 1.0))
        s.adapt_factor_exponent = float(ctl.get("adaptFactorExponent", 0.8))
        s.lower, s.upper = float(ctl.get("lower", -np.inf)), float(ctl.get("upper", np.inf))
        s.propcov_darr = -1
        if sc.type == T.SAMPLER_SLICE:  # MCMC_samplers.R:869-874
            s.scale = float(ctl.get("sliceWidth", 1.0))
            s.tries = int(ctl.get("sliceMaxSteps", 100))
            s.pad = int(ctl.get("maxContractions", 1000))
            if s.scale <= 0 or s.tries < 1 or s.pad < 1:
                raise ValueError("slice sampler: sliceWidth, sliceMaxSteps and maxContractions must be positive")
        if s.log_scale and s.reflective:
            raise ValueError("cannot use reflective RW sampler on a log scale (i.e. with options log=TRUE and reflective=TRUE")
        if s.adapt_factor_exponent < 0:
            raise ValueError("cannot use RW sampler with adaptFactorExponent control parameter less than 0")
        if s.scale < 0:
            raise ValueError("cannot use RW sampler with scale control parameter less than 0")
        if block:
            if d > T.MAX_BLOCK_D:
                raise ValueError("RW_block dimension %d exceeds the supported maximum %d" % (d, T.MAX_BLOCK_D))
            pc = ctl.get("propCov", "identity")
            if not isinstance(pc, str):
                pc = np.asarray(pc, dtype=np.float64)
                if pc.shape != (d, d):
                    raise ValueError("propCov matrix must have dimension %d x %d" % (d, d))
                if not np.allclose(pc, pc.T):
                    raise ValueError("propCov matrix must be symmetric")
                s.propcov_darr = b.add_darr(pc.reshape(-1, order="F"))
        s.dep_off = len(b.deps)
        span = d if block else Tn
        # stage 0: the target's own prior nodes (a block may span several declarations of the variable)
        if block:
            for dcl in model.decl_of(var):
                nodes = np.nonzero((dcl.pos >= pos0) & (dcl.pos < pos0 + span))[0]
                if not nodes.size:
                    continue
                dep0 = T.Dep()
                dep0.decl, dep0.stage = self.decl_ids[dcl.index], 0
                if np.all(np.diff(nodes) == 1):
                    dep0.sel, dep0.a, dep0.b = T.SEL_RANGE, int(nodes[0]), int(nodes[-1]) + 1
                else:
                    dep0.sel = T.SEL_CSR
                    dep0.a, dep0.b = b.add_iarr(np.array([0, nodes.size])), b.add_iarr(nodes)
                b.deps.append(dep0)
        else:
            node0 = int(np.nonzero(decl.pos == pos0)[0][0])
            dep0 = T.Dep()
            dep0.decl, dep0.stage = self.decl_ids[decl.index], 0
            if Tn == 1:
                dep0.sel, dep0.a, dep0.b = T.SEL_RANGE, node0, node0 + 1
            else:
                dep0.sel, dep0.a = T.SEL_SAME, node0
            b.deps.append(dep0)
        # stage 1: stochastic dependents.  Samplers of non-predictive targets leave predictive nodes out of their calculations
        # (the reference's default MCMCusePredictiveDependenciesInCalculations = FALSE, P/R/options.R:190, MCMC_build.R:156-176);
        # a sampler whose target is itself predictive keeps its whole Markov blanket (MCMC_utils.R:441-451)
        predictive = getattr(self.conf, "predictive_decls", set())
        skip_predictive = decl.index not in predictive
        for dep in model.dependents(var):
            if skip_predictive and dep.index in predictive and dep is not decl:
                continue
            tgt = self._target_of_nodes(dep, var, pos0, span)
            if tgt is None:
                continue
            if isinstance(tgt, str):  # node reads several elements: every node depends on every target
                if not block and Tn > 1:
                    raise ModelError("internal: non-separable dependents in a vectorised sampler")
                reads = [p for v, p in dep.reads() if v == var]
                hit = False
                for p in reads:
                    if isinstance(p, tuple):
                        hit |= (p[1] < pos0 + span) and (p[1] + p[2] > pos0)
                    else:
                        pa = np.asarray(p)
                        hit |= bool(np.any((pa >= pos0) & (pa < pos0 + span)))
                if not hit:
                    continue
                sel = (T.SEL_ALL, 0, 0)
            elif block:
                inside = (tgt >= 0) & (tgt < span)
                if not np.any(inside):
                    continue
                sel = self._selection(np.where(inside, 0, -1), dep.n, 1)
            else:
                if not np.any((tgt >= 0) & (tgt < Tn)):
                    continue
                sel = self._selection(tgt, dep.n, Tn)
            if dep.var == var and not dep.is_data:
                if block:
                    # x[t] ~ dnorm(x[t-1], ..): inside the block the prior nodes are the stage-0 entries above; a node OUTSIDE
                    # the block whose prior reads a block element would silently drop out of the MH ratio -> refuse
                    if np.any((dep.pos < pos0) | (dep.pos >= pos0 + span)):
                        raise ModelError("%s: RW_block must cover every node of a declaration that reads the variable itself" % var)
                    continue
                raise ModelError("%s: a node's own declaration reads the variable; not supported" % var)
            de = T.Dep()
            de.decl, de.stage = self.decl_ids[dep.index], 1
            de.sel, de.a, de.b = sel
            if sc.conj is not None:
                if dep.index not in sc.conj:
                    raise ModelError("internal: conjugate dependent missing")
                de.link, de.param = sc.conj[dep.index]
            b.deps.append(de)
        s.n_dep = len(b.deps) - s.dep_off
        b.samplers.append(s)
        self.sampler_map.append((owner if owner is not None else sc, np.asarray(positions)))

    def _separable(self, sc):
        """Can the targets of this scalar sampler be updated as one parallel sweep?  True when no
        stochastic node depends on two different targets and no target's prior reads another target
        (conditional independence given the rest), so that the sweep equals the reference's sequential visits."""
        model, var, pos = self.model, sc.var, sc.positions
        if len(pos) == 1:
            return True
        if not np.all(np.diff(pos) == 1):
            return False
        decls = [d for d in model.decl_of(var) if np.any(np.isin(pos, d.pos))]
        if len(decls) != 1:
            return False
        for dep in model.dependents(var):
            if dep is decls[0]:
                return False
            tgt = self._target_of_nodes(dep, var, int(pos[0]), len(pos))
            if isinstance(tgt, str):
                return False
        return True

    def run(self):
        model, b = self.model, self.b
        for d in model.decls:
            self.decl_ids[d.index] = self._add_decl(d)
        for sc in self.conf.samplerConfs:
            if sc.type == T.SAMPLER_RW_BLOCK:
                if not np.all(np.diff(sc.positions) == 1):
                    raise NotImplementedError("RW_block targets must be contiguous elements of one variable")
                self._emit_sampler(sc, sc.positions, False)
            elif self._separable(sc):
                self._emit_sampler(sc, sc.positions, True)
            else:
                for p in sc.positions:
                    self._emit_sampler(sc, np.array([p]), False)
            # posterior_predictive: model$simulate(simNodes) walks the whole downstream network in topological order
            # (MCMC_samplers.R:91, 97) -- here one table sampler per member declaration, parents first, right after the root's
            for di in getattr(sc, "network", []) if sc.type == T.SAMPLER_POSTERIOR_PREDICTIVE else []:
                dd = model.decls[di]
                sub = SamplerConf(T.SAMPLER_POSTERIOR_PREDICTIVE, dd.var, dd.pos, sc.control)
                if self._separable(sub):
                    self._emit_sampler(sub, sub.positions, True, owner=sc)
                else:
                    for p in sub.positions:
                        self._emit_sampler(sub, np.array([p]), False, owner=sc)
        # the device records ONE set of columns: monitors, then those of monitors2 that are not among them; run_chains splits
        # the rows and columns into mvSamples / mvSamples2 again
        self.cols1, self.cols2, self.colnames2 = [], [], []
        seen = {}
        for which, names in ((1, self.conf.monitors), (2, getattr(self.conf, "monitors2", []))):
            for name_ in names:
                if name_ not in model.var_slot:
                    raise ModelError("monitor %s is not a stochastic non-data variable of the model" % name_)
                shape = model.var_shape[name_]
                size = int(np.prod(shape)) if shape else 1
                if name_ not in seen:
                    seen[name_] = len(b.monitors)
                    b.monitors.extend(range(model.var_slot[name_], model.var_slot[name_] + size))
                cols = list(range(seen[name_], seen[name_] + size))
                if which == 1:
                    self.cols1.extend(cols)
                    self.colnames.extend(model.element_names(name_, np.arange(size)))
                else:
                    self.cols2.extend(cols)
                    self.colnames2.extend(model.element_names(name_, np.arange(size)))
        return b


# ----------------------------------------------------------------------------- buildMCMC / compileNimble / runMCMC
class MCMC:
    """Uncompiled MCMC object (buildMCMC's return value).  Only its compiled form can run here."""
    name = "MCMC"

    def __init__(self, conf):
        self.conf, self.model = conf, conf.model
        self.monitors, self.monitors2 = list(conf.monitors), list(getattr(conf, "monitors2", []))
        self.thinFromConfVec = [conf.thin, getattr(conf, "thin2", conf.thin)]

    def run(self, *a, **k):
        raise RuntimeError("uncompiled MCMC execution is not available in the B200 back end; call compileNimble() first")


def buildMCMC(conf, **kw):
    if isinstance(conf, Model):
        conf = configureMCMC(conf, **kw)
    if not isinstance(conf, MCMCconf):
        raise TypeError("buildMCMC: argument must be an MCMC configuration or a model")
    if "enableWAIC" in kw:
        conf.enableWAIC = bool(kw["enableWAIC"])
    return MCMC(conf)


class SampleMatrix(np.ndarray):
    """as.matrix(mvSamples): rows = saved iterations, columns = monitored scalar nodes (types_util.R:199-214)."""

    def __new__(cls, a, colnames):
        obj = np.asarray(a).view(cls)
        obj.colnames = list(colnames)
        return obj

    def __array_finalize__(self, obj):
        self.colnames = getattr(obj, "colnames", None)

    def col(self, name_):
        return np.asarray(self[:, self.colnames.index(name_)])


class CompiledMCMC:
    """Cmcmc: owns the lowered table and, once the chain count is known, the device-resident model."""
    name = "MCMC"

    def __init__(self, mcmc, backend=None, device=0, first_chain=0):
        self.Robject, self.model, self.conf = mcmc, mcmc.model, mcmc.conf
        self.backend = backend if backend is not None else _capi.default_backend()
        self.device, self.first_chain = device, first_chain
        self.lowering = _Lowering(self.model, self.conf)
        self.builder = self.lowering.run()
        self.colnames, self.colnames2 = self.lowering.colnames, self.lowering.colnames2
        self.mvSamples2 = None
        self.thinFromConfVec = mcmc.thinFromConfVec
        self.dm = None
        self.mvSamples = None
        self.check_initial_logprobs = True  # initializeModel's NaN / -Inf warnings (MCMC_build.R:241)
        self.enableWAIC = bool(getattr(self.conf, "enableWAIC", False))

    def _ensure(self, nchains):
        if self.dm is None or self.dm.C != nchains:
            if self.dm is not None:
                self.dm.close()
            self.dm = _capi.DeviceModel(self.backend, self.builder, nchains, self.first_chain, self.device)
            if self.enableWAIC:
                self.dm.waic_enable(True)
            if _OPTIONS["MCMCsaveHistory"] and self.backend.has_device:
                self.dm.history_enable(1024)
        return self.dm

    def _history(self, k, what):
        if self.dm is None or not _OPTIONS["MCMCsaveHistory"]:
            raise RuntimeError("Please set 'nimbleOptions(MCMCsaveHistory = TRUE)' before building the MCMC.")   # MCMC_samplers.R:426
        sc = self.conf.samplerConfs[k]
        parts = [self.dm.sampler_history(i, what) for i, (owner, _) in enumerate(self.lowering.sampler_map) if owner is sc]
        return np.concatenate(parts, axis=1)

    def getScaleHistory(self, k):
        """samplerFunctions[[k]]$getScaleHistory(): [adaptations][targets][chains] (0-based sampler index k)."""
        return self._history(k, 0)

    def getAcceptanceHistory(self, k):
        return self._history(k, 1)

    def getPropCovHistory(self, k):
        """[adaptations][d][d][chains] for an RW_block sampler with d <= 10 (maxDimCovHistory)."""
        h = self._history(k, 2)
        d = int(round(np.sqrt(h.shape[1])))
        return h.reshape(h.shape[0], d, d, h.shape[2]).transpose(0, 2, 1, 3)   # device layout is column-major

    def getWAIC(self):
        """mcmc$getWAIC(): the online WAIC of everything sampled since the last reset (MCMC_build.R:323-326, MCMC_WAIC.R:269-300)."""
        if not self.enableWAIC or self.dm is None:
            raise RuntimeError("WAIC was disabled based on the 'enableWAIC' argument to 'configureMCMC' or 'buildMCMC'")
        return self.dm.waic_get()

    def getTimes(self):
        """mcmc$getTimes() after run(time = TRUE) (MCMC_build.R:323-326): seconds per configured sampler, in samplerConfs order
        (a sampler lowered to several table samplers reports their sum)."""
        if self.dm is None or not getattr(self, "_timed", False):
            raise RuntimeError("getTimes: run the MCMC with time = True first")
        ms = self.dm.sampler_times()
        out = np.zeros(len(self.conf.samplerConfs))
        index = {id(sc): i for i, sc in enumerate(self.conf.samplerConfs)}
        for k, (sc, _) in enumerate(self.lowering.sampler_map):
            out[index[id(sc)]] += ms[k] * 1e-3
        return out

    def run_chains(self, inits_matrix, niter, nburnin=0, thin=None, reset=True, seed=0, thin2=None, time=False):
        """All chains at once.  inits_matrix: [nchains][n_slots] (ignored when reset is False)."""
        if niter < 0:
            raise ValueError("cannot specify niter < 0")
        if nburnin < 0:
            raise ValueError("cannot specify nburnin < 0")
        if nburnin > niter:
            raise ValueError("cannot specify nburnin > niter")
        thin = self.thinFromConfVec[0] if thin in (None, -1) else thin
        if thin < 1:
            raise ValueError("cannot use thin < 1")
        if thin != int(thin):
            raise ValueError("cannot use non-integer thin")
        if not reset and nburnin != 0:
            raise ValueError("cannot specify nburnin when using reset = FALSE.")
        dm = self._ensure(inits_matrix.shape[0])
        if time or getattr(self, "_timed", False):
            dm.set_timing(bool(time))
            self._timed = bool(time)
        if reset:
            dm.set_values(inits_matrix)
            if np.any(np.isnan(inits_matrix)):
                self._simulate_missing_inits(dm, inits_matrix, int(seed))
        if self.check_initial_logprobs:
            self._initialize_model_checks(dm)
        low = self.lowering
        if not low.cols2:
            dm.run(int(niter), int(nburnin), int(thin), reset, int(seed))
            self.mvSamples = dm.samples()
            return self.mvSamples
        # monitors2 / thin2 (MCMC_build.R:294-303): record the union of the columns every gcd(thin, thin2) iterations, then
        # take each set's own rows -- sample k of a set sits at post-burn-in iteration k * its thinning interval
        thin2 = self.thinFromConfVec[1] if thin2 in (None, -1) else thin2
        if thin2 < 1 or thin2 != int(thin2):
            raise ValueError("cannot use thin2 < 1")
        g = int(np.gcd(int(thin), int(thin2)))
        dm.run(int(niter), int(nburnin), g, reset, int(seed))
        allrows = dm.samples()
        # this call's rows are the last (niter - nburnin) // g of the buffer; thinning restarts with every run() call, as the
        # reference's loop counter does (MCMC_build.R:279-303), so a continuation selects inside its own rows and appends
        n_call = (int(niter) - int(nburnin)) // g
        mine = allrows[:, allrows.shape[1] - n_call:]
        n1, n2 = (int(niter) - int(nburnin)) // int(thin), (int(niter) - int(nburnin)) // int(thin2)
        r1 = (np.arange(1, n1 + 1) * (int(thin) // g) - 1).astype(np.int64)
        r2 = (np.arange(1, n2 + 1) * (int(thin2) // g) - 1).astype(np.int64)
        new1 = np.ascontiguousarray(mine[:, r1][:, :, low.cols1])
        new2 = np.ascontiguousarray(mine[:, r2][:, :, low.cols2])
        if reset or self.mvSamples is None or self.mvSamples2 is None:
            self.mvSamples, self.mvSamples2 = new1, new2
        else:   # reset = FALSE: mvSamples and mvSamples2 grow (MCMC_build.R:257-263)
            self.mvSamples = np.concatenate([self.mvSamples, new1], axis=1)
            self.mvSamples2 = np.concatenate([self.mvSamples2, new2], axis=1)
        return self.mvSamples

    def _simulate_missing_inits(self, dm, inits_matrix, seed):
        """initializeModel (P/R/initializeModel.R:78-139): NA non-data nodes are simulated from their distributions, parents
        first.  Refused up front when a variable with missing values has no device variate generator."""
        model = self.model
        missing = [v for v, s0 in model.var_slot.items()
                   if np.any(np.isnan(inits_matrix[:, s0:s0 + (int(np.prod(model.var_shape[v])) if model.var_shape[v] else 1)]))]
        bad = [v for v in missing if not model.simulable(v)]
        if bad:
            raise ModelError("initial values are required for %s: missing values can only be simulated for dnorm, dgamma, dbeta, "
                             "dexp, dlogis, dbin and dpois nodes" % ", ".join(bad))
        order = [self.lowering.decl_ids[d.index] for v in model.topo_vars() for d in model.decl_of(v)]
        dm.initialize_model(order, seed)

    def _initialize_model_checks(self, dm):
        """The warnings of initializeModel (P/R/initializeModel.R:97-139): every declaration is calculated once for every
        chain before sampling starts; NaN / -Inf log-probabilities are reported, sampling goes on regardless."""
        import warnings
        for di, decl in enumerate(self.model.decls):
            lp = dm.calculate(di)
            kind = "data node" if decl.is_data else "stochastic node"
            if np.any(np.isnan(lp)):
                warnings.warn("problem initializing %s %s: logProb is NA or NaN (%d of %d chains)."
                              % (kind, decl.var, int(np.isnan(lp).sum()), lp.size), RuntimeWarning, stacklevel=3)
            elif np.any(lp == -np.inf):
                warnings.warn("problem initializing %s %s: logProb is -Inf (%d of %d chains)."
                              % (kind, decl.var, int((lp == -np.inf).sum()), lp.size), RuntimeWarning, stacklevel=3)

    def run(self, niter, nburnin=0, thin=-1, reset=True, inits=None, seed=0, **_ignored):
        """mcmc$run(niter, ...) for a single chain (compatibility route)."""
        init = self.model.init_vector(inits)[None, :]
        if not reset and self.dm is not None:
            init = self.dm.get_values()
        self.run_chains(init, niter, nburnin, thin, reset, seed)

    def as_matrix(self, chain=0):
        return SampleMatrix(self.mvSamples[chain], self.colnames)


def compileNimble(*objs, backend=None, device=0, first_chain=0, **_ignored):
    """compileNimble(model, mcmc): returns the compiled MCMC (and the model, if it was passed too)."""
    out = []
    for o in objs:
        if isinstance(o, MCMC):
            out.append(CompiledMCMC(o, backend, device, first_chain))
        elif isinstance(o, Model):
            out.append(o)
        else:
            raise TypeError("compileNimble: cannot compile object of type %s" % type(o).__name__)
    return out[0] if len(out) == 1 else tuple(out)


def runMCMC(mcmc, niter=10000, nburnin=0, thin=None, inits=None, nchains=1, setSeed=False, samplesAsCodaMCMC=False,
            summary=False, progressBar=False, thin2=None, WAIC=False, **_ignored):
    """runMCMC(mcmc, niter, nburnin, thin, inits, nchains, setSeed): P/R/MCMC_run.R:82-213.

    inits: a dict (same values for every chain), a list of nchains dicts, or a callable returning a dict,
    evaluated once per chain in chain order exactly as the reference does (MCMC_run.R:135-142).
    setSeed: False, True (seed 0, per-chain streams keyed by chain index) or an int seed."""
    if getattr(mcmc, "name", None) != "MCMC":
        raise TypeError("mcmc argument must be a NIMBLE MCMC algorithm")
    if not isinstance(mcmc, CompiledMCMC):
        raise TypeError("mcmc argument must be a *compiled* NIMBLE MCMC algorithm")
    if isinstance(inits, (list, tuple)) and len(inits) != nchains and len(inits) != 1:
        raise ValueError("inits list must be of length nchains")
    # MCMC_run.R:130-134: TRUE -> seed 0 (plus the chain index), a number -> that seed; `1 == True` in Python, so test the type
    seed = 0 if (setSeed is None or isinstance(setSeed, (bool, np.bool_))) else int(setSeed)
    if setSeed is False:
        seed = int(np.random.SeedSequence().entropy % (1 << 63))
    per_chain = []
    for i in range(nchains):
        if callable(inits):
            these = inits()
        elif isinstance(inits, (list, tuple)):
            these = inits[i if len(inits) > 1 else 0]
        else:
            these = inits
        per_chain.append(these)
    samples = mcmc.run_chains(mcmc.model.init_matrix(per_chain), niter, nburnin, thin, True, seed, thin2=thin2)
    mats = [SampleMatrix(samples[c], mcmc.colnames) for c in range(nchains)]
    result = mats[0] if nchains == 1 else {"chain%d" % (c + 1): mats[c] for c in range(nchains)}
    ret = {"samples": result}
    has2 = bool(mcmc.lowering.cols2)
    if has2:  # MCMC_run.R:145-151, 183-207: a second list of matrices, `samples2`
        mats2 = [SampleMatrix(mcmc.mvSamples2[c], mcmc.colnames2) for c in range(nchains)]
        ret["samples2"] = mats2[0] if nchains == 1 else {"chain%d" % (c + 1): mats2[c] for c in range(nchains)}
    if summary and not has2 and mcmc.backend.has_device:
        # samplesSummary on the device (nimb200_samples_summary): nothing but 5 numbers per column crosses the bus;
        # all chains pooled, as the `all.chains` entry of the reference's summary (MCMC_run.R:196-207)
        tab = mcmc.dm.samples_summary()
        keys = ("Mean", "Median", "St.Dev.", "95%CI_low", "95%CI_upp")
        ret["summary"] = {name: {k: float(tab[i, j]) for i, k in enumerate(keys)} for j, name in enumerate(mcmc.colnames)}
    elif summary:
        from .ess import samples_summary
        summ = samples_summary(np.concatenate(mats, axis=0), mcmc.colnames)
        if has2:
            summ2 = samples_summary(np.concatenate(mats2, axis=0), mcmc.colnames2)
            summ = {**summ, **summ2} if isinstance(summ, dict) else summ
        ret["summary"] = summ
    if WAIC:  # MCMC_run.R:153-182: computed over all chains
        if not mcmc.enableWAIC:
            raise ValueError("mcmc argument must have been created with 'enableWAIC = TRUE' in order to calculate WAIC.")
        ret["WAIC"] = mcmc.getWAIC()
    return ret["samples"] if len(ret) == 1 else ret
