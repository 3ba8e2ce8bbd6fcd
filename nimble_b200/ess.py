"""Effective sample size and posterior summaries for the measurement harness.

The reference defines MCMC efficiency as min over monitored columns of coda::effectiveSize / run time
(P/R/MCMC_autoBlock.R:297-306).  `coda` is not vendored in the reference tree, so its published estimator
is restated here once and applied to both the GPU samples and the CPU baseline samples:
effectiveSize(x) = n * var(x) / spectrum0.ar(x), with spectrum0.ar = var.pred / (1 - sum(ar))^2 from an
AR(p) fit by Yule-Walker with the order chosen by AIC (R's ar.yw, order.max = min(n-1, floor(10 log10 n))).
"""
import numpy as np


def _ar_yw(x):
    x = np.asarray(x, dtype=np.float64)
    n = x.size
    order_max = int(min(n - 1, np.floor(10 * np.log10(n))))
    xc = x - x.mean()
    r = np.array([np.dot(xc[: n - k], xc[k:]) / n for k in range(order_max + 1)])
    if r[0] == 0:
        return 0, np.zeros(0), 0.0
    # Levinson-Durbin
    vars_ = np.empty(order_max + 1)
    coefs = np.zeros((order_max + 1, order_max + 1))
    vars_[0] = r[0]
    for k in range(1, order_max + 1):
        acc = r[k] - np.dot(coefs[k - 1, 1:k], r[1:k][::-1])
        phi = acc / vars_[k - 1]
        coefs[k, k] = phi
        coefs[k, 1:k] = coefs[k - 1, 1:k] - phi * coefs[k - 1, 1:k][::-1]
        vars_[k] = vars_[k - 1] * (1 - phi * phi)
    aic = n * np.log(np.maximum(vars_, 1e-300)) + 2 * np.arange(order_max + 1) + 2
    order = int(np.argmin(aic))
    var_pred = vars_[order] * n / (n - (order + 1))
    return order, coefs[order, 1:order + 1], var_pred


def effective_size(x):
    """coda::effectiveSize for one column."""
    x = np.asarray(x, dtype=np.float64)
    n = x.size
    if n < 4:
        return 0.0
    v = x.var(ddof=1)
    if v == 0:
        return 0.0
    order, ar, var_pred = _ar_yw(x)
    spec = var_pred / (1 - ar.sum()) ** 2
    if spec <= 0 or not np.isfinite(spec):
        return 0.0
    return float(n * v / spec)


def effective_size_batch(x):
    """coda::effectiveSize of many series at once: x[series][n] -> ess[series].  Same estimator as effective_size() (AR order
    by AIC over Yule-Walker fits from a Levinson-Durbin recursion), with the recursion carried over all series together."""
    x = np.asarray(x, dtype=np.float64)
    S, n = x.shape
    if n < 4:
        return np.zeros(S)
    order_max = int(min(n - 1, np.floor(10 * np.log10(n))))
    xc = x - x.mean(1, keepdims=True)
    v = (xc * xc).sum(1) / (n - 1)
    r = np.empty((order_max + 1, S))
    for k in range(order_max + 1):
        r[k] = np.einsum("sn,sn->s", xc[:, : n - k], xc[:, k:]) / n
    ok = r[0] > 0
    r0 = np.where(ok, r[0], 1.0)
    vars_ = np.empty((order_max + 1, S))
    coefs = np.zeros((order_max + 1, order_max + 1, S))
    vars_[0] = r0
    for k in range(1, order_max + 1):
        acc = r[k].copy()
        for j in range(1, k):
            acc -= coefs[k - 1, j] * r[k - j]
        phi = acc / vars_[k - 1]
        coefs[k, k] = phi
        for j in range(1, k):
            coefs[k, j] = coefs[k - 1, j] - phi * coefs[k - 1, k - j]
        vars_[k] = vars_[k - 1] * (1 - phi * phi)
    aic = n * np.log(np.maximum(vars_, 1e-300)) + 2 * np.arange(order_max + 1)[:, None] + 2
    order = np.argmin(aic, axis=0)
    sidx = np.arange(S)
    var_pred = vars_[order, sidx] * n / (n - (order + 1))
    arsum = np.zeros(S)
    for j in range(1, order_max + 1):
        arsum += np.where(j <= order, coefs[order, j, sidx], 0.0)
    spec = var_pred / (1 - arsum) ** 2
    good = ok & (v > 0) & (spec > 0) & np.isfinite(spec)
    return np.where(good, n * v / np.where(good, spec, 1.0), 0.0)


def min_ess(samples):
    """samples: [nchains][nrows][ncols] -> (min over columns of the sum over chains of ESS, per-column sums).
    Columns with zero ESS are dropped, as in MCMC_autoBlock.R:303-306."""
    samples = np.asarray(samples)
    C, n, k = samples.shape
    tot = np.zeros(k)
    step = max(1, (1 << 24) // max(1, n * k))        # chains per batch: ~128 MB of doubles at a time
    for c0 in range(0, C, step):
        blk = samples[c0:c0 + step]
        e = effective_size_batch(np.ascontiguousarray(blk.transpose(0, 2, 1)).reshape(-1, n))
        tot += e.reshape(blk.shape[0], k).sum(0)
    nz = tot[tot > 0]
    return (float(nz.min()) if nz.size else 0.0), tot


def samples_summary(samples, colnames):
    """samplesSummary (P/R/MCMC_utils.R:315-341): Mean, Median, St.Dev., 95% CI."""
    s = np.asarray(samples)
    out = {}
    for j, name in enumerate(colnames):
        col = s[:, j]
        out[name] = {"Mean": float(col.mean()), "Median": float(np.median(col)), "St.Dev.": float(col.std(ddof=1)),
                     "95%CI_low": float(np.quantile(col, 0.025)), "95%CI_upp": float(np.quantile(col, 0.975))}
    return out
