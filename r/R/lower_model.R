## lower_model.R -- lowers a NIMBLE model + MCMC configuration to the node table of include/nimb200.h.
##
## UNVERIFIED: R is not installed in the build image; this file has never been parsed or run.  It states, in R and against
## NIMBLE's own objects, the lowering that nimble_b200/model.py + nimble_b200/mcmc.py implement from BUGS text and that every
## test of this repository exercises (the table it must produce is the same).  Where the two disagree, the Python side and
## the header are authoritative.  Sources it reads in the reference:
##   model$modelDef$declInfo[[i]]  (P/R/BUGS_BUGSdecl.R:111-170): targetVarName, distributionName, valueExprReplaced,
##                                 unrolledIndicesMatrix, replacementNameExprs / symbolicParentNodesReplaced
##   reparameterisation table       P/R/distributions_inputList.R:20-149 (getDistributionInfo(dist)$altParams)
##   links already inverted         P/R/BUGS_modelDef.R:197, P/R/BUGS_mathCompatibility.R:3-12
##   model$getDependencies          P/R/BUGS_model.R:980-1060 (includePredictive = FALSE for non-predictive targets,
##                                 the reference's MCMCusePredictiveDependenciesInCalculations = FALSE default, options.R:190)
##   conf$samplerConfs              P/R/MCMC_configuration.R:2-42; conjugacy results in control$dep_* (MCMC_conjugacy.R:523-649)

NIMB <- list(
  dist = c(dnorm = 1L, dgamma = 2L, dpois = 3L, dbinom = 4L, dbin = 4L, dbern = 4L, dbeta = 5L, dlogis = 6L, dcat = 7L, dexp = 8L),
  op = c(const = 1L, data = 2L, cvar = 3L, cvarn = 4L, inprod = 5L, dataj = 6L, cvarj = 7L,
         `+` = 16L, `-` = 17L, `*` = 18L, `/` = 19L, `^` = 20L, pow = 20L,
         neg = 32L, exp = 33L, log = 34L, sqrt = 35L, ilogit = 36L, expit = 36L, logit = 37L, inv = 38L, sq = 39L, abs = 40L,
         log1p = 41L, icloglog = 42L, step = 43L),
  idx = c(i = 0L, div = 1L, mod = 2L, iarr = 3L),
  sel = c(all = 0L, same = 1L, stride = 2L, seg = 3L, csr = 4L, range = 5L),
  link = c(identity = 0L, additive = 1L, multiplicative = 2L, linear = 3L),
  sampler = c(RW = 1L, RW_block = 2L, binary = 3L, conjugate_dgamma = 4L, conjugate_dnorm = 5L, conjugate_dbeta = 6L,
              posterior_predictive = 7L, slice = 8L, categorical = 9L),
  param = c(rate = 16L, tau = 17L))

## ---- table builder: arrays de-duplicated by identity of content, programs appended to one op list
new_builder <- function() {
  b <- new.env()
  b$darr <- list(); b$iarr <- list(); b$ops <- matrix(0L, 4, 0); b$ops_k <- numeric(0)
  b$decl <- matrix(0L, 18, 0); b$samplers_i <- matrix(0L, 14, 0); b$samplers_d <- matrix(0, 4, 0); b$deps <- matrix(0L, 8, 0)
  b$monitor_slots <- integer(0); b$colnames <- character(0)
  b
}
add_darr <- function(b, x) {
  x <- as.double(x)
  for (k in seq_along(b$darr)) if (length(b$darr[[k]]) == length(x) && identical(b$darr[[k]], x)) return(k - 1L)
  b$darr[[length(b$darr) + 1L]] <- x
  length(b$darr) - 1L
}
add_iarr <- function(b, x) { b$iarr[[length(b$iarr) + 1L]] <- as.integer(x); length(b$iarr) - 1L }
add_prog <- function(b, ops) {            # ops: list of c(op, a, b, c, k)
  off <- ncol(b$ops)
  for (o in ops) { b$ops <- cbind(b$ops, as.integer(o[1:4])); b$ops_k <- c(b$ops_k, o[[5]]) }
  c(off, length(ops))
}

## ---- state layout: every non-data stochastic variable gets a run of slots, elements in column-major order (R's own)
state_layout <- function(model) {
  vars <- unique(model$modelDef$maps$nodeName_2_varName[model$getNodeNames(stochOnly = TRUE, includeData = FALSE)])
  slot <- 0L; out <- list()
  for (v in vars) {
    size <- length(model[[v]])
    out[[v]] <- list(slot = slot, dim = dim(model[[v]]) %||% size)
    slot <- slot + size
  }
  list(vars = out, n_slots = slot)
}
`%||%` <- function(a, b) if (is.null(a)) b else a
elem_pos <- function(layout, var, idx) {   # 0-based position of var[idx...] inside the variable (column-major)
  d <- layout$vars[[var]]$dim
  if (length(idx) == 0) return(0L)
  pos <- 0L; mult <- 1L
  for (k in seq_along(idx)) { pos <- pos + (idx[[k]] - 1L) * mult; mult <- mult * d[k] }
  pos
}

## ---- expressions: R call tree (after the reference's own replacement of deterministic nodes) -> postfix ops.
## `env` maps the declaration's index variables to integer vectors of length n (one entry per unrolled node instance).
flatten_expr <- function(b, e, env, n, model, layout) {
  if (is.numeric(e)) return(list(c(NIMB$op[["const"]], 0, 0, 0, as.double(e))))
  if (is.name(e)) {
    nm <- as.character(e)
    if (!is.null(env[[nm]])) return(list(c(NIMB$op[["data"]], add_darr(b, env[[nm]]), NIMB$idx[["i"]], 0, 0)))
    if (!is.null(layout$vars[[nm]])) return(list(c(NIMB$op[["cvar"]], layout$vars[[nm]]$slot, 0, 0, 0)))
    v <- model[[nm]]                                       # constant or (scalar) data
    return(list(c(NIMB$op[["const"]], 0, 0, 0, as.double(v))))
  }
  f <- as.character(e[[1]])
  if (f == "(") return(flatten_expr(b, e[[2]], env, n, model, layout))
  if (f == "[") {
    var <- as.character(e[[2]])
    idx <- lapply(as.list(e)[-(1:2)], function(ix) eval(ix, env))          # integer vectors (length n or 1)
    if (!is.null(layout$vars[[var]])) {                    # latent variable element(s)
      pos <- elem_pos(layout, var, idx)
      if (length(unique(pos)) == 1L) return(list(c(NIMB$op[["cvar"]], layout$vars[[var]]$slot + pos[1], 0, 0, 0)))
      pos <- rep_len(pos, n)
      if (all(diff(pos) == 1L)) return(list(c(NIMB$op[["cvarn"]], layout$vars[[var]]$slot + pos[1], NIMB$idx[["i"]], 0, 0)))
      return(list(c(NIMB$op[["cvarn"]], layout$vars[[var]]$slot, NIMB$idx[["iarr"]], add_iarr(b, pos), 0)))
    }
    vals <- do.call(`[`, c(list(model[[var]]), lapply(idx, function(z) cbind(rep_len(z, n)))))   # data / constants gathered per node
    vals <- model[[var]][do.call(cbind, lapply(idx, rep_len, n))]
    if (length(unique(vals)) == 1L) return(list(c(NIMB$op[["const"]], 0, 0, 0, vals[1])))
    return(list(c(NIMB$op[["data"]], add_darr(b, vals), NIMB$idx[["i"]], 0, 0)))
  }
  if (f == "inprod" || (f == "sum" && is.call(e[[2]]) && identical(e[[2]][[1]], as.name("*")))) {
    ab <- if (f == "sum") as.list(e[[2]])[2:3] else as.list(e)[2:3]         # sum(a * b) is inprod(a, b), MCMC_conjugacy.R:1478-1486
    lat <- which(vapply(ab, function(z) as.character(z[[2]]) %in% names(layout$vars), TRUE))
    stopifnot(length(lat) == 1L)
    lv <- ab[[lat]]; dv <- ab[[3L - lat]]
    rng <- eval(lv[[3]], env)                                               # beta[lo:hi]
    X <- t(vapply(seq_len(n), function(i) {                                 # X[i, lo:hi] row per node instance
      env_i <- lapply(env, function(z) z[min(i, length(z))])
      as.double(eval(dv, c(env_i, as.list(model))))
    }, numeric(length(rng))))
    return(list(c(NIMB$op[["inprod"]], add_darr(b, t(X)), layout$vars[[as.character(lv[[2]])]]$slot + rng[1] - 1L, length(rng), length(rng))))
  }
  args <- lapply(as.list(e)[-1], flatten_expr, b = b, env = env, n = n, model = model, layout = layout)
  if (f == "-" && length(args) == 1L) return(c(args[[1]], list(c(NIMB$op[["neg"]], 0, 0, 0, 0))))
  if (f == "^" && is.numeric(e[[3]]) && e[[3]] == 2) return(c(args[[1]], list(c(NIMB$op[["sq"]], 0, 0, 0, 0))))
  code <- NIMB$op[[f]]
  if (is.null(code)) stop("unsupported function in model code: ", f, "() -- there is no CPU fallback")
  c(do.call(c, args), list(c(code, 0, 0, 0, 0)))
}

## canonical ("Rdist") parameters: dnorm(mean, sd), dgamma(shape, scale), dbinom(size, prob), dlogis(location, scale),
## dexp(rate), dpois(lambda), dbeta(shape1, shape2) -- distributions_inputList.R:20-149
canonical_params <- function(dist, args) {
  q <- function(x) substitute(x)
  switch(dist,
    dnorm = list(args$mean, if (!is.null(args$sd)) args$sd else if (!is.null(args$var)) call("sqrt", args$var) else call("/", 1, call("sqrt", args$tau))),
    dgamma = list(args$shape, if (!is.null(args$scale)) args$scale else call("/", 1, args$rate)),
    dpois = list(args$lambda),
    dbern = list(1, args$prob), dbin = list(args$size, args$prob), dbinom = list(args$size, args$prob),
    dbeta = list(args$shape1, args$shape2),
    dlogis = list(args$location, if (!is.null(args$scale)) args$scale else call("/", 1, args$rate)),
    dexp = list(if (!is.null(args$rate)) args$rate else call("/", 1, args$scale)),
    stop("distribution ", dist, " is outside the supported set"))
}

lower_decl <- function(b, model, layout, di) {
  info <- model$modelDef$declInfo[[di]]
  if (info$type != "stoch") return(NULL)
  dist <- info$distributionName
  um <- info$unrolledIndicesMatrix                        # one row per node instance, one column per index variable
  n <- max(1L, nrow(um))
  env <- if (ncol(um)) as.list(as.data.frame(um)) else list()
  var <- info$targetVarName
  is_data <- all(model$isData(info$targetNodeName %||% var))
  args <- as.list(info$valueExprReplaced)[-1]
  par <- canonical_params(dist, args)
  d <- integer(18)                                          # struct nimb200_decl, field by field
  d[1] <- NIMB$dist[[dist]]; d[2] <- n
  if (is_data) { d[3] <- 0L; d[4] <- add_darr(b, values(model, model$expandNodeNames(info$targetExprReplaced))) }
  else { d[3] <- 1L; d[4] <- layout$vars[[var]]$slot + elem_pos(layout, var, lapply(seq_len(ncol(um)), function(k) um[1, k])) }
  d[5] <- length(par)
  for (k in seq_along(par)) {
    pr <- add_prog(b, flatten_expr(b, par[[k]], env, n, model, layout))
    d[5 + k] <- pr[1]; d[8 + k] <- pr[2]
  }
  d[12] <- 0L                                               # family: the device recognises the fused families itself from
  b$decl <- cbind(b$decl, d)                                # the programs when fam_arg is filled; left generic here (see
  ncol(b$decl) - 1L                                         # nimble_b200/mcmc.py::_match_* for the four matchers to port)
}

## dependents of a target as (declaration, selection) entries: ALL / SAME / STRIDE / SEG / CSR / RANGE (include/nimb200.h)
selection_of <- function(b, tgt, n, Tn) {
  inside <- tgt >= 0L & tgt < Tn
  if (Tn == 1L) {
    if (all(inside)) return(c(NIMB$sel[["all"]], 0L, 0L))
    nodes <- which(inside) - 1L
    if (length(nodes) && all(diff(nodes) == 1L)) return(c(NIMB$sel[["range"]], nodes[1], nodes[length(nodes)] + 1L))
    return(c(NIMB$sel[["csr"]], add_iarr(b, c(0L, length(nodes))), add_iarr(b, nodes)))
  }
  if (all(inside)) {
    if (n == Tn && all(tgt == seq_len(n) - 1L)) return(c(NIMB$sel[["same"]], 0L, 0L))
    if (n %% Tn == 0L && all(tgt == (seq_len(n) - 1L) %/% (n %/% Tn))) return(c(NIMB$sel[["stride"]], n %/% Tn, 0L))
    if (all(diff(tgt) >= 0L)) return(c(NIMB$sel[["seg"]], add_iarr(b, findInterval(seq_len(Tn + 1L) - 1L - 0.5, tgt)), 0L))
  }
  nodes <- which(inside) - 1L
  ord <- nodes[order(tgt[nodes + 1L])]
  c(NIMB$sel[["csr"]], add_iarr(b, findInterval(seq_len(Tn + 1L) - 1L - 0.5, tgt[ord + 1L])), add_iarr(b, ord))
}

lower_model <- function(mcmc) {
  conf <- mcmc$Robject$conf %||% mcmc$conf
  model <- conf$model
  layout <- state_layout(model)
  b <- new_builder()
  decl_id <- integer(0)
  for (di in seq_along(model$modelDef$declInfo)) { id <- lower_decl(b, model, layout, di); if (!is.null(id)) decl_id[as.character(di)] <- id }
  predictive <- model$getNodeNames(predictiveOnly = TRUE)
  for (sc in conf$samplerConfs) {
    type <- sub("^sampler_", "", sc$name)
    base <- sub("^(conjugate_d[a-z]+).*$", "\\1", type)
    code <- NIMB$sampler[[base]]
    if (is.null(code)) stop("sampler type '", sc$name, "' is not part of the B200 sampling path; there is no CPU fallback")
    tnodes <- model$expandNodeNames(sc$target)
    var <- model$getVarNames(nodes = tnodes)
    s_i <- integer(14); s_d <- c(sc$control$scale %||% 1, sc$control$adaptFactorExponent %||% 0.8,
                                 sc$control$lower %||% -Inf, sc$control$upper %||% Inf)
    s_i[1] <- code; s_i[2] <- if (base == "RW_block") 1L else length(tnodes)
    s_i[3] <- layout$vars[[var]]$slot + min(vapply(tnodes, function(nd) elem_pos(layout, var, as.list(model$modelDef$maps$nodeName_2_unrolledIndices[[nd]])), 1L))
    s_i[4] <- if (base == "RW_block") length(tnodes) else 1L
    s_i[5] <- ncol(b$deps)
    s_i[7] <- as.integer(isTRUE(sc$control$log)); s_i[8] <- as.integer(isTRUE(sc$control$reflective))
    s_i[9] <- as.integer(sc$control$adaptive %||% TRUE); s_i[10] <- as.integer(sc$control$adaptInterval %||% 200L)
    s_i[11] <- as.integer(isTRUE(sc$control$adaptScaleOnly)); s_i[12] <- as.integer(sc$control$tries %||% 1L); s_i[13] <- -1L
    ## stage 0: the target's own declaration; stage 1: its stochastic dependents (predictive ones left out unless the target is predictive)
    keep_pred <- all(tnodes %in% predictive)
    deps <- model$getDependencies(tnodes, self = FALSE, stochOnly = TRUE, includePredictive = keep_pred)
    own <- unique(model$modelDef$maps$graphID_2_declID[model$expandNodeNames(tnodes, returnType = "ids")])
    b$deps <- cbind(b$deps, c(decl_id[[as.character(own[1])]], if (length(tnodes) == 1L) NIMB$sel[["range"]] else NIMB$sel[["same"]],
                              0L, if (length(tnodes) == 1L) 1L else 0L, 0L, 0L, 0L, 0L))
    for (dd in unique(model$modelDef$maps$graphID_2_declID[model$expandNodeNames(deps, returnType = "ids")])) {
      dnodes <- model$expandNodeNames(model$modelDef$declInfo[[dd]]$targetExprReplaced)
      tgt <- vapply(dnodes, function(nd) {                 # which target each node instance of the dependent declaration reads
        hit <- which(tnodes %in% model$getParents(nd, stochOnly = TRUE, immediateOnly = FALSE))
        if (length(hit) == 1L) hit - 1L else if (length(hit)) -2L else -1L
      }, 1L)
      sel <- if (any(tgt == -2L)) c(NIMB$sel[["all"]], 0L, 0L) else selection_of(b, tgt, length(dnodes), s_i[2])
      link <- 0L; param <- 0L
      if (grepl("^conjugate", type)) {                     # realised link and linked parameter, from the conjugacy system
        dep_dist <- model$getDistribution(dnodes[1])
        lk <- names(Filter(function(z) any(dnodes %in% z), sc$control[grep("^dep_", names(sc$control))]))
        link <- NIMB$link[[sub("^dep_[a-z]+_", "", lk[1])]] %||% 0L
        param <- if (base == "conjugate_dgamma" && dep_dist == "dgamma") NIMB$param[["rate"]] else
                 if (base == "conjugate_dgamma" && dep_dist == "dnorm") NIMB$param[["tau"]] else
                 if (base == "conjugate_dbeta") 1L else 0L
      }
      b$deps <- cbind(b$deps, c(decl_id[[as.character(dd)]], sel, 1L, link, param, 0L))
    }
    s_i[6] <- ncol(b$deps) - s_i[5]
    b$samplers_i <- cbind(b$samplers_i, s_i); b$samplers_d <- cbind(b$samplers_d, s_d)
  }
  for (mon in mcmc$monitors %||% conf$monitors) {
    size <- length(model[[mon]])
    b$monitor_slots <- c(b$monitor_slots, layout$vars[[mon]]$slot + seq_len(size) - 1L)
    b$colnames <- c(b$colnames, model$expandNodeNames(mon, returnScalarComponents = TRUE))
  }
  ## parents-first order of the latent declarations, for initializeModel's simulation of NA values
  order <- unlist(lapply(model$getNodeNames(stochOnly = TRUE, includeData = FALSE), function(nd)
    decl_id[[as.character(model$modelDef$maps$graphID_2_declID[model$expandNodeNames(nd, returnType = "ids")])]]))
  list(n_slots = layout$n_slots, darr = b$darr, iarr = b$iarr, ops = b$ops, ops_k = b$ops_k, decl = b$decl,
       samplers_i = b$samplers_i, samplers_d = b$samplers_d, deps = b$deps, monitor_slots = as.integer(b$monitor_slots),
       colnames = b$colnames, decl_order = as.integer(unique(order)), layout = layout)
}

## one chain's inits list -> state vector (NA where no value is given: simulated on the device, initializeModel.R:78-139)
flatten_inits <- function(tab, inits) {
  out <- rep(NA_real_, tab$n_slots)
  for (v in intersect(names(inits), names(tab$layout$vars))) {
    lv <- tab$layout$vars[[v]]
    out[lv$slot + seq_along(inits[[v]])] <- as.double(inits[[v]])
  }
  out
}
