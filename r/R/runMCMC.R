## runMCMC.R -- drop-in replacement for nimble::runMCMC on the B200 back end (same signature and return value as
## P/R/MCMC_run.R:82-213).  UNVERIFIED (no R in the build image); the tested twin is nimble_b200.mcmc.runMCMC.
##
## What changes: the reference loops over chains (MCMC_run.R:127-153), calling inits() / model$setInits() immediately
## before each chain's run.  Here the initial values of ALL chains are materialised first -- inits() is called in the same
## order and under the same set.seed() semantics (MCMC_run.R:130-142) -- then every chain advances in ONE device run and the
## result is split into chain1..chainK.  Attach this package after nimble so that it masks nimble::runMCMC.

runMCMC <- function(mcmc, niter = 10000, nburnin = 0, thin, thin2, nchains = 1, inits, setSeed = FALSE,
                    progressBar = FALSE, samples = TRUE, samplesAsCodaMCMC = FALSE, summary = FALSE, WAIC = FALSE,
                    device = 0L, first_chain = 0L) {
  if (!identical(nimble:::nfGetDefVar(mcmc, "name"), "MCMC")) stop("mcmc argument must be a NIMBLE MCMC algorithm")   # MCMC_run.R:99
  if (WAIC && !isTRUE(mcmc$enableWAIC)) stop("mcmc argument must have been created with 'enableWAIC = TRUE' in order to calculate WAIC.")
  if (missing(thin)) thin <- mcmc$thinFromConfVec[1]
  tab <- lower_model(mcmc)
  ## seeds: FALSE -> a fresh seed; TRUE -> seed 0; a number -> that seed (streams are keyed by (seed, global chain index), so
  ## one seed serves all chains -- R's RNG is never used on the device)
  seed <- if (identical(setSeed, FALSE)) sample.int(.Machine$integer.max, 1) else if (isTRUE(setSeed)) 0 else as.numeric(setSeed[1])
  init <- matrix(NA_real_, tab$n_slots, nchains)
  for (i in seq_len(nchains)) {
    if (isTRUE(setSeed)) set.seed(i) else if (is.numeric(setSeed) && length(setSeed) == nchains) set.seed(setSeed[i])      # MCMC_run.R:130-134
    these <- if (missing(inits)) list() else if (is.function(inits)) inits() else
             if (is.list(inits) && length(inits) > 0 && is.list(inits[[1]])) inits[[if (length(inits) > 1) i else 1]] else inits
    init[, i] <- flatten_inits(tab, c(these, mcmc$Robject$model$origInits)[!duplicated(names(c(these, mcmc$Robject$model$origInits)))])
  }
  ptr <- .Call("CALL_nimb200_build", tab, as.integer(nchains), as.integer(first_chain), as.integer(device))
  on.exit(.Call("CALL_nimb200_clear", ptr), add = TRUE)
  if (WAIC) .Call("CALL_nimb200_waic_enable", ptr, TRUE)                      # online accumulation on the device (MCMC_WAIC.R:246-268)
  .Call("CALL_nimb200_run", ptr, init, tab$decl_order, as.numeric(niter), as.numeric(nburnin), as.integer(thin), TRUE, as.numeric(seed))
  s <- aperm(.Call("CALL_nimb200_samples", ptr, length(tab$monitor_slots), as.integer(nchains)), c(2, 1, 3))
  mats <- lapply(seq_len(nchains), function(k) { m <- matrix(s[, , k], nrow = dim(s)[1]); colnames(m) <- tab$colnames; m })
  if (samplesAsCodaMCMC) mats <- lapply(mats, coda::as.mcmc)
  out <- if (nchains == 1) mats[[1]] else setNames(mats, paste0("chain", seq_len(nchains)))                                 # MCMC_run.R:183-189
  waic <- NULL
  if (WAIC) {                                                                 # c(WAIC, lppd, pWAIC, n) over all chains pooled
    w <- .Call("CALL_nimb200_waic_get", ptr)
    waic <- list(WAIC = w[1], lppd = w[2], pWAIC = w[3])
  }
  if (!summary && !WAIC) return(out)
  res <- list(samples = out)
  if (summary) {
    ## samplesSummary on the device (five numbers per column cross the bus): per chain, and all chains pooled (MCMC_run.R:196-207)
    dev_summary <- function(lo, hi) {
      m <- t(.Call("CALL_nimb200_samples_summary", ptr, as.integer(lo), as.integer(hi), length(tab$monitor_slots)))
      dimnames(m) <- list(tab$colnames, c("Mean", "Median", "St.Dev.", "95%CI_low", "95%CI_upp"))
      m
    }
    res$summary <- if (nchains == 1) dev_summary(0L, 1L) else
      c(setNames(lapply(seq_len(nchains), function(k) dev_summary(k - 1L, k)), paste0("chain", seq_len(nchains))),
        list(all.chains = dev_summary(0L, nchains)))
  }
  if (WAIC) res$WAIC <- waic
  res
}

## Compatibility route (SURVEY 8b-A): an object with the duck type the UNMODIFIED runMCMC loop touches -- $run(niter, nburnin =,
## thin =, reset =, chain =), $mvSamples convertible by as.matrix, $thinFromConfVec -- that runs ONE chain per run() call on
## the device.  Correct but unbatched; reset = FALSE continues the resident chain and appends (MCMC_build.R:243-263).
b200_mcmc <- function(mcmc, device = 0L) {
  tab <- lower_model(mcmc)
  self <- new.env()
  self$ptr <- .Call("CALL_nimb200_build", tab, 1L, 0L, as.integer(device))
  self$thinFromConfVec <- mcmc$thinFromConfVec
  self$Robject <- mcmc$Robject %||% mcmc
  self$run <- function(niter, nburnin = 0, thin = self$thinFromConfVec[1], reset = TRUE, chain = 1, ...) {
    init <- if (reset) matrix(flatten_inits(tab, as.list(self$Robject$model$origInits)), ncol = 1) else NULL
    .Call("CALL_nimb200_run", self$ptr, init, tab$decl_order, as.numeric(niter), as.numeric(nburnin), as.integer(thin),
          as.logical(reset), as.numeric(chain - 1))
    s <- .Call("CALL_nimb200_samples", self$ptr, length(tab$monitor_slots), 1L)
    self$mvSamples <- structure(t(matrix(s, nrow = dim(s)[1])), dimnames = list(NULL, tab$colnames))
    invisible(NULL)
  }
  self$getTimes <- function() .Call("CALL_nimb200_times", self$ptr, ncol(tab$samplers_i))
  ## nimbleOptions(MCMCsaveHistory = TRUE): call enableHistory() before run(); sampler is the 1-based index in conf$samplerConfs
  self$enableHistory <- function(max_adaptations = 1000L) .Call("CALL_nimb200_history_enable", self$ptr, as.integer(max_adaptations))
  self$getScaleHistory <- function(sampler) .Call("CALL_nimb200_history", self$ptr, as.integer(sampler - 1L), 0L, 1)
  self$getAcceptanceHistory <- function(sampler) .Call("CALL_nimb200_history", self$ptr, as.integer(sampler - 1L), 1L, 1)
  self$getPropCovHistory <- function(sampler, d) {
    h <- .Call("CALL_nimb200_history", self$ptr, as.integer(sampler - 1L), 2L, d * d)
    aperm(array(h, c(d, d, length(h) / (d * d))), c(3, 2, 1))                 # [adaptation, d, d], as MCMC_samplers.R:772-780 returns it
  }
  class(self) <- "b200_mcmc"
  self
}
as.matrix.b200_mcmc_samples <- function(x, ...) unclass(x)
