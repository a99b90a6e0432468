# hmm_gpu.R -- drop-in bodies for the reference's HMM functions (same names, arguments, defaults and
# return values); only the per-cell / per-chromosome loops are replaced by one batched .Call.
# SOURCE ONLY (no R where this repository is built); see INTEGRATION.md.
#
# Replaces: R/callHSCN.R:36-108 (HaplotypeHMM), :135-172 (assignHaplotypeHMM), :217-295
# (.callHaplotypeSpecificCN_), R/callASCN.R:9-11 (logspace_add), :41-103 (alleleHMM), :128-161 (assignalleleHMM) and the
# per-cell loops of callAlleleSpecificCN (:382-414) and callAlleleSpecificCNfromHSCN (:556-588) via .sgb_ascn_batch.
# R/RcppExports.R (viterbi, logspace_addcpp) is unchanged.

# GPUs to shard the cells over: options(signals.gpus = k) wins; otherwise `ncores` (what the reference hands to
# parallel::mclapply, R/callHSCN.R:245), capped by the number of visible devices.
.sgb_ngpus <- function(ncores = 1) {
  want <- getOption("signals.gpus", default = ncores)
  have <- .Call("_signals_device_count", PACKAGE = "signals")
  max(1L, min(as.integer(want), as.integer(have)))
}

.sgb_seg_off <- function(cell_id, chr) {
  n <- length(chr)
  if (n == 0) return(0)
  change <- c(TRUE, (chr[-1] != chr[-n]) | (cell_id[-1] != cell_id[-n]))
  c(which(change) - 1, n)
}

# like the reference (unique(CNBAF$chr) + filter, R/callHSCN.R:150-154) this only makes sense when the rows of a
# chromosome are contiguous; a table where they are not is an error rather than silently more sequences
.sgb_check_contiguous <- function(chr) {
  runs <- sum(c(TRUE, chr[-1] != chr[-length(chr)]))
  if (length(chr) > 0 && runs != length(unique(chr))) stop("rows of each chromosome must be contiguous")
}

.sgb_decode <- function(model, n, x, binstates, seg_off, maxCN, loherror, selftransitionprob, eps,
                        likelihood, rho, Abias, want_l = FALSE, ngpus = 1L) {
  # integer and numeric columns go over as they are (no as.integer copy): the library takes either
  .Call("_signals_hmm_decode_batch", as.integer(model), n, x, binstates, as.numeric(seg_off),
        likelihood, as.integer(maxCN), loherror, eps, rho, Abias, selftransitionprob, want_l, as.integer(ngpus),
        PACKAGE = "signals")
}

# vectorised twin of logspace_addcpp (R/callASCN.R:9-11).  One call = one launch over the whole vector; calling the
# scalar logspace_addcpp per element (the reference's idiom, R/callHSCN.R:68-70) would be a launch per element.
logspace_add <- function(logx, logy) {
  n <- max(length(logx), length(logy))
  .Call("_signals_logspace_add_vec", rep_len(as.numeric(logx), n), rep_len(as.numeric(logy), n), PACKAGE = "signals")
}

HaplotypeHMM <- function(n, x, binstates, minor_cn, loherror = 0.02, selftransitionprob = 0.999,
                         eps = 1e-12, likelihood = "binomial", rho = 0.0, Abias = 0.0, viterbiver = "cpp") {
  stopifnot(all(minor_cn == seq(0, length(minor_cn) - 1)))          # every driver passes 0:maxCN
  r <- .sgb_decode(0L, n, x, binstates, c(0, length(n)), length(minor_cn) - 1, loherror,
                   selftransitionprob, eps, likelihood, rho, Abias, want_l = TRUE)
  l <- matrix(r[[2]], length(n), length(minor_cn))
  res <- r[[1]]
  if (viterbiver == "R") {
    tp <- if (selftransitionprob == 0) matrix(1 / length(minor_cn), length(minor_cn), length(minor_cn)) else {
      m <- matrix((1 - selftransitionprob) / (length(minor_cn) - 1), length(minor_cn), length(minor_cn)); diag(m) <- selftransitionprob; m }
    res <- viterbiR(l, log(tp), observations = seq_len(length(binstates)))
  }
  list(minorcn = res, l = l)
}

alleleHMM <- function(n, x, binstates, minor_cn, loherror = 0.02, selftransitionprob = 0.999,
                      eps = 1e-12, rho = 0.0, likelihood = "binomial") {
  stopifnot(all(minor_cn == seq(0, length(minor_cn) - 1)))
  r <- .sgb_decode(1L, n, x, binstates, c(0, length(n)), length(minor_cn) - 1, loherror,
                   selftransitionprob, eps, likelihood, rho, 0.0, want_l = TRUE)
  list(minorcn = r[[1]], l = matrix(r[[2]], length(n), length(minor_cn)))
}

assignHaplotypeHMM <- function(CNBAF, minor_cn, eps = 1e-12, loherror = 0.02, selftransitionprob = 0.999,
                               pb = NULL, likelihood = "binomial", rho = 0.0, Abias = 0.0, viterbiver = "cpp") {
  if (!is.null(pb)) pb$tick()$print()
  .sgb_check_contiguous(as.character(CNBAF$chr))
  seg <- .sgb_seg_off(rep("", nrow(CNBAF)), as.character(CNBAF$chr))
  CNBAF$state_min <- .sgb_decode(0L, CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, seg, length(minor_cn) - 1,
                                 loherror, selftransitionprob, eps, likelihood, rho, Abias)[[1]]
  as.data.frame(data.table::as.data.table(CNBAF))
}

assignalleleHMM <- function(CNBAF, minor_cn, eps = 1e-12, loherror = 0.02, selftransitionprob = 0.999,
                            pb = NULL, rho = 0.0, likelihood = "binomial") {
  if (!is.null(pb)) pb$tick()$print()
  .sgb_check_contiguous(as.character(CNBAF$chr))
  seg <- .sgb_seg_off(rep("", nrow(CNBAF)), as.character(CNBAF$chr))
  CNBAF$state_min <- .sgb_decode(1L, CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, seg, length(minor_cn) - 1,
                                 loherror, selftransitionprob, eps, likelihood, rho, 0.0)[[1]]
  as.data.frame(data.table::as.data.table(CNBAF))
}

.callHaplotypeSpecificCN_ <- function(CNBAF, eps = 1e-12, loherror = 0.02, maxCN = 10, selftransitionprob = 0.999,
                                      progressbar = TRUE, ncores = 1, likelihood = "binomial", rho = 0.0,
                                      Abias = 0.0, viterbiver = "cpp") {
  # progressbar / viterbiver are accepted and ignored: one call decodes every cell; ncores picks the number of GPUs
  if (inherits(CNBAF, "sgb_table")) {                                               # columns already on the device
    tab <- CNBAF
    CNBAF <- data.table::copy(tab$CNBAF)
    CNBAF$state_min <- .sgb_table_decode(tab, 0L, maxCN, loherror, selftransitionprob, eps, likelihood, rho, Abias)
  } else {
    CNBAF <- data.table::as.data.table(CNBAF) %>% .[order(cell_id, chr, start)]    # R/callHSCN.R:241-242
    seg <- .sgb_seg_off(CNBAF$cell_id, as.character(CNBAF$chr))
    CNBAF$state_min <- .sgb_decode(0L, CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, seg, maxCN, loherror,
                                   selftransitionprob, eps, likelihood, rho, Abias, ngpus = .sgb_ngpus(ncores))[[1]]
  }
  alleleCN <- CNBAF %>%                                                             # R/callHSCN.R:282-292 unchanged
    .[, state_min := fifelse(state_min < 0, 0, state_min)] %>%
    .[, A := state - state_min] %>% .[, B := state_min] %>%
    .[, B := fifelse(B < 0, 0, B)] %>% .[, A := fifelse(A < 0, 0, A)] %>%
    .[, B := fifelse(B > state, state, B)] %>% .[, A := fifelse(A > state, state, A)]
  as.data.frame(add_states(alleleCN))
}

# Several passes on ONE upload of the table.  callHaplotypeSpecificCN calls .callHaplotypeSpecificCN_ twice on the same CNBAF
# (R/callHSCN.R:883-893 with the default loherror, :925-935 with the inferred one and the fitted rho); the three columns and
# their order do not change in between.  With
#   tab <- .sgb_table(CNBAF)                      # sort (R/callHSCN.R:241-242), offsets, one conversion + upload
#   hscn_data <- .callHaplotypeSpecificCN_(tab, loherror = loherror, ...)            # pass 1
#   hscn_data <- .callHaplotypeSpecificCN_(tab, loherror = infloherror, likelihood = likelihood, rho = bbfit$rho, ...)   # pass 2
#   .sgb_table_free(tab)                          # (or leave it to the garbage collector)
# the second pass costs the kernels and the copy of state_min.  A `tab` is list(handle, nrows, CNBAF sorted); a plain data
# frame still works everywhere (it is uploaded for that one call).
.sgb_table <- function(CNBAF) {
  CNBAF <- data.table::as.data.table(CNBAF) %>% .[order(cell_id, chr, start)]
  seg <- .sgb_seg_off(CNBAF$cell_id, as.character(CNBAF$chr))
  h <- .Call("_signals_cols_upload", CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, as.numeric(seg), PACKAGE = "signals")
  structure(list(handle = h[[1]], nrows = h[[2]], CNBAF = CNBAF), class = "sgb_table")
}
.sgb_table_decode <- function(tab, model, maxCN, loherror, selftransitionprob, eps, likelihood, rho, Abias) {
  .Call("_signals_cols_decode", tab$handle, tab$nrows, as.integer(model), as.character(likelihood), as.integer(maxCN),
        as.numeric(loherror), as.numeric(eps), as.numeric(rho), as.numeric(Abias), as.numeric(selftransitionprob),
        PACKAGE = "signals")
}
.sgb_table_free <- function(tab) invisible(.Call("_signals_cols_free", tab$handle, PACKAGE = "signals"))

# The per-cell loops of callAlleleSpecificCN (R/callASCN.R:382-414) and callAlleleSpecificCNfromHSCN (:556-588):
#   alleleCN <- rbindlist(lapply / mclapply(unique(CNBAF$cell_id), function(cell) assignalleleHMM(filter(CNBAF, cell_id == cell), ...)))
#               %>% .[order(cell_id, chr, start)]
# become
#   alleleCN <- .sgb_ascn_batch(CNBAF, maxCN, eps, infloherror, selftransitionprob, likelihood, rho, ncores)
# (callAlleleSpecificCN: likelihood = likelihood, rho = bbfit$rho; ...fromHSCN: hscn$likelihood$likelihood / $rho).
# Everything after the loop (:416-447, :590-621) is unchanged.
.sgb_ascn_batch <- function(CNBAF, maxCN, eps, loherror, selftransitionprob, likelihood, rho, ncores = 1) {
  if (inherits(CNBAF, "sgb_table")) {
    alleleCN <- data.table::copy(CNBAF$CNBAF)
    alleleCN$state_min <- .sgb_table_decode(CNBAF, 1L, maxCN, loherror, selftransitionprob, eps, likelihood, rho, 0.0)
    return(alleleCN)
  }
  alleleCN <- data.table::as.data.table(CNBAF) %>% .[order(cell_id, chr, start)]
  seg <- .sgb_seg_off(alleleCN$cell_id, as.character(alleleCN$chr))
  alleleCN$state_min <- .sgb_decode(1L, alleleCN$totalcounts, alleleCN$alleleB, alleleCN$state, seg, maxCN, loherror,
                                    selftransitionprob, eps, likelihood, rho, 0.0, ngpus = .sgb_ngpus(ncores))[[1]]
  alleleCN
}

# Device versions of two whole-table passes the drivers make (optional accelerations; same values):
#   .sgb_table_columns(state, state_min)  ->  A, B, state_min, LOH, phase, state_phase, state_BAF of the clamp block
#                                             R/callHSCN.R:282-290 + add_states R/util.R:1078-1097 (integer codes;
#                                             phase: c("Balanced","A","B")[code + 1], state_phase likewise)
#   .sgb_infloherror(...)                 ->  weighted.mean(BAF, totalcounts) over the A-Hom rows, R/callHSCN.R:898-902
.sgb_table_columns <- function(state, state_min) {
  r <- .Call("_signals_postprocess", as.integer(state), as.integer(state_min), PACKAGE = "signals")
  names(r) <- c("A", "B", "state_min", "LOH", "phase", "state_phase", "state_BAF")
  r$LOH <- ifelse(r$LOH == 1L, "LOH", "NO")
  r$phase <- c("Balanced", "A", "B")[r$phase + 1L]
  r$state_phase <- c("Balanced", "A-Gained", "B-Gained", "A-Hom", "B-Hom")[r$state_phase + 1L]
  r
}

.sgb_infloherror <- function(totalcounts, alleleB, state, state_min, maxCN, maxloherror) {
  r <- .Call("_signals_pass_stats", as.integer(totalcounts), as.integer(alleleB), as.integer(state),
             as.integer(state_min), as.integer(maxCN), PACKAGE = "signals")
  min(r[[1]][1] / r[[1]][2], maxloherror)
}
