# hmm_gpu.R -- drop-in bodies for the reference's HMM functions (same names, arguments, defaults and
# return values); only the per-cell / per-chromosome loops are replaced by one batched .Call.
# SOURCE ONLY (no R where this repository is built); see INTEGRATION.md.
#
# Replaces: R/callHSCN.R:36-108 (HaplotypeHMM), :135-172 (assignHaplotypeHMM), :217-295
# (.callHaplotypeSpecificCN_), R/callASCN.R:41-103 (alleleHMM), :128-161 (assignalleleHMM) and the
# per-cell loop of callAlleleSpecificCN (:382-414).  R/RcppExports.R (viterbi, logspace_addcpp) is unchanged.

.sgb_seg_off <- function(cell_id, chr) {
  n <- length(chr)
  if (n == 0) return(0)
  change <- c(TRUE, (chr[-1] != chr[-n]) | (cell_id[-1] != cell_id[-n]))
  c(which(change) - 1, n)
}

.sgb_decode <- function(model, n, x, binstates, seg_off, maxCN, loherror, selftransitionprob, eps,
                        likelihood, rho, Abias, want_l = FALSE) {
  .Call("_signals_hmm_decode_batch", as.integer(model), n, x, binstates, as.numeric(seg_off),
        likelihood, as.integer(maxCN), loherror, eps, rho, Abias, selftransitionprob, want_l,
        PACKAGE = "signals")
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
  seg <- .sgb_seg_off(rep("", nrow(CNBAF)), as.character(CNBAF$chr))
  CNBAF$state_min <- .sgb_decode(0L, CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, seg, length(minor_cn) - 1,
                                 loherror, selftransitionprob, eps, likelihood, rho, Abias)[[1]]
  as.data.frame(data.table::as.data.table(CNBAF))
}

assignalleleHMM <- function(CNBAF, minor_cn, eps = 1e-12, loherror = 0.02, selftransitionprob = 0.999,
                            pb = NULL, rho = 0.0, likelihood = "binomial") {
  if (!is.null(pb)) pb$tick()$print()
  seg <- .sgb_seg_off(rep("", nrow(CNBAF)), as.character(CNBAF$chr))
  CNBAF$state_min <- .sgb_decode(1L, CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, seg, length(minor_cn) - 1,
                                 loherror, selftransitionprob, eps, likelihood, rho, 0.0)[[1]]
  as.data.frame(data.table::as.data.table(CNBAF))
}

.callHaplotypeSpecificCN_ <- function(CNBAF, eps = 1e-12, loherror = 0.02, maxCN = 10, selftransitionprob = 0.999,
                                      progressbar = TRUE, ncores = 1, likelihood = "binomial", rho = 0.0,
                                      Abias = 0.0, viterbiver = "cpp") {
  # progressbar / ncores / viterbiver are accepted and ignored: one launch decodes every cell
  CNBAF <- data.table::as.data.table(CNBAF) %>% .[order(cell_id, chr, start)]      # R/callHSCN.R:241-242
  seg <- .sgb_seg_off(CNBAF$cell_id, as.character(CNBAF$chr))
  CNBAF$state_min <- .sgb_decode(0L, CNBAF$totalcounts, CNBAF$alleleB, CNBAF$state, seg, maxCN, loherror,
                                 selftransitionprob, eps, likelihood, rho, Abias)[[1]]
  alleleCN <- CNBAF %>%                                                             # R/callHSCN.R:282-292 unchanged
    .[, state_min := fifelse(state_min < 0, 0, state_min)] %>%
    .[, A := state - state_min] %>% .[, B := state_min] %>%
    .[, B := fifelse(B < 0, 0, B)] %>% .[, A := fifelse(A < 0, 0, A)] %>%
    .[, B := fifelse(B > state, state, B)] %>% .[, A := fifelse(A > state, state, A)]
  as.data.frame(add_states(alleleCN))
}
