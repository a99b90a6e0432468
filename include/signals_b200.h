/* signals_b200.h -- C ABI of libsignals_b200.so: the B200 (sm_100a) implementation of the
 * per-cell haplotype- / allele-specific copy-number HMM of shahcompbio/signals.
 *
 * This is the drop-in boundary.  Every entry point replaces a specific piece of the reference
 * (citations are file:line into the reference repository); INTEGRATION.md shows the .Call shim
 * a maintainer adds on the R side (rpkg/src/r_shim.c) so the R API stays unchanged.
 *
 * Conventions
 *   - plain C types only; the caller owns every buffer; the library never retains caller memory
 *   - every function returns 0 on success, non-zero on failure; sgb_last_error() gives the text
 *     (the R shim turns it into Rf_error, mirroring BEGIN_RCPP/END_RCPP, src/RcppExports.cpp:16,23)
 *   - there is NO CPU implementation behind these symbols: a missing / unusable CUDA device is an error
 *   - work runs on the CUDA device that is current in the calling thread (cudaSetDevice / sgb_init)
 *   - not re-entrant: one call at a time per device (R is single threaded; `ncores` is ignored)
 */
#ifndef SIGNALS_B200_H
#define SIGNALS_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SGB_API __attribute__((visibility("default")))
#else
#define SGB_API
#endif

#define SGB_MODEL_HSCN 0          /* HaplotypeHMM  R/callHSCN.R:36-108 */
#define SGB_MODEL_ASCN 1          /* alleleHMM     R/callASCN.R:41-103 */
#define SGB_LIK_BINOMIAL 0        /* stats::dbinom        R/callHSCN.R:62 */
#define SGB_LIK_BETABINOMIAL 1    /* VGAM::dbetabinom     R/callHSCN.R:64 */
#define SGB_DECODE_REFERENCE 0    /* the reference's decode rule, src/code.cpp:35-42 (default, bit-exact) */
#define SGB_DECODE_CHAINED 1      /* textbook Viterbi backtrace over uint8 backpointers (opt-in extra) */

/* Arguments of HaplotypeHMM / alleleHMM that reach the kernels (R/callHSCN.R:36-46, R/callASCN.R:41-49).
 * minor_cn is always 0..n_states-1 (R/callHSCN.R:232: minor_cn <- seq(0, maxCN, 1)); n_states <= 32. */
typedef struct sgb_params {
    int32_t model;               /* SGB_MODEL_*  */
    int32_t likelihood;          /* SGB_LIK_*    */
    int32_t n_states;            /* maxCN + 1    */
    int32_t decode;              /* SGB_DECODE_* */
    double loherror;
    double eps;
    double rho;
    double Abias;
    double selftransitionprob;
} sgb_params;

typedef struct sgb_plan sgb_plan; /* sequence index of one batch, resident on the device */

/* ---- library / device ------------------------------------------------------------------ */
/* Select CUDA device `device` (>= 0) for this thread and create the library context on it.
 * Optional: every entry point creates the context of the current device on first use.
 * Replaces nothing in the reference (useDynLib(signals), NAMESPACE:107, needs no device). */
SGB_API int sgb_init(int device);
/* Release every context (streams, device buffers).  Counterpart of library unload. */
SGB_API int sgb_shutdown(void);
/* Text of the last error raised in the calling thread ("" if none). */
SGB_API const char* sgb_last_error(void);
/* Number of CUDA devices visible to the process (0 and an error text if CUDA is unusable). */
SGB_API int sgb_device_count(void);

/* ---- the batched hot path -------------------------------------------------------------- */
/* Replaces the per-cell / per-chromosome loops of .callHaplotypeSpecificCN_ (R/callHSCN.R:244-280
 * -> assignHaplotypeHMM :150-167 -> HaplotypeHMM :36-108 -> viterbi src/code.cpp:11-45) and of
 * callAlleleSpecificCN (R/callASCN.R:382-414 -> assignalleleHMM :143-156 -> alleleHMM :41-103).
 *
 * The N rows are the long CNBAF table already ordered by (cell_id, chr, start) as R/callHSCN.R:241-242
 * leaves it.  n = totalcounts, x = alleleB, cn = state (int32 when cn_itemsize == 4, uint8 when 1).
 * Sequence s (one cell x chromosome) covers rows seg_off[s] .. seg_off[s+1]-1, nseg sequences.
 * state_out[N] receives the decoded minor-CN state per row (the `minorcn` vector of each HMM call,
 * concatenated as assignHaplotypeHMM :164 does).  loglik_out, if not NULL, receives the bins x states
 * matrix `l` of every sequence (R/callHSCN.R:107): for sequence s a T x n_states column-major block
 * at loglik_out + seg_off[s]*n_states.
 * Host buffers may be pageable or pinned; inputs are copied to the device in chunks that overlap
 * with the kernels.  Counts must be in [0, 2^28), cn in [0, 255]; violations are reported as errors. */
SGB_API int sgb_hmm_decode(const sgb_params* params, int64_t N, const int32_t* n, const int32_t* x, const void* cn,
                   int32_t cn_itemsize, int64_t nseg, const int64_t* seg_off, uint8_t* state_out,
                   double* loglik_out);

/* Same work sharded over the first `n_devices` CUDA devices of this process (contiguous blocks of
 * sequences balanced by bin count; one host thread per device; each device writes its slice of
 * state_out).  B200 equivalent of parallel::mclapply(mc.cores = ncores), R/callHSCN.R:244-260. */
SGB_API int sgb_hmm_decode_multi(const sgb_params* params, int32_t n_devices, int64_t N, const int32_t* n,
                         const int32_t* x, const void* cn, int32_t cn_itemsize, int64_t nseg,
                         const int64_t* seg_off, uint8_t* state_out);

/* The same batched decode for columns of ANY element type, one type code per column -- what an R caller holds:
 * INTSXP columns are SGB_T_I32, REALSXP columns SGB_T_F64 (integer-valued doubles; NA / NaN / fractions are an error),
 * and state_out can be written as doubles directly (the reference returns state_min as numeric, R/callHSCN.R:169).
 * n_devices > 1 shards like sgb_hmm_decode_multi.  Pageable memory is fine: worker threads stage every chunk through
 * page-locked buffers, narrowing counts to uint16 (int32 when a value does not fit) and states to uint8, so 5 bytes per
 * cell-bin cross PCIe whatever the caller's types are (page-locked int32 / uint8 columns are copied directly instead:
 * on a host with ~60 GB/s of memory bandwidth the DMA engine reading them is faster than a staging pass).
 * Replaces the same reference lines as sgb_hmm_decode. */
#define SGB_T_U8 1
#define SGB_T_U16 2
#define SGB_T_I32 4
#define SGB_T_F64 8
SGB_API int sgb_hmm_decode_typed(const sgb_params* params, int64_t N, const void* n, int32_t n_type, const void* x,
                                 int32_t x_type, const void* cn, int32_t cn_type, int64_t nseg, const int64_t* seg_off,
                                 void* state_out, int32_t out_type, int32_t n_devices);

/* One upload, several passes.  callHaplotypeSpecificCN decodes the same sorted table twice -- pass 1 at the default
 * loherror, pass 2 at the inferred one and possibly beta-binomial (R/callHSCN.R:883-935, both through
 * .callHaplotypeSpecificCN_, R/callHSCN.R:217-295) -- and callAlleleSpecificCN a third time (R/callASCN.R:372-414); the
 * reference hands the three columns to every pass anew.  sgb_cols_upload converts (any element type, validated like
 * sgb_hmm_decode_typed) and copies them to the current device ONCE; sgb_cols_decode runs one pass on the resident columns
 * with any parameters (model, likelihood, loherror, rho, n_states, decode rule) and writes the states into a host vector
 * of any element type; sgb_cols_free releases the device memory.  9 bytes per cell-bin stay on the device between passes. */
typedef struct sgb_cols sgb_cols;
SGB_API int sgb_cols_upload(int64_t N, const void* n, int32_t n_type, const void* x, int32_t x_type, const void* cn,
                            int32_t cn_type, int64_t nseg, const int64_t* seg_off, sgb_cols** cols_out);
SGB_API int sgb_cols_decode(sgb_cols* cols, const sgb_params* params, void* state_out, int32_t out_type);
SGB_API int sgb_cols_free(sgb_cols* cols);

/* Device-resident variant: inputs and outputs are device pointers on the current device.
 * sgb_plan_create sorts the sequences by length on the host, uploads the index and reserves the
 * packed-record scratch; the plan can be reused for any inputs with the same seg_off.
 * sgb_hmm_decode_device only enqueues work on `stream` (a cudaStream_t, NULL = legacy default stream)
 * and returns; sgb_plan_check synchronises that stream and reports input-validation errors.
 * One exception: the packed-record scratch (8 B per cell-bin, needed by the beta-binomial likelihood, by decode = chained and
 * by columns that are not 16-byte aligned) is reserved by sgb_plan_create only when it is certain to be needed; otherwise the
 * FIRST decode call that needs it allocates it (an implicit device-wide synchronisation, not allowed during
 * stream capture).  Call sgb_plan_reserve_packed after sgb_plan_create to take that allocation up front. */
SGB_API int sgb_plan_create(int64_t N, int64_t nseg, const int64_t* seg_off, int32_t n_states, int32_t decode,
                    sgb_plan** plan_out);
SGB_API int sgb_plan_reserve_packed(sgb_plan* plan);
/* The default schedule of decode = reference calls (k_emit + k_recur) keeps the emission
 * log-likelihoods of the live states in device scratch between its two kernels: at most 32 B per cell-bin and
 * 4-state plane (96 B for maxCN = 11, usually ~41), capped at SGB_E_BUDGET_MB (default 8192) by decoding very large
 * batches in slices.  It is allocated by the first decode call of a plan, or up front by sgb_plan_reserve_split. */
SGB_API int sgb_plan_reserve_split(sgb_plan* plan);
/* diagnostic: after a split-schedule call, the number of 32-sequence groups and the bytes of emission tiles that
 * k_emit wrote and k_recur read (of the last slice); zeros after a fused-schedule call */
SGB_API int sgb_plan_split_info(sgb_plan* plan, void* stream, int64_t* ngroups_out, int64_t* e_bytes_out);
SGB_API int sgb_plan_destroy(sgb_plan* plan);
SGB_API int sgb_hmm_decode_device(const sgb_params* params, sgb_plan* plan, const int32_t* d_n, const int32_t* d_x,
                          const void* d_cn, int32_t cn_itemsize, uint8_t* d_state_out, double* d_loglik_out,
                          void* stream);
SGB_API int sgb_plan_check(sgb_plan* plan, void* stream);
/* Optional kernel timing for the device-resident path: when enabled, every sgb_hmm_decode_device call
 * records CUDA events on its stream around k_pack and k_decode (ring of the last 64 calls);
 * sgb_plan_read_timing returns the most recent durations in milliseconds, oldest first. */
SGB_API int sgb_plan_enable_timing(sgb_plan* plan, int32_t on);
SGB_API int sgb_plan_read_timing(sgb_plan* plan, int32_t max_records, float* pack_ms, float* decode_ms,
                                 int32_t* n_out);
/* three-phase view of the same records: preparation (k_pack, or the split schedule's partition kernels), k_emit
 * (0 on the fused schedule), and k_decode / k_recur */
SGB_API int sgb_plan_read_timing3(sgb_plan* plan, int32_t max_records, float* prep_ms, float* emit_ms, float* decode_ms,
                                  int32_t* n_out);
/* bytes of device scratch the plan holds (plan arrays, optional backpointers, and -- only once a call has
 * needed them -- the packed records of the k_pack path) */
SGB_API int64_t sgb_plan_scratch_bytes(const sgb_plan* plan);
/* How the plan's most recent sgb_hmm_decode_device call fed k_decode (no reference counterpart; diagnostic):
 * 2 = the caller's columns were read in place (binomial models, uint8 cn, n/x 16-byte and cn 4-byte aligned:
 * 2 kernels per call, k_setup + k_decode), 3 = in place after narrowing an int32 cn column to uint8 (k_setup +
 * k_narrow_cn + k_decode), 1 / 0 = through k_pack's packed records (TMA / plain loads: 3 kernels per call),
 * 4 = split schedule (k_emit + k_recur: the columns are read in place by k_emit, any element type), -1 = no call yet. */
SGB_API int sgb_plan_last_staging(const sgb_plan* plan);

/* ---- the two Rcpp exports ---------------------------------------------------------------- */
/* viterbi(emission, transition, observations)  src/code.cpp:11-45, R/RcppExports.R:8-10.
 * emission: nrow_e x S column-major (rows = bins, as the R matrix); transition: S x S column-major log
 * probabilities indexed [from, to]; observations: T 1-based row indices into emission.
 * path_out[T]: 0-based states.  init_mode 0 = -log(S) (C++ viterbi), 1 = log(1/S) (viterbiR, R/viterbi.R:3).
 * NaN emissions are treated as -Inf (the reference indexes with NA there: undefined).  S <= 128. */
SGB_API int sgb_viterbi(const double* emission, int64_t nrow_e, const double* transition, const int32_t* observations,
                int64_t T, int32_t S, int32_t init_mode, int32_t decode, int32_t* path_out);

/* Chained-decode viterbi for callers that must not allocate device memory (fixed-arena services,
 * latency-critical loops that cannot take cudaMalloc's device-wide stall -- see signals_b200/csrc/noalloc.cu).  This entry point has no counterpart
 * in the reference.  Same arithmetic as sgb_viterbi(..., decode = SGB_DECODE_CHAINED) but every pointer is a
 * DEVICE pointer on the current device, the work is only enqueued on `stream`, and the uint8 backpointers live in
 * a 4 MiB array built into the library instead of a library-allocated buffer: T * S must not exceed 4,194,304
 * (larger problems return an error).  It calls no cudaMalloc / cudaFree, but it does make ordinary non-allocating
 * runtime calls and a kernel launch, so it may not run inside a cudaLaunchHostFunc callback.  The first call on a device resolves the
 * array's address: make it before any stream capture.  Calls that overlap on one GPU race on the shared array.
 * Observations must already be valid 1-based row indices (they cannot be checked on the host). */
SGB_API int sgb_viterbi_chained_noalloc(const double* d_emission, int64_t nrow_e, const double* d_transition,
                                        const int32_t* d_observations, int64_t T, int32_t S, int32_t init_mode,
                                        int32_t* d_path_out, void* stream);

/* logspace_addcpp(logx, logy)  src/code.cpp:5-8, R/RcppExports.R:4-6 ; and its vectorised R twin
 * logspace_add (R/callASCN.R:9-11).  Evaluated on the device with the same deterministic math as the
 * emission kernel (a scalar call is a 1-element launch). */
SGB_API int sgb_logspace_add(double logx, double logy, double* out);
SGB_API int sgb_logspace_add_vec(const double* logx, const double* logy, int64_t n, double* out);

/* ---- post-decode table columns ---------------------------------------------- */
/* Clamp block R/callHSCN.R:282-290 (same arithmetic as R/callASCN.R:424-429) and the numeric part of
 * add_states R/util.R:1078-1097.  cn/path are the `state` column and the decoded state_min of N rows.
 * phase: 0 Balanced 1 A 2 B; state_phase: 0 Balanced 1 A-Gained 2 B-Gained 3 A-Hom 4 B-Hom; loh: 1 = "LOH". */
SGB_API int sgb_postprocess(int64_t N, const void* cn, int32_t cn_itemsize, const uint8_t* path, int32_t* A, int32_t* B,
                    int32_t* state_min, uint8_t* loh, uint8_t* phase, uint8_t* state_phase, double* state_baf);

/* ---- between-pass statistics ------------------------------------------------ */
/* The whole-table reductions the drivers make between the two HMM passes, as exact integer sums computed on
 * the device from the decoded paths: sums_out[0..2] = sum(alleleB), sum(totalcounts), row count over the
 * "A-Hom" rows (B == 0 < A after the clamp block) so that infloherror = weighted.mean(BAF, totalcounts) =
 * sums_out[0] / sums_out[1]  (R/callHSCN.R:898-902, R/callASCN.R:334-343); ab_hist_out[A*n_states + B] = number
 * of rows in state A|B with B > 0, from which fitBB picks its modal state (R/callHSCN.R:620-621). */
SGB_API int sgb_pass_stats(int64_t N, const int32_t* n, const int32_t* x, const void* cn, int32_t cn_itemsize,
                           const uint8_t* path, int32_t n_states, int64_t* sums_out, int64_t* ab_hist_out);

/* ---- the table passes either side of the hot path ------------------------------------ */
/* order(cell_id, chr, start) of the long table (R/callHSCN.R:241-242) and the (cell, chromosome) sequence boundaries
 * that replace the per-cell dplyr::filter scans (R/callHSCN.R:247,266) and the per-chromosome loops (:150-154).
 * cell / chr are integer codes in 0..n_cells-1 / 0..n_chr-1 whose numeric order is the order R gives the strings
 * (match(x, sort(unique(x))) - 1 in the C locale: chr "1","10",...,"2",...,"X"); start >= 0.
 * perm_out[N]: row i of the sorted table is input row perm_out[i] (stable).  seg_off_out[nseg+1] (capacity N+1). */
SGB_API int sgb_table_order(int64_t N, const int32_t* cell, const int32_t* chr, const int32_t* start, int32_t n_cells,
                            int32_t n_chr, int32_t* perm_out, int64_t* nseg_out, int64_t* seg_off_out);
/* The aggregation and filters of combineBAFCN (R/combinedata.R:309-325) on the joined CNbins x haplotypes rows:
 * [totalcounts > filtern], sums of alleleA / alleleB / totalcounts by (chr, start, end, cell_id, state, copy), cells with
 * <= minbins bins dropped, then cells whose smallest chromosome (chrX counted as 100) has <= minbinschr bins.  The key is
 * (cell, chr, start): end, state and copy are functions of it after the join (rows of one key that disagree on `state`
 * are an error).  Outputs (capacity N; seg_off N+1) are sorted by (cell, chr, start) -- the order
 * .callHaplotypeSpecificCN_ sorts into anyway -- with first_row_out[i] = the first input row of output row i, so the
 * caller can carry any other column (end, copy) over itself; BAF = alleleB_out / totalcounts_out. */
SGB_API int sgb_combine_bafcn(int64_t N, const int32_t* cell, const int32_t* chr, const int32_t* start, const int32_t* state,
                              const int32_t* alleleA, const int32_t* alleleB, const int32_t* totalcounts, int32_t n_cells,
                              int32_t n_chr, double filtern, int32_t minbins, int32_t minbinschr, int32_t chrX_code,
                              int64_t* nrow_out, int32_t* first_row_out, int32_t* cell_out, int32_t* chr_out,
                              int32_t* start_out, int32_t* state_out, int32_t* alleleA_out, int32_t* alleleB_out,
                              int32_t* totalcounts_out, int64_t* nseg_out, int64_t* seg_off_out);
/* diagnostic: device time (CUDA events, ms) of the kernels of the last sgb_table_order / sgb_combine_bafcn call, copies excluded */
SGB_API int sgb_table_last_ms(double* ms_out);
/* All the whole-table sums the drivers need between the two HMM passes, over the rows with row_keep[i] != 0 (NULL = all
 * rows; the cells that pass the first-pass filter, R/callHSCN.R:883-893), exact (integer accumulation, 128-bit for the
 * squares, converted once): out9 = { sum alleleB, sum totalcounts, rows } over the A-Hom rows (infloherror, :895-899), then
 * { rows, sum m, sum n, sum m^2, sum m n, sum n^2 } over the rows in state selA|selB with m = alleleB, n = totalcounts
 * (tarones_Z, :609-616: p = sum m / sum n, S = (sum m^2 - 2 p sum mn + p^2 sum n^2) / (p (1 - p)),
 * Z = (S - sum n) / sqrt(2 (sum n^2 - sum n))); ab_hist_out[A*n_states + B] as in sgb_pass_stats. */
SGB_API int sgb_pass_stats2(int64_t N, const int32_t* n, const int32_t* x, const void* cn, int32_t cn_itemsize,
                            const uint8_t* path, const uint8_t* row_keep, int32_t n_states, int32_t selA, int32_t selB,
                            double* out9, int64_t* ab_hist_out);
/* qc_per_cell's average_distance (R/util.R:667-691): per cell, over its allele-specific states with more than 10 bins,
 * |B/(A+B) - median(BAF)| weighted by n / (rows outside state 0|0).  cell_of_row: cell code of every row.
 * avg_out[n_cells]: NaN for a cell without a qualifying state (the reference has no row for it). */
SGB_API int sgb_qc_per_cell(int64_t N, const int32_t* n, const int32_t* x, const void* cn, int32_t cn_itemsize,
                            const uint8_t* path, const int32_t* cell_of_row, int32_t n_cells, double* avg_out);
/* fix_assignments (R/callHSCN.R:1072-1286), as written: singleton bins take their successor's B (the reference's
 * LLprev := fifelse(is.na(LLprev), 0, LLafter) makes the likelihood comparison always true), then per run of equal A|B
 * (inside a sequence for binomial, across the table for beta-binomial, as the reference's rleid calls differ) the
 * densities of the run's allele totals under B/state and A/state decide whether A and B swap.  Densities are dbinom /
 * dbetabinom values (not logs), so both underflow to 0 -- and nothing changes -- for long runs, as in R. */
SGB_API int sgb_fix_assignments(int64_t N, const int32_t* state, const int32_t* A, const int32_t* B, const int32_t* alleleA,
                                const int32_t* alleleB, int64_t nseg, const int64_t* seg_off, double loherror,
                                int32_t likelihood, double rho, int32_t* A_out, int32_t* B_out);

/* ---- diagnostics -------------------------------------------------------------------------- */
/* Evaluate one deterministic math primitive on the device (tests pin device bits == host bits).
 * fn: 0 exp, 1 log, 2 log1p on [0,1], 3 log1p, 4 lgamma, 5 logspace_add(in, in2), 6 log(in!) */
SGB_API int sgb_math_eval(int32_t fn, const double* in, const double* in2, int64_t n, double* out);
/* Time `iters` launches of an FP64 FMA throughput micro-kernel; *tflops_out = achieved DFMA TFLOP/s
 * (the measured FP64 roofline denominator). */
SGB_API int sgb_fp64_peak(int32_t iters, double* tflops_out);

#ifdef __cplusplus
}
#endif
#endif
