// kernels.h -- host-side launch interface of the sm_100a kernels (hmm_kernels.cu).
// Internal to the library; the public boundary is include/signals_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "hmm_core.cuh"

// Per-call constants computed ON THE DEVICE by k_setup (so they carry device bits):
// scalar parameters plus the per-(total CN, state) tables for total CN < SGB_CN_TABLE.
struct SgbDevTables {
    SgbScalars sp;
    double lp[SGB_CN_TABLE * SGB_MAX_STATES];
    double lq[SGB_CN_TABLE * SGB_MAX_STATES];
    double a[SGB_CN_TABLE * SGB_MAX_STATES];
    double b[SGB_CN_TABLE * SGB_MAX_STATES];
    double l0[SGB_CN_TABLE * SGB_MAX_STATES];
    int kind[SGB_CN_TABLE * SGB_MAX_STATES];
    // beta-binomial only: lgamma(a + k), lgamma(b + k), k < SGB_LGK, row index = cn * S + state (the call's S)
    double lgk[SGB_CN_TABLE * SGB_MAX_STATES * 2 * SGB_LGK];
};

// A "plan" describes the (cell, chromosome) sequences of one batch, sorted by length
// (longest first) and grouped 32 to a warp; see DESIGN.md "HBM layout".
// Packed buffer: group g owns a tile of roundup(T_g, SGB_CH) x 32 records starting at goff[g],
// record (bin t, lane l) at goff[g] + t*32 + l  (bin-major, lanes interleaved: a warp reads 256
// contiguous bytes per bin and one 1-D bulk copy moves SGB_CH bins of all 32 sequences).
#ifndef SGB_CH
#define SGB_CH 4               // bins per staged chunk
#endif
struct SgbPlanDev {
    const long long* goff;     // [ngroups] first record of the group's tile in the packed buffer
    const long long* ostart;   // [nseq] first row of the sequence in the caller's n/x/cn/state arrays
    const int* len;            // [nseq] number of bins
    long long nseq;
    int ngroups;               // ceil(nseq / 32)
};

struct SgbModelArgs {
    int model, lik_bb, S, decode;
    double loherror, rho;
};

// error bits raised by k_pack (validated on the device so the host never touches the counts)
#define SGB_ERR_NEGATIVE_COUNT 1
#define SGB_ERR_COUNT_TOO_LARGE 2
#define SGB_ERR_BAD_CN 4
#define SGB_ERR_NOT_INTEGER 8          // a double-typed column holds NaN or a non-integer value
#define SGB_ERR_INTERNAL 16            // a scratch-size bound was violated (a bug, never an input problem)
#define SGB_ERR_BAD_KEY 32             // a cell / chromosome code outside its declared range, or a negative start
#define SGB_ERR_KEY_CONFLICT 64        // rows with the same (cell, chr, start) disagree on `state`

cudaError_t sgbk_setup(SgbDevTables* d_tabs, const SgbModelArgs& m, double eps, double Abias, double stp,
                       cudaStream_t st);
cudaError_t sgbk_pack(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize,
                      const SgbPlanDev& plan, unsigned long long* d_packed, int* d_err, cudaStream_t st);
// the caller's columns as k_decode's RAW staging reads them (rows = length of the columns, err = device error word)
struct SgbRawIn {
    const int32_t* n;
    const int32_t* x;
    const void* cn;
    int cn_itemsize;
    long long rows;
    int* err;
};
// int32 cn column -> validated uint8 copy (d_cn 16-byte aligned, d_out 4-byte aligned)
cudaError_t sgbk_narrow_cn(const int32_t* d_cn, long long rows, uint8_t* d_out, int* d_err, cudaStream_t st);
// stage the decode will really use for this model / these columns: 2 = RAW (no k_pack needed), 1 = TMA, 0 = LDG
int sgbk_resolve_stage(const SgbModelArgs& m, int stage, const SgbRawIn& raw);
cudaError_t sgbk_decode(const SgbPlanDev& plan, const unsigned long long* d_packed, const SgbRawIn& raw,
                        const SgbDevTables* d_tabs, const SgbModelArgs& m, uint8_t* d_state_out, uint8_t* d_bp,
                        int* d_last_state, unsigned int* d_counter, int sm_count, int stage, cudaStream_t st);
// bytes of backpointers kept per bin by decode = chained: one 16-byte record up to 16 states, two beyond
static __host__ __device__ inline int sgb_bp_stride(int S) { return S <= 16 ? 16 : 32; }
cudaError_t sgbk_backtrace(const SgbPlanDev& plan, const uint8_t* d_bp, const int* d_last_state, int S,
                           uint8_t* d_state_out, cudaStream_t st);
cudaError_t sgbk_loglik(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize,
                        const SgbPlanDev& plan, const SgbDevTables* d_tabs, const SgbModelArgs& m,
                        double* d_l, cudaStream_t st);
cudaError_t sgbk_viterbi(const double* d_emission, long long nrow_e, const double* d_trans, const int32_t* d_obs,
                         long long T, int S, int init_mode, int decode, uint8_t* d_bp, int32_t* d_path,
                         cudaStream_t st);
cudaError_t sgbk_logspace_add(const double* d_a, const double* d_b, long long n, double* d_out, cudaStream_t st);
cudaError_t sgbk_math_eval(int fn, const double* d_in, const double* d_in2, long long n, double* d_out,
                           cudaStream_t st);
cudaError_t sgbk_postprocess(const void* d_cn, int cn_itemsize, const uint8_t* d_path, long long N, int32_t* d_A,
                             int32_t* d_B, int32_t* d_state_min, uint8_t* d_loh, uint8_t* d_phase,
                             uint8_t* d_state_phase, double* d_state_baf, cudaStream_t st);
cudaError_t sgbk_pass_stats(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize,
                            const uint8_t* d_path, long long N, int S, unsigned long long* d_out, cudaStream_t st);
size_t sgbk_viterbi_bp_bytes(long long T, int S);

// ---- split schedule (split_kernels.cuh): k_emit (lanes over bins) + k_recur (lane per sequence, TMA-fed) ----
#define SGB_IN_I32_U8 0                   // n, x int32; cn uint8
#define SGB_IN_I32_I32 1                  // n, x, cn int32 (R integer columns)
#define SGB_IN_U16_U8 2                   // n, x uint16; cn uint8 (host-narrowed staging)
#define SGB_IN_F64 3                      // n, x, cn double (R numeric columns); must be integer valued
struct SgbColsIn {
    const void* n;
    const void* x;
    const void* cn;
    int kind;
    long long rows;
    int* err;
};
struct SgbSplitDev {
    uint8_t* cls;              // [nseq] planes needed by host-plan sequence i (0 = empty)
    int* perm;                 // [gcap * 32] partitioned position -> host-plan sequence, -1 = empty lane
    unsigned int* part;        // [ntile * 9] per-tile class histograms, then bases
    int* g_npl;                // [gcap]
    int* g_tmax;               // [gcap]
    long long* g_eoff;         // [gcap] first 32-byte sector of the group's tile
    unsigned long long* g_size;// [gcap] tile size in sectors
    int* ngroups;              // [1]
    unsigned int* counters;    // [2] dynamic work counters (k_emit, k_recur)
    double* E;                 // emission tiles
    long long e_cap_sectors;
    int gcap;
    int ntile;
};
// scratch sizes for a range of `nseq` plan entries whose lengths sum to `rows` and whose longest has `maxlen` bins
struct SgbSplitSizes {
    size_t cls, perm, part, g_i32, g_i64, e_bytes;
    int gcap, ntile;
};
SgbSplitSizes sgbk_split_sizes(long long nseq, long long rows, int maxlen, int S);
// plan entries [0, nseq) of `plan` (len-descending order); d_tabs must already hold the call's constants
cudaError_t sgbk_split_decode(const SgbPlanDev& plan, long long nseq, const SgbColsIn& in, const SgbDevTables* d_tabs,
                              const SgbModelArgs& m, const SgbSplitDev& W, uint8_t* d_state_out, int sm_count,
                              cudaStream_t st, cudaEvent_t ev_prep, cudaEvent_t ev_mid);
void sgb_internal_set_error(const char* msg);      // sets the calling thread's sgb_last_error() text (capi.cu)
cudaError_t sgbk_fp64_peak(int iters, int sm_count, cudaStream_t st, double* tflops);

// ---- table passes (table_kernels.cu): sort / segment, combineBAFCN reduce, between-pass sums, qc_per_cell, fix_assignments
size_t sgbk_table_scratch_bytes(long long n, int ncell, int nchr);
cudaError_t sgbk_table(long long n, const int32_t* d_cell, const int32_t* d_chr, const int32_t* d_start, const int32_t* d_state,
                       const int32_t* d_a, const int32_t* d_b, const int32_t* d_n, int ncell, int nchr, int do_combine, double filtern,
                       int minbins, int minbinschr, int chrX, void* d_scratch, int32_t* o_first, int32_t* o_cell, int32_t* o_chr,
                       int32_t* o_start, int32_t* o_state, int32_t* o_a, int32_t* o_b, int32_t* o_n, long long* o_seg_off,
                       long long* h_counts, int* d_err, cudaStream_t st);
int sgbk_pass_stats2_blocks(long long N);
cudaError_t sgbk_pass_stats2(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize, const uint8_t* d_path,
                             const uint8_t* d_rowkeep, long long N, int S, int selA, int selB, double* d_part, double* d_out9,
                             unsigned long long* d_hist, cudaStream_t st);
size_t sgbk_qc_scratch_bytes(long long n);
cudaError_t sgbk_qc_per_cell(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize, const uint8_t* d_path,
                             const int32_t* d_cell_of_row, long long N, int ncell, void* d_scratch, double* d_avg, int* d_err, cudaStream_t st);
cudaError_t sgbk_fix_assignments(const int32_t* d_state, const int32_t* d_A, const int32_t* d_B, const int32_t* d_alleleA,
                                 const int32_t* d_alleleB, const uint8_t* d_seqhead, long long N, double loherror, int lik_bb, double rho,
                                 int32_t* d_A1, int32_t* d_B1, int32_t* d_A2, int32_t* d_B2, void* d_scratch, cudaStream_t st);
size_t sgbk_fix_scratch_bytes(long long N);
