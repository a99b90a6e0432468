// hmm_kernels.cu -- sm_100a kernels for the signals HSCN/ASCN HMM hot path.
//
//   k_setup        per-call constants (log p, log1p(-p), beta-binomial shapes, transitions) on the device
//   k_decode       fused emission + max-plus forward recursion + the reference's decode rule
//                  (one lane per (cell, chromosome) sequence, 32 sequences per warp); reads the caller's
//                  columns in place (binomial models) or k_pack's records
//   k_narrow_cn    int32 copy-number column -> validated uint8 copy (so R-style input can be read in place)
//   k_pack         caller's flat int32/uint8 columns -> 8-byte records, sequences sorted by length
//                  (beta-binomial, chained decode, unaligned columns)
//   k_backtrace    opt-in textbook backtrace over uint8 backpointers (decode = chained)
//   k_loglik       bins x states log-likelihood matrix `l` (HaplotypeHMM()$l, alleleHMM()$l)
//   k_viterbi      stand-alone viterbi(emission, transition, observations), dense S x S transition
//   k_logspace_add, k_math_eval, k_postprocess   small element-wise kernels
//
// Reference functions replaced: src/code.cpp:5-45, R/callHSCN.R:36-108,135-172,217-295,
// R/callASCN.R:41-103,128-161 (see hmm_core.cuh for the arithmetic and DESIGN.md for the design).
#include "kernels.h"

__device__ const double g_logtab[SGB_LOGTAB_LEN] = SGB_LOGTAB_INIT;
__device__ const double g_exptab[SGB_EXPTAB_LEN] = SGB_EXPTAB_INIT;
__device__ const double g_lftab[SGB_LFTAB_LEN] = SGB_LFTAB_INIT;

#define SGB_FULL 0xffffffffu
#define SGB_COUNT_MASK ((1ull << SGB_COUNT_BITS) - 1ull)

// ------------------------------------------------------------------ k_setup
__global__ void k_setup(SgbDevTables* T, SgbModelArgs m, double eps, double Abias, double stp)
{
    __shared__ double s_bb;
    if (threadIdx.x == 0) {
        SgbScalars sp = sgb_scalars(m.S, eps, Abias, stp, m.lik_bb ? m.rho : 0.0, g_logtab);
        T->sp = sp;
        s_bb = sp.bb_s;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < SGB_CN_TABLE * SGB_MAX_STATES; idx += blockDim.x) {
        int c = idx / SGB_MAX_STATES, s = idx % SGB_MAX_STATES;
        SgbStateConst sc = sgb_state_const(c, s, m.loherror, m.lik_bb ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL,
                                           m.rho, s_bb, g_logtab);
        T->lp[idx] = sc.lp; T->lq[idx] = sc.lq; T->a[idx] = sc.a; T->b[idx] = sc.b; T->l0[idx] = sc.l0;
        T->kind[idx] = sc.kind;
        if (m.lik_bb && s < m.S && sc.kind == SGB_KIND_NORMAL) {
            double* row = T->lgk + (size_t)(c * m.S + s) * 2 * SGB_LGK;
            for (int k = 0; k < SGB_LGK; k++) {
                row[k] = sgb_lgamma_pos(sgb_add(sc.a, (double)k), g_logtab);
                row[SGB_LGK + k] = sgb_lgamma_pos(sgb_add(sc.b, (double)k), g_logtab);
            }
        }
    }
}

cudaError_t sgbk_setup(SgbDevTables* d_tabs, const SgbModelArgs& m, double eps, double Abias, double stp,
                       cudaStream_t st)
{
    k_setup<<<1, 256, 0, st>>>(d_tabs, m, eps, Abias, stp);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ k_pack
// Caller's flat columns -> 8-byte records (n:28 | x:28 | cn:8) in the group-tiled, lane-interleaved
// layout k_decode streams (see SgbPlanDev).  One warp per group of 32 sequences; per block of 32 bins it
// reads each sequence's n / x / cn with lanes over consecutive bins (coalesced 128 B / 32 B segments),
// transposes through a padded shared-memory tile and writes 256 B per bin, lanes over sequences.
// Counts are validated here, on the device, so the host never touches the per-bin data.
__global__ void __launch_bounds__(128) k_pack(const int32_t* __restrict__ n, const int32_t* __restrict__ x,
                                              const void* __restrict__ cn, int cn_itemsize, SgbPlanDev plan,
                                              unsigned long long* __restrict__ packed, int* err)
{
    __shared__ unsigned long long tile[4][32][33];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // work item = (group, block of 32 bins); items are enumerated group-major by a per-warp walk
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    int bad = 0;
    for (long long g = warp; g < plan.ngroups; g += nwarps) {
        const long long r0 = g * 32;
        // lane j keeps row j's length and start; the row loop broadcasts them with shuffles
        const bool has = r0 + lane < plan.nseq;
        const int my_len = has ? plan.len[r0 + lane] : 0;
        const long long my_o = has ? plan.ostart[r0 + lane] : 0;
        const int Tg = __shfl_sync(SGB_FULL, my_len, 0);
        const int Tpad = ((Tg + SGB_CH - 1) / SGB_CH) * SGB_CH;
        const long long base = plan.goff[g];
        for (int t0 = 0; t0 < Tpad; t0 += 32) {
            const int t = t0 + lane;
            // all 96 loads of the block are issued before any is consumed (the kernel is latency-, not issue-bound)
            int nv[32], xv[32], cv[32];
            unsigned okmask = 0;
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const int lenj = __shfl_sync(SGB_FULL, my_len, j);
                const long long o = __shfl_sync(SGB_FULL, my_o, j) + t;
                const bool ok = t < lenj;
                okmask |= ok ? (1u << j) : 0u;
                nv[j] = ok ? __ldg(n + o) : 0;
                xv[j] = ok ? __ldg(x + o) : 0;
                cv[j] = ok ? ((cn_itemsize == 1) ? (int)__ldg((const uint8_t*)cn + o) : __ldg((const int32_t*)cn + o)) : 0;
            }
#pragma unroll
            for (int j = 0; j < 32; j++) {
                if (nv[j] < 0 || xv[j] < 0) bad |= SGB_ERR_NEGATIVE_COUNT;
                if (nv[j] >= (1 << SGB_COUNT_BITS) || xv[j] >= (1 << SGB_COUNT_BITS)) bad |= SGB_ERR_COUNT_TOO_LARGE;
                if (cv[j] < 0 || cv[j] > 255) bad |= SGB_ERR_BAD_CN;
                const unsigned long long rec = ((unsigned long long)(unsigned)nv[j] & SGB_COUNT_MASK) |
                                               (((unsigned long long)(unsigned)xv[j] & SGB_COUNT_MASK) << SGB_COUNT_BITS) |
                                               ((unsigned long long)(cv[j] & 255) << 56);
                tile[wid][j][lane] = ((okmask >> j) & 1u) ? rec : 0ull;
            }
            __syncwarp();
            const int nb = (Tpad - t0) < 32 ? (Tpad - t0) : 32;
#pragma unroll 8
            for (int jb = 0; jb < nb; jb++) packed[base + (long long)(t0 + jb) * 32 + lane] = tile[wid][lane][jb];
            __syncwarp();
        }
    }
    if (bad) atomicOr(err, bad);
}

cudaError_t sgbk_pack(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize,
                      const SgbPlanDev& plan, unsigned long long* d_packed, int* d_err, cudaStream_t st)
{
    if (plan.nseq == 0) return cudaSuccess;
    long long blocks = ((long long)plan.ngroups + 3) / 4;
    if (blocks > 148 * 6) blocks = 148 * 6;
    k_pack<<<(unsigned)blocks, 128, 0, st>>>(d_n, d_x, d_cn, cn_itemsize, plan, d_packed, d_err);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ k_narrow_cn
// int32 copy-number column (what R hands over) -> uint8, validated, so that k_decode can read the columns in
// place: 5 B per cell-bin of traffic instead of k_pack's 17.
__global__ void __launch_bounds__(256) k_narrow_cn(const int32_t* __restrict__ cn, long long rows,
                                                   uint8_t* __restrict__ out, int* err)
{
    const long long nvec = rows >> 2;
    int bad = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        const int4 v = __ldcs((const int4*)cn + i);
        bad |= (v.x | v.y | v.z | v.w) & ~255;
        ((uint32_t*)out)[i] = (uint32_t)(v.x & 255) | ((uint32_t)(v.y & 255) << 8) | ((uint32_t)(v.z & 255) << 16) |
                              ((uint32_t)(v.w & 255) << 24);
    }
    if (blockIdx.x == 0 && threadIdx.x < (rows & 3)) {
        const int v = cn[(nvec << 2) + threadIdx.x];
        bad |= v & ~255;
        out[(nvec << 2) + threadIdx.x] = (uint8_t)v;
    }
    if (bad) atomicOr(err, SGB_ERR_BAD_CN);
}

cudaError_t sgbk_narrow_cn(const int32_t* d_cn, long long rows, uint8_t* d_out, int* d_err, cudaStream_t st)
{
    if (rows <= 0) return cudaSuccess;
    long long blocks = ((rows >> 2) + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    k_narrow_cn<<<(unsigned)blocks, 256, 0, st>>>(d_cn, rows, d_out, d_err);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ k_decode
struct SgbDecodeArgs {
    SgbPlanDev plan;
    const unsigned long long* packed;
    const SgbDevTables* tabs;
    uint8_t* state_out;
    uint8_t* bp;            // chained only: [(row)*S + state]
    int* last_state;        // chained only: [nseq]
    unsigned int* counter;  // dynamic group scheduler
    int S, lik;
    double loherror, rho;
    // SGB_STAGE_RAW only: the caller's columns, read in place (no k_pack pass)
    const int32_t* n;
    const int32_t* x;
    const void* cn;
    long long rows;                     // length of the columns: copies are clamped to it
    int* err;
};

#ifndef SGB_DECODE_THREADS
#define SGB_DECODE_THREADS 256
#endif
#define SGB_DECODE_WARPS (SGB_DECODE_THREADS / 32)
#ifndef SGB_LF_SMEM
#define SGB_LF_SMEM 1024         // log-factorial entries cached in shared memory (the rest of the 2048 stay in L1/L2)
#endif
#ifndef SGB_DECODE_MINBLOCKS
#define SGB_DECODE_MINBLOCKS 2   // resident CTAs per SM the register allocation is bounded for
#endif
#ifndef SGB_BB_PAIR
#define SGB_BB_PAIR 0            // beta-binomial kernels evaluate ONE item per lane per round: the two-item body overflowed the
                                 // instruction cache (icc hit rate 78 %, fetch stalls dominant; 5.18 -> 3.68 ms on ASCN beta-binomial)
#endif
#define SGB_PAIR_OK(likbb) (!(likbb) || SGB_BB_PAIR)
#define SGB_STAGE_LDG 0        // records read straight from global memory (one LDG.64 per lane per bin)
#define SGB_STAGE_TMA 1        // records staged through shared memory by 1-D bulk copies (cp.async.bulk + mbarrier)
#define SGB_STAGE_RAW 2        // no packed records at all: every lane streams its own sequence out of the caller's n / x /
                               // cn columns in aligned 4-bin blocks (16-byte cp.async copies) through a 3-block ring in
                               // shared memory (binomial models, uint8 cn, 16-byte aligned columns)
#define SGB_RAW_WORDS (3 * 32 * 9)      // ring: 3 blocks x 32 lanes x (4 words n + 4 words x + 1 word cn)

// src_bytes < 16 zero-fills the rest of the destination (used at the end of the caller's columns)
__device__ __forceinline__ void sgb_cp_async16(void* dst, const void* src, unsigned src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void sgb_cp_async4(void* dst, const void* src, unsigned src_bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void sgb_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void sgb_cp_async_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
// log-factorial entries cached in shared memory: the raw ring needs the room, larger counts read the L1/L2 copy
__host__ __device__ constexpr int sgb_lf_smem(int model, int stage)
{
    return stage == SGB_STAGE_RAW ? (model == SGB_MODEL_ASCN ? 256 : 512) : SGB_LF_SMEM;
}

#include "sgb_async.cuh"

// per-warp shared-memory scratch, in doubles (see the pointer carving in k_decode)
//   E  [32][P]  emissions of the current bin, lane-major with odd pitch P = SMAX+1 (conflict-free columns)
//   T1 [32][P]  forward column of every lane's sequence
//   bD/bL/bG    per-lane bin fields for the item queue ({x, n-x}, log-choose terms, lgamma(s+n))
//   bI [32]     cprime | in_support << 16
//   q           work queue, SMAX*32 uint32 entries (pre-scaled offsets, see below)
//   stage       2 x SGB_CH x 32 records + 2 mbarriers (TMA staging only)
static inline size_t decode_warp_doubles(int SMAX, int model, int lik_bb, int stage)
{
    const size_t P = (size_t)SMAX + 1;
    size_t d = 2 * 32 * P + 16 + (size_t)SMAX * 16;
    d += 64 + (model == SGB_MODEL_ASCN ? 64 : 32) + (lik_bb ? 32 : 0);
    if (stage == SGB_STAGE_TMA) d += 2 * 32 * SGB_CH + 2;
    if (stage == SGB_STAGE_RAW) d += SGB_RAW_WORDS / 2;
    return (d + 1) & ~(size_t)1;
}
static inline size_t decode_smem_bytes(int SMAX, int S, int model, int lik_bb, int stage)
{
    size_t d = SGB_LOGTAB_LEN + SGB_EXPTAB_LEN + sgb_lf_smem(model, stage);
    d += (size_t)SGB_CN_TABLE * S * (lik_bb ? 5 : 2);
    d += (size_t)SGB_DECODE_WARPS * decode_warp_doubles(SMAX, model, lik_bb, stage);
    return d * sizeof(double) + ((size_t)SGB_CN_TABLE * S + SGB_CN_TABLE) * sizeof(int);
}

// Rare variant of the recursion (selftransitionprob below 1/S, so tr_self < tr_off), kept out of line so it
// does not sit in the hot loop's instruction footprint.  Returns the decoded state of the previous bin.
// Its two running maxima live in the padding slot of the lane's T1 / E columns (index SMAX), the argmax travels
// in the return value (back | i1 << 8), so the hot path carries no extra registers for it.
template <int SMAX>
__device__ __noinline__ int decode_step_general(double* colT, double* colE, int S, double init, double tr_self,
                                                double tr_off, int i1_in, int first)
{
    SgbSeqState<SMAX> st;
    double e[SMAX];
#pragma unroll
    for (int i = 0; i < SMAX; i++) { e[i] = colE[i]; st.T1[i] = colT[i]; }
    st.m1 = colT[SMAX]; st.m2 = colE[SMAX]; st.i1 = i1_in;
    int back = 0;
    if (first) sgb_seq_init<SMAX>(st, e, S, init);
    else back = sgb_seq_step_general<SMAX>(st, e, S, tr_self, tr_off);
#pragma unroll
    for (int i = 0; i < SMAX; i++) colT[i] = st.T1[i];
    colT[SMAX] = st.m1; colE[SMAX] = st.m2;
    return back | (st.i1 << 8);
}

// Fused emission + forward recursion + decode.  One lane per (cell, chromosome) sequence, 32 sequences
// of similar length per warp, warps fetch groups from a global counter (longest first).
// The kernel is bound by instruction issue and shared-memory bandwidth, so per bin step a warp
//  (0) takes each lane's n / x / cn from shared memory -- STAGE RAW: a 3-block ring that every lane fills from
//      its own rows of the caller's columns with aligned 16-byte cp.async copies; STAGE TMA: packed records
//      moved by 1-D bulk copies one chunk (SGB_CH bins x 32 sequences, 2 KB) ahead, an mbarrier per stage;
//      STAGE LDG: packed records by plain coalesced loads;
//  (1) looks the log-factorials up and publishes the lane's per-bin terms ({x, n-x}, log-choose, lgamma(s+n));
//  (2) keeps a queue of all live (lane, state) pairs, compacted across the 32 lanes by a prefix sum -- a lane
//      with copy number 2 does not idle while a lane with copy number 9 works -- and rebuilt only on bins
//      where some lane's copy number changed;
//  (3) evaluates the queue, two entries per lane per round (one for beta-binomial), into E[lane][state];
//  (4) each lane runs the O(S) max-plus step of its sequence over the live states only and stores the
//      decoded state of the previous bin.
template <int SMAX, int MODEL, int LIKBB, int CHAINED, int STAGE>
__global__ void __launch_bounds__(SGB_DECODE_THREADS, SGB_DECODE_MINBLOCKS) k_decode(SgbDecodeArgs A)
{
    extern __shared__ __align__(16) double smem[];
    const int S = A.S;
    const int CTS = SGB_CN_TABLE * S;
    constexpr int P = SMAX + 1;
    constexpr int NL = (MODEL == SGB_MODEL_ASCN) ? 64 : 32;
    constexpr int WD0 = 2 * 32 * P + 16 + SMAX * 16 + 64 + NL + (LIKBB ? 32 : 0) +
                        (STAGE == SGB_STAGE_TMA ? 2 * 32 * SGB_CH + 2 : 0) + (STAGE == SGB_STAGE_RAW ? SGB_RAW_WORDS / 2 : 0);
    constexpr int WARP_DOUBLES = (WD0 + 1) & ~1;
    constexpr int LFN = sgb_lf_smem(MODEL, STAGE);
    double* sLT = smem;
    double* sET = sLT + SGB_LOGTAB_LEN;
    double* sLF = sET + SGB_EXPTAB_LEN;
    double* sLP = sLF + LFN;
    double* sLQ = sLP + CTS;
    double* sA = sLQ + CTS;
    double* sB = sA + (LIKBB ? CTS : 0);
    double* sL0 = sB + (LIKBB ? CTS : 0);
    double* sW = sL0 + (LIKBB ? CTS : 0);
    int* sKind = (int*)(sW + SGB_DECODE_WARPS * WARP_DOUBLES);
    int* sKM = sKind + CTS;                 // per total-CN bit mask of the states on the lean path

    for (int i = threadIdx.x; i < SGB_LOGTAB_LEN; i += blockDim.x) sLT[i] = g_logtab[i];
    for (int i = threadIdx.x; i < SGB_EXPTAB_LEN; i += blockDim.x) sET[i] = g_exptab[i];
    for (int i = threadIdx.x; i < LFN; i += blockDim.x) sLF[i] = g_lftab[i];
    for (int i = threadIdx.x; i < CTS; i += blockDim.x) {
        int c = i / S, s = i - c * S, src = c * SGB_MAX_STATES + s;
        sLP[2 * i] = A.tabs->lp[src];           // {log p, log(1-p)} pairs: one 16-byte load per item
        sLP[2 * i + 1] = A.tabs->lq[src];
        sKind[i] = A.tabs->kind[src];
        if (LIKBB) { sA[i] = A.tabs->a[src]; sB[i] = A.tabs->b[src]; sL0[i] = A.tabs->l0[src]; }
    }
    if (threadIdx.x < SGB_CN_TABLE) {
        int mask = 0;
        for (int s = 0; s < S; s++)
            if (A.tabs->kind[threadIdx.x * SGB_MAX_STATES + s] == SGB_KIND_NORMAL) mask |= (int)(1u << s);
        sKM[threadIdx.x] = mask;
    }
    __shared__ SgbScalars s_sp;             // read-mostly scalars live in shared memory, not in registers
    __shared__ SgbTables s_TB;
    if (threadIdx.x == 0) {
        s_sp = A.tabs->sp;
        s_TB.LT = sLT; s_TB.ET = sET; s_TB.LP = sLP; s_TB.LQ = sLP + 1; s_TB.pstride = 2; s_TB.A = sA; s_TB.B = sB; s_TB.L0 = sL0;
        s_TB.LGK = A.tabs->lgk; s_TB.KIND = sKind; s_TB.S = S; s_TB.lik = A.lik; s_TB.loherror = A.loherror; s_TB.rho = A.rho;
    }
    const SgbScalars& sp = s_sp;
    const SgbTables& TB = s_TB;

    const int lane = threadIdx.x & 31;
    const int wid = threadIdx.x >> 5;
    double* myE = sW + wid * WARP_DOUBLES;                    // E [lane][P]
    double* myT = myE + 32 * P;                               // T1[lane][P]
    double2* bD = (double2*)(myT + 32 * P);                   // {x, n-x}
    double* bL = (double*)(bD + 32);                          // HSCN lcx[32]; ASCN {lcx, lcy}[32]
    double* bG = bL + NL;                                     // beta-binomial: lgamma(s+n)[32]
    int* bI = (int*)(bG + (LIKBB ? 32 : 0));                  // cprime | in_support << 16
    unsigned int* q = (unsigned int*)(bI + 32);
    unsigned long long* stg = (unsigned long long*)(q + SMAX * 32);     // [2][SGB_CH][32] records
    unsigned long long* bar = stg + 2 * 32 * SGB_CH;                    // [2] mbarriers
    unsigned int* sraw = (unsigned int*)stg;                            // RAW ring: n [3][32][4], x [3][32][4], cn words [3][32]
    if (STAGE == SGB_STAGE_TMA && lane == 0) {
        sgb_mbar_init(&bar[0], 1);
        sgb_mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncthreads();

    const double NEG_INF = -sgb_inf();
    const double* gLGK = A.tabs->lgk;       // small-count lgamma cache (global memory, L1-resident: ~12 KB live)
    double* colE = myE + lane * P;          // this lane's own columns
    double* colT = myT + lane * P;
    unsigned use0 = 0, use1 = 0;            // completed phases of the two stage barriers

    for (;;) {
        int g = 0;
        if (lane == 0) g = (int)atomicAdd(A.counter, 1u);
        g = __shfl_sync(SGB_FULL, g, 0);
        if (g >= A.plan.ngroups) break;
        const long long r = (long long)g * 32 + lane;
        int len = 0;
        long long ostart = 0;
        if (r < A.plan.nseq) {
            len = A.plan.len[r];
            ostart = A.plan.ostart[r];
        }
        const unsigned long long* gtile = (STAGE == SGB_STAGE_RAW) ? nullptr : A.packed + A.plan.goff[g];   // [bin][lane] records
        const int Tmax = __shfl_sync(SGB_FULL, len, 0);                   // lane 0 holds the group's longest sequence
        double m1 = NEG_INF;            // running column maximum of this lane's sequence
        double m2 = NEG_INF;            // second maximum / argmax: only the general and chained variants use them
        int i1 = 0;
        unsigned long long rec_next = 0ull;
        // one 1-D bulk copy moves chunk c (SGB_CH bins x 32 lanes, contiguous in the tile) into stage c&1
        auto issue_chunk = [&](int c) {
            if (lane == 0) {
                unsigned long long* b = &bar[c & 1];
                sgb_mbar_expect_tx(b, SGB_CH * 32 * 8);
                sgb_tma_load_1d(stg + (c & 1) * 32 * SGB_CH, gtile + (long long)c * SGB_CH * 32, SGB_CH * 32 * 8, b);
            }
        };
        // RAW: block B of a lane = rows [a0 + 4B, a0 + 4B + 4) of the caller's columns, a0 = ostart rounded down to a
        // multiple of 4, so every copy is 16-byte aligned whatever the sequence's first row is (the rows before ostart
        // belong to the previous sequence and are ignored).  Ring slot B % 3; the bin a lane reads at step t sits in block
        // (sh + t) >> 2, sh = ostart & 3, i.e. block t/4 or t/4 + 1: two blocks resident, a third in flight.  Every lane
        // commits exactly one cp.async group per block, empty or not.
        const int sh = (int)(ostart & 3);
        const long long a0 = ostart - sh;
        const int span = sh + len;                                   // rows of this lane's sequence, counted from a0
        const long long tail = A.rows - a0;                          // rows of the columns, counted from a0
        const int avail = tail > 0x7fffffffll ? 0x7fffffff : (int)tail;
        auto issue_raw = [&](int B, int slot) {
            if (4 * B < span) {
                const int left = avail - 4 * B;                      // > 0: the sequence lies inside the columns
                const long long row = a0 + 4ll * B;
                unsigned int* d = sraw + (slot * 32 + lane) * 4;
                sgb_cp_async16(d, A.n + row, left >= 4 ? 16u : (unsigned)left * 4u);
                sgb_cp_async16(d + 3 * 32 * 4, A.x + row, left >= 4 ? 16u : (unsigned)left * 4u);
                sgb_cp_async4(sraw + 2 * 3 * 32 * 4 + slot * 32 + lane, (const uint8_t*)A.cn + row, left >= 4 ? 4u : (unsigned)left);
            }
            sgb_cp_async_commit();
        };
        int slot_m = 0;            // ring slot of block t/4
        if (STAGE == SGB_STAGE_TMA) {
            if (Tmax > 0) issue_chunk(0);
        } else if (STAGE == SGB_STAGE_RAW) {
            if (Tmax > 0) { issue_raw(0, 0); issue_raw(1, 1); }
        } else {
            rec_next = (Tmax > 0) ? __ldcs(gtile + lane) : 0ull;
        }
        int back = 0;
        int Lprev = SMAX;          // states >= L hold -Inf in this lane's E column; refreshed when L shrinks
        int Lw_prev = SMAX;        // warp-wide maximum live-state count of the previous bin
        int Lw = 0, tot = 0;       // ... of the current bin, and the length of the item queue (both kept while it is reused)
        int prev_key = -1;         // (copy number, validity, active) the queue entries of this lane were built from
        for (int t = 0; t < Tmax; t++) {
            const bool act = t < len;
            if (STAGE == SGB_STAGE_RAW && (t & 3) == 0) {
                if (t) slot_m = slot_m == 2 ? 0 : slot_m + 1;
                issue_raw((t >> 2) + 2, slot_m == 0 ? 2 : slot_m - 1);      // into the slot block t/4 - 1 has left
                sgb_cp_async_wait1();       // all but the group just committed have landed: blocks t/4 and t/4 + 1
                __syncwarp();
            }
            unsigned long long rec = 0ull;      // packed-record stages only
            if (STAGE == SGB_STAGE_TMA) {
                const int ci = t / SGB_CH, w = t % SGB_CH;
                if (w == 0) {
                    if (ci & 1) { sgb_mbar_wait(&bar[1], use1 & 1); use1++; }
                    else { sgb_mbar_wait(&bar[0], use0 & 1); use0++; }
                    if ((ci + 1) * SGB_CH < Tmax) issue_chunk(ci + 1);
                }
                rec = stg[((ci & 1) * SGB_CH + w) * 32 + lane];
            } else if (STAGE == SGB_STAGE_LDG) {
                rec = rec_next;
                if (t + 1 < Tmax) rec_next = __ldcs(gtile + (long long)(t + 1) * 32 + lane);
            }
            // ---- (1) per-bin quantities of this lane's sequence, published for the item queue: ONE divergent region
            int L = 0, cprime = 0, ins = 0;
            if (act) {
                int n, x, cnv;                  // this lane's bin: total count, B-allele count, total copy number
                if (STAGE == SGB_STAGE_RAW) {
                    const int u = sh + (t & 3);
                    int slot = slot_m + (u >> 2);
                    slot = slot >= 3 ? slot - 3 : slot;
                    const unsigned int* rN = sraw + (slot * 32 + lane) * 4 + (u & 3);
                    n = (int)rN[0]; x = (int)rN[3 * 32 * 4];
                    cnv = (int)((sraw[2 * 3 * 32 * 4 + slot * 32 + lane] >> (8 * (u & 3))) & 255u);
                    // the checks k_pack makes on the packed path; a bad bin is neutralised, the call fails afterwards
                    if ((unsigned)(n | x) >> SGB_COUNT_BITS) {
                        atomicOr(A.err, (n | x) < 0 ? SGB_ERR_NEGATIVE_COUNT : SGB_ERR_COUNT_TOO_LARGE);
                        n = 0; x = 0; cnv = 0;
                    }
                } else {
                    n = (int)(rec & SGB_COUNT_MASK); x = (int)((rec >> SGB_COUNT_BITS) & SGB_COUNT_MASK);
                    cnv = (int)(rec >> 56);
                }
                double lcx = 0.0, lcy = 0.0, lgs = 0.0;
                cprime = sgb_cprime(MODEL, cnv);
                ins = x <= n;
                L = sgb_live_states(MODEL, cprime, S);
                if (ins) {
                    double fn, fx, fy;
                    if (n < LFN) { fn = sLF[n]; fx = sLF[x]; fy = sLF[n - x]; }
                    else if (n < SGB_NLF) { fn = g_lftab[n]; fx = g_lftab[x]; fy = g_lftab[n - x]; }
                    else { fn = sgb_logfact(n, g_lftab, sLT); fx = sgb_logfact(x, g_lftab, sLT); fy = sgb_logfact(n - x, g_lftab, sLT); }
                    lcx = sgb_sub(sgb_sub(fn, fx), fy);
                    if (MODEL == SGB_MODEL_ASCN) lcy = sgb_sub(sgb_sub(fn, fy), fx);
                    if (LIKBB) lgs = sgb_lgamma_pos(sgb_add(sp.bb_s, (double)n), sLT);
                }
                for (int i = L; i < Lprev; i++) colE[i] = NEG_INF;      // rare: only when L shrinks
                Lprev = L;
                bD[lane] = make_double2((double)x, (double)(n - x));
                if (MODEL == SGB_MODEL_ASCN) ((double2*)bL)[lane] = make_double2(lcx, lcy);
                else bL[lane] = lcx;
                if (LIKBB) bG[lane] = lgs;
            }
            // ---- (2) queue of live (lane, state) pairs.  Entry (32 bit): E-tile index lane*P + state (bits 0-11) |
            //      table row cprime*S + state (bits 12-21) | lane (bits 22-26) | lean-path flag (bit 27): every
            //      address an item needs is one shift/mask away, and consecutive states differ by a constant.
            //      A queue entry depends on (lane, copy number, count validity) only, and copy number is piecewise
            //      constant along a chromosome: the queue of the previous bin is reused verbatim unless some lane's
            //      key changed (warp-uniform vote), which skips the scan and the entry stores on most bins.
            const int key = act ? (cprime | (ins << 16) | (1 << 17)) : 0;
            const bool rebuild = __any_sync(SGB_FULL, key != prev_key);
            prev_key = key;
            if (rebuild) {
                Lw = __reduce_max_sync(SGB_FULL, L);
                unsigned need = 0;          // bit m: state m of this lane's bin is queued
                int fastmask = 0;
                if (act) {
                    fastmask = (cprime < SGB_CN_TABLE && ins && sp.fast_ok) ? sKM[cprime] : 0;
                    bI[lane] = cprime | (ins << 16);
                    need = L >= 32 ? 0xffffffffu : (1u << L) - 1u;
                }
                const int cnt = __popc(need);
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    int v = __shfl_up_sync(SGB_FULL, incl, d);
                    if (lane >= d) incl += v;
                }
                tot = __shfl_sync(SGB_FULL, incl, 31);
                const int qpos = incl - cnt;
                const unsigned allfast = ((fastmask & need) == need) ? (1u << 27) : 0u;
                const unsigned trow = (cprime < SGB_CN_TABLE) ? (unsigned)(cprime * S) : 0u;
                const unsigned ebase = (unsigned)(lane * P) | (trow << 12) | ((unsigned)lane << 22) | allfast;
#pragma unroll
                for (int m = 0; m < SMAX; m++)
                    if (m < L) q[qpos + m] = ebase + (unsigned)m * 0x1001u;
                // pad the queue to a multiple of 32 with copies of its first entry (re-evaluating an item and storing
                // the same value twice is harmless), so the rounds below need no per-lane validity predicates
                __syncwarp();
                if (tot > 0 && tot + lane < ((tot + 31) & ~31)) q[tot + lane] = q[0];
                tot = (tot + 31) & ~31;
            }
            __syncwarp();
            // ---- (3) the queue: two entries per lane per round while at least 64 remain, else one
            for (int base = 0; base < tot;) {
                if (SGB_PAIR_OK(LIKBB) && tot - base >= 64) {
                    const int j0 = base + lane, j1 = j0 + 32;
                    const unsigned en0 = q[j0], en1 = q[j1];
                    const int ol0 = (en0 >> 22) & 31, ol1 = (en1 >> 22) & 31;
                    const int e0 = en0 & 0xfff, e1 = en1 & 0xfff;               // E-tile indices
                    const int t0 = (en0 >> 12) & 0x3ff, t1 = (en1 >> 12) & 0x3ff;     // table rows
                    const double2 d0 = bD[ol0], d1 = bD[ol1];
                    double lx0, ly0 = 0.0, lx1, ly1 = 0.0;
                    if (MODEL == SGB_MODEL_ASCN) {
                        const double2 c0 = ((const double2*)bL)[ol0], c1 = ((const double2*)bL)[ol1];
                        lx0 = c0.x; ly0 = c0.y; lx1 = c1.x; ly1 = c1.y;
                    } else {
                        lx0 = bL[ol0]; lx1 = bL[ol1];
                    }
                    const double gs0 = LIKBB ? bG[ol0] : 0.0, gs1 = LIKBB ? bG[ol1] : 0.0;
                    double r0, r1;
                    if (en0 & en1 & (1u << 27)) {
                        const double2 pq0 = ((const double2*)sLP)[t0], pq1 = ((const double2*)sLP)[t1];
                        sgb_emission_fast2<MODEL, LIKBB>(d0.x, d0.y, lx0, ly0, gs0, pq0.x, pq0.y, LIKBB ? sA[t0] : 0.0,
                                                         LIKBB ? sB[t0] : 0.0, LIKBB ? sL0[t0] : 0.0, d1.x, d1.y, lx1, ly1, gs1,
                                                         pq1.x, pq1.y, LIKBB ? sA[t1] : 0.0, LIKBB ? sB[t1] : 0.0,
                                                         LIKBB ? sL0[t1] : 0.0, LIKBB ? gLGK + (long)t0 * 2 * SGB_LGK : sLT,
                                                         LIKBB ? gLGK + (long)t1 * 2 * SGB_LGK : sLT, sp, sLT, sET, r0, r1);
                    } else {
                        const int bi0 = bI[ol0], bi1 = bI[ol1];
                        // rare: keep the hot loop small -- the out-of-line general path handles lean items too (same bits)
                        r0 = sgb_emit_item_slow<MODEL, LIKBB>(bi0 & 0xffff, bi0 >> 16, d0.x, d0.y, lx0, ly0, gs0, e0 - ol0 * P, TB, sp);
                        r1 = sgb_emit_item_slow<MODEL, LIKBB>(bi1 & 0xffff, bi1 >> 16, d1.x, d1.y, lx1, ly1, gs1, e1 - ol1 * P, TB, sp);
                    }
                    myE[e0] = r0;
                    myE[e1] = r1;
                    base += 64;
                } else {
                    const unsigned en0 = q[base + lane];
                    const int ol0 = (en0 >> 22) & 31, e0 = en0 & 0xfff, t0 = (en0 >> 12) & 0x3ff;
                    const double2 d0 = bD[ol0];
                    double lx0, ly0 = 0.0;
                    if (MODEL == SGB_MODEL_ASCN) {
                        const double2 c0 = ((const double2*)bL)[ol0];
                        lx0 = c0.x; ly0 = c0.y;
                    } else {
                        lx0 = bL[ol0];
                    }
                    const double gs0 = LIKBB ? bG[ol0] : 0.0;
                    double r0;
                    if (en0 & (1u << 27)) {
                        const double2 pq0 = ((const double2*)sLP)[t0];
                        r0 = sgb_emission_fast<MODEL, LIKBB>(d0.x, d0.y, lx0, ly0, gs0, pq0.x, pq0.y, LIKBB ? sA[t0] : 0.0,
                                                             LIKBB ? sB[t0] : 0.0, LIKBB ? sL0[t0] : 0.0,
                                                             LIKBB ? gLGK + (long)t0 * 2 * SGB_LGK : sLT, sp, sLT, sET);
                    } else {
                        const int bi0 = bI[ol0];
                        r0 = sgb_emit_item_slow<MODEL, LIKBB>(bi0 & 0xffff, bi0 >> 16, d0.x, d0.y, lx0, ly0, gs0, e0 - ol0 * P, TB, sp);
                    }
                    myE[e0] = r0;
                    base += 32;
                }
            }
            __syncwarp();
            // ---- (4) forward recursion + decode of this lane's sequence (T1 column lives in shared memory)
            const int nlive = (t == 0) ? SMAX : (Lw > Lw_prev ? Lw : Lw_prev);
            Lw_prev = Lw;
            if (act) {
                if (CHAINED == 0 && sp.diag_dominant) {
                    if (t == 0) {
                        sgb_seq_init_mem<SMAX>(colT, 1, colE, 1, S, sp.init, m1);
                        if (len == 1) A.state_out[ostart] = 0;
                    } else {
                        back = sgb_seq_step_mem<SMAX>(colT, 1, colE, 1, m1, sp.tr_self, sp.tr_off, nlive);
                        A.state_out[ostart + t - 1] = (uint8_t)back;
                        if (t == len - 1) A.state_out[ostart + t] = (uint8_t)back;
                    }
                } else if (CHAINED == 0) {
                    // general transition (selftransitionprob < 1/S)
                    const int bi = decode_step_general<SMAX>(colT, colE, S, sp.init, sp.tr_self, sp.tr_off, i1, t == 0);
                    back = bi & 0xff; i1 = bi >> 8;
                    if (t == 0) {
                        if (len == 1) A.state_out[ostart] = 0;
                    } else {
                        A.state_out[ostart + t - 1] = (uint8_t)back;
                        if (t == len - 1) A.state_out[ostart + t] = (uint8_t)back;
                    }
                } else {
                    // chained decode: register-resident step, backpointers of every state to global scratch
                    SgbSeqState<SMAX> st;
                    double e[SMAX];
#pragma unroll
                    for (int i = 0; i < SMAX; i++) { e[i] = colE[i]; st.T1[i] = colT[i]; }
                    st.m1 = m1; st.m2 = m2; st.i1 = i1;
                    if (t == 0) {
                        sgb_seq_init<SMAX>(st, e, S, sp.init);
                    } else {
                        unsigned char bpv[SMAX] = {};
                        sgb_seq_step_chained<SMAX>(st, e, S, sp.tr_self, sp.tr_off, bpv, nlive);
                        // the bin's backpointers as ONE 16-byte record (two for more than 16 states): a byte store per state
                        // and lane was 12 x 32 scattered sectors per warp and bin, and the kernel ran at 7.6 ms for 57 M bins
                        constexpr int BPW = SMAX <= 16 ? 4 : 8;
                        unsigned w[BPW];
#pragma unroll
                        for (int q = 0; q < BPW; q++) {
                            unsigned v = 0;
#pragma unroll
                            for (int b = 0; b < 4; b++) if (4 * q + b < SMAX) v |= (unsigned)bpv[4 * q + b] << (8 * b);
                            w[q] = v;
                        }
                        uint4* dst = (uint4*)(A.bp + (ostart + t) * (long long)(4 * BPW));
                        dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
                        if (BPW == 8) dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
                    }
#pragma unroll
                    for (int i = 0; i < SMAX; i++) colT[i] = st.T1[i];
                    m1 = st.m1; m2 = st.m2; i1 = st.i1;
                }
            }
            __syncwarp();
        }
        if (CHAINED && r < A.plan.nseq) A.last_state[r] = i1;
    }
}

// resident CTAs per SM of one instantiation (0 if it does not fit in shared memory at all)
template <int SMAX, int MODEL, int LIKBB, int CHAINED, int STAGE>
static int decode_occupancy(int S)
{
    // per instantiation and per device (the shared-memory attribute below is a per-device setting); the answer
    // depends on S only.  Host threads of different devices touch different entries; two threads of the same
    // device would store the same values.
    constexpr int MAXDEV = 64;
    static int cached_S[MAXDEV], cached_occ[MAXDEV];                // zero-initialised: key 0 = not computed yet
    int dev = 0;
    cudaGetDevice(&dev);
    const int slot = dev & (MAXDEV - 1);
    if (cached_S[slot] == S + 1) return cached_occ[slot];
    size_t smem = decode_smem_bytes(SMAX, S, MODEL, LIKBB, STAGE);
    auto kern = k_decode<SMAX, MODEL, LIKBB, CHAINED, STAGE>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, SGB_DECODE_THREADS, smem) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    cached_occ[slot] = occ;
    cached_S[slot] = S + 1;
    return occ;
}

template <int SMAX, int MODEL, int LIKBB, int CHAINED, int STAGE>
static cudaError_t launch_decode(const SgbDecodeArgs& A, int sm_count, cudaStream_t st)
{
    int occ = decode_occupancy<SMAX, MODEL, LIKBB, CHAINED, STAGE>(A.S);
    if (STAGE == SGB_STAGE_TMA) {
        // the staging buffers cost shared memory: if they would lower the number of resident CTAs, stream the
        // records with plain coalesced loads instead (same arithmetic, same results)
        int occ_ldg = decode_occupancy<SMAX, MODEL, LIKBB, CHAINED, SGB_STAGE_LDG>(A.S);
        if (occ_ldg > occ) return launch_decode<SMAX, MODEL, LIKBB, CHAINED, SGB_STAGE_LDG>(A, sm_count, st);
    }
    if (occ < 1) return cudaErrorInvalidConfiguration;
    size_t smem = decode_smem_bytes(SMAX, A.S, MODEL, LIKBB, STAGE);
    auto kern = k_decode<SMAX, MODEL, LIKBB, CHAINED, STAGE>;
    int blocks = (A.plan.ngroups + SGB_DECODE_WARPS - 1) / SGB_DECODE_WARPS;
    int cap = sm_count * occ;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kern<<<blocks, SGB_DECODE_THREADS, smem, st>>>(A);
    return cudaGetLastError();
}

template <int SMAX>
static cudaError_t dispatch_decode(const SgbDecodeArgs& A, const SgbModelArgs& m, int stage, int sm_count, cudaStream_t st)
{
    const int ch = m.decode == SGB_DECODE_CHAINED;
    if (ch) stage = SGB_STAGE_LDG;      // the opt-in chained mode keeps the plain global-load path
    // beta-binomial: the small-count lgamma cache is read through L1; the TMA ring's shared memory would shrink
    // that cache (measured 5.25 ms vs 4.81 ms on the ASCN beta-binomial config), so it streams records with LDG
    if (m.lik_bb) stage = SGB_STAGE_LDG;
#define SGB_CASE(MD, LK, CH, SG) \
    if (m.model == MD && m.lik_bb == LK && ch == CH && stage == SG) return launch_decode<SMAX, MD, LK, CH, SG>(A, sm_count, st);
    SGB_CASE(0, 0, 0, 2) SGB_CASE(1, 0, 0, 2)
    SGB_CASE(0, 0, 0, 1) SGB_CASE(0, 1, 0, 1) SGB_CASE(1, 0, 0, 1) SGB_CASE(1, 1, 0, 1)
    SGB_CASE(0, 0, 0, 0) SGB_CASE(0, 1, 0, 0) SGB_CASE(1, 0, 0, 0) SGB_CASE(1, 1, 0, 0)
    SGB_CASE(0, 0, 1, 0) SGB_CASE(0, 1, 1, 0) SGB_CASE(1, 0, 1, 0) SGB_CASE(1, 1, 1, 0)
#undef SGB_CASE
    return cudaErrorInvalidValue;
}

// RAW staging exists for the binomial, reference-decode kernels and must not cost a resident CTA
template <int SMAX>
static int resolve_stage_t(const SgbModelArgs& m, int stage)
{
    if (m.decode == SGB_DECODE_CHAINED || m.lik_bb) return SGB_STAGE_LDG;
    if (stage != SGB_STAGE_RAW) return stage;
    const int occ_raw = m.model == SGB_MODEL_ASCN ? decode_occupancy<SMAX, 1, 0, 0, SGB_STAGE_RAW>(m.S)
                                                  : decode_occupancy<SMAX, 0, 0, 0, SGB_STAGE_RAW>(m.S);
    const int occ_ldg = m.model == SGB_MODEL_ASCN ? decode_occupancy<SMAX, 1, 0, 0, SGB_STAGE_LDG>(m.S)
                                                  : decode_occupancy<SMAX, 0, 0, 0, SGB_STAGE_LDG>(m.S);
    return (occ_raw >= 1 && occ_raw >= occ_ldg) ? SGB_STAGE_RAW : SGB_STAGE_TMA;
}

int sgbk_resolve_stage(const SgbModelArgs& m, int stage, const SgbRawIn& raw)
{
    if (stage == SGB_STAGE_RAW) {
        // 16-byte cp.async copies of 4-bin blocks: count columns 16-byte aligned, a uint8 cn column word-aligned
        const unsigned long long al = (unsigned long long)raw.n | (unsigned long long)raw.x;
        if ((al & 15ull) || ((unsigned long long)raw.cn & 3ull) || raw.cn_itemsize != 1) stage = SGB_STAGE_TMA;
    }
    if (m.S <= 12) return resolve_stage_t<12>(m, stage);
    if (m.S <= 16) return resolve_stage_t<16>(m, stage);
    return resolve_stage_t<SGB_MAX_STATES>(m, stage);
}

cudaError_t sgbk_decode(const SgbPlanDev& plan, const unsigned long long* d_packed, const SgbRawIn& raw,
                        const SgbDevTables* d_tabs, const SgbModelArgs& m, uint8_t* d_state_out, uint8_t* d_bp,
                        int* d_last_state, unsigned int* d_counter, int sm_count, int stage, cudaStream_t st)
{
    if (plan.nseq == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(d_counter, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return e;
    SgbDecodeArgs A;
    A.plan = plan; A.packed = d_packed; A.tabs = d_tabs; A.state_out = d_state_out; A.bp = d_bp;
    A.last_state = d_last_state; A.counter = d_counter; A.S = m.S;
    A.lik = m.lik_bb ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL;
    A.loherror = m.loherror; A.rho = m.rho;
    A.n = raw.n; A.x = raw.x; A.cn = raw.cn; A.rows = raw.rows; A.err = raw.err;
    if (m.S <= 12) return dispatch_decode<12>(A, m, stage, sm_count, st);
    if (m.S <= 16) return dispatch_decode<16>(A, m, stage, sm_count, st);
    if (m.S <= SGB_MAX_STATES) return dispatch_decode<SGB_MAX_STATES>(A, m, stage, sm_count, st);
    return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------ k_backtrace (decode = chained)
__global__ void k_backtrace(SgbPlanDev plan, const uint8_t* __restrict__ bp, const int* __restrict__ last_state,
                            int S, uint8_t* __restrict__ state_out)
{
    long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= plan.nseq) return;
    const int len = plan.len[r];
    if (len <= 0) return;
    const long long o = plan.ostart[r];
    int cur = last_state[r];
    state_out[o + len - 1] = (uint8_t)cur;
    for (int t = len - 1; t > 0; t--) {
        cur = bp[(o + t) * (long long)sgb_bp_stride(S) + cur];
        state_out[o + t - 1] = (uint8_t)cur;
    }
}

cudaError_t sgbk_backtrace(const SgbPlanDev& plan, const uint8_t* d_bp, const int* d_last_state, int S,
                           uint8_t* d_state_out, cudaStream_t st)
{
    if (plan.nseq == 0) return cudaSuccess;
    k_backtrace<<<(unsigned)((plan.nseq + 127) / 128), 128, 0, st>>>(plan, d_bp, d_last_state, S, d_state_out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ k_loglik  (K4: the `l` matrix)
// Output layout: for the sequence starting at row o with T bins, a T x S column-major block at l + o*S
// (what R holds as HaplotypeHMM()$l for that chromosome).
__global__ void __launch_bounds__(256) k_loglik(const int32_t* __restrict__ n, const int32_t* __restrict__ x,
                                                const void* __restrict__ cn, int cn_itemsize, SgbPlanDev plan,
                                                const SgbDevTables* __restrict__ tabs, SgbModelArgs m,
                                                double* __restrict__ l)
{
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const SgbScalars sp = tabs->sp;
    const int S = m.S;
    for (long long r = warp; r < plan.nseq; r += nwarps) {
        const long long o = plan.ostart[r];
        const int len = plan.len[r];
        for (int t = lane; t < len; t += 32) {
            int cv = (cn_itemsize == 1) ? (int)((const uint8_t*)cn)[o + t] : ((const int32_t*)cn)[o + t];
            SgbBin b = sgb_bin_prepare(m.model, m.lik_bb, n[o + t], x[o + t], cv, sp.bb_s, g_lftab, g_logtab);
            const int L = sgb_live_states(m.model, b.cprime, S);
            for (int i = 0; i < S; i++) {
                double e = -sgb_inf();
                if (i < L) {
                    SgbStateConst sc = sgb_state_const(b.cprime, i, m.loherror,
                                                       m.lik_bb ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL, m.rho,
                                                       sp.bb_s, g_logtab);
                    e = sgb_emission(m.model, m.lik_bb, b, i, sc, sp, g_logtab, g_exptab);
                }
                l[o * S + t + (long long)len * i] = e;
            }
        }
    }
}

cudaError_t sgbk_loglik(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize,
                        const SgbPlanDev& plan, const SgbDevTables* d_tabs, const SgbModelArgs& m, double* d_l,
                        cudaStream_t st)
{
    if (plan.nseq == 0) return cudaSuccess;
    long long blocks = (plan.nseq + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_loglik<<<(unsigned)blocks, 256, 0, st>>>(d_n, d_x, d_cn, cn_itemsize, plan, d_tabs, m, d_l);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ k_viterbi  (K3: src/code.cpp:11-45)
// One warp for the sequence; lane i owns target states i, i+32, ...; the S x S log transition matrix sits
// in shared memory transposed to [from][to] rows so a warp reads it conflict-free; the previous T1 column
// is double-buffered in shared memory; the column argmax is a shuffle reduction ordered by
// (value descending, index ascending) = "first maximum".  uint8 backpointers go to global scratch only in
// chained mode.  NaN emissions are treated as -Inf (the reference's behaviour there is undefined).
#define SGB_VIT_MAX_STATES 128
__global__ void __launch_bounds__(32) k_viterbi(const double* __restrict__ em, long long nrow_e,
                                                const double* __restrict__ tr, const int32_t* __restrict__ obs,
                                                long long T, int S, int init_mode, int decode,
                                                uint8_t* __restrict__ bp, int32_t* __restrict__ path)
{
    extern __shared__ __align__(16) double vsm[];
    double* trS = vsm;                 // [from k][to i]
    double* T1a = trS + (size_t)S * S;
    double* T1b = T1a + S;
    const int lane = threadIdx.x;
    const double NEG_INF = -sgb_inf();
    for (int idx = lane; idx < S * S; idx += 32) {
        int k = idx % S, i = idx / S;          // tr is column-major [from=k, to=i]
        trS[k * S + i] = tr[idx];
    }
    const double init = (init_mode == 0) ? -sgb_log((double)S, g_logtab)
                                         : sgb_log(sgb_div(1.0, (double)S), g_logtab);
    {
        const long long row = (long long)obs[0] - 1;
        for (int i = lane; i < S; i += 32) {
            double e = em[row + nrow_e * i];
            if (e != e) e = NEG_INF;
            T1a[i] = sgb_add(init, e);
        }
    }
    __syncwarp();
    double* cur = T1a;
    double* nxt = T1b;
    int back = 0;
    for (long long t = 1; t < T; t++) {
        const long long row = (long long)obs[t] - 1;
        double lbest = NEG_INF;
        int lidx = 0x7fffffff, lt2 = 0;
        for (int i = lane; i < S; i += 32) {
            double e = em[row + nrow_e * i];
            if (e != e) e = NEG_INF;
            double best = sgb_add(sgb_add(cur[0], trS[i]), e);
            int arg = 0;
            for (int k = 1; k < S; k++) {
                double v = sgb_add(sgb_add(cur[k], trS[k * S + i]), e);
                if (v > best) { best = v; arg = k; }
            }
            nxt[i] = best;
            if (decode == SGB_DECODE_CHAINED) bp[t * S + i] = (uint8_t)arg;
            if (lidx == 0x7fffffff || best > lbest) { lbest = best; lidx = i; lt2 = arg; }
        }
        // warp argmax: larger value wins, equal values -> smaller state index ("first maximum")
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            double ov = __shfl_xor_sync(SGB_FULL, lbest, off);
            int oi = __shfl_xor_sync(SGB_FULL, lidx, off);
            int ot = __shfl_xor_sync(SGB_FULL, lt2, off);
            bool take = (oi != 0x7fffffff) && (lidx == 0x7fffffff || ov > lbest || (ov == lbest && oi < lidx));
            if (take) { lbest = ov; lidx = oi; lt2 = ot; }
        }
        back = lt2;
        if (decode == SGB_DECODE_REFERENCE && lane == 0) path[t - 1] = back;
        __syncwarp();
        double* tmp = cur; cur = nxt; nxt = tmp;
    }
    if (decode == SGB_DECODE_REFERENCE) {
        if (lane == 0) path[T - 1] = (T > 1) ? back : 0;
    } else {
        // final column argmax, then follow the backpointers
        double lbest = NEG_INF;
        int lidx = 0x7fffffff;
        for (int i = lane; i < S; i += 32) {
            double v = cur[i];
            if (lidx == 0x7fffffff || v > lbest) { lbest = v; lidx = i; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            double ov = __shfl_xor_sync(SGB_FULL, lbest, off);
            int oi = __shfl_xor_sync(SGB_FULL, lidx, off);
            bool take = (oi != 0x7fffffff) && (lidx == 0x7fffffff || ov > lbest || (ov == lbest && oi < lidx));
            if (take) { lbest = ov; lidx = oi; }
        }
        __syncwarp();
        if (lane == 0) {
            int c = lidx;
            path[T - 1] = c;
            for (long long t = T - 1; t > 0; t--) {
                c = bp[t * S + c];
                path[t - 1] = c;
            }
        }
    }
}

size_t sgbk_viterbi_bp_bytes(long long T, int S) { return (size_t)T * (size_t)S; }

cudaError_t sgbk_viterbi(const double* d_emission, long long nrow_e, const double* d_trans, const int32_t* d_obs,
                         long long T, int S, int init_mode, int decode, uint8_t* d_bp, int32_t* d_path,
                         cudaStream_t st)
{
    if (S < 1 || S > SGB_VIT_MAX_STATES) return cudaErrorInvalidValue;
    if (T <= 0) return cudaSuccess;
    size_t smem = ((size_t)S * S + 2 * (size_t)S) * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(k_viterbi, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_viterbi<<<1, 32, smem, st>>>(d_emission, nrow_e, d_trans, d_obs, T, S, init_mode, decode, d_bp, d_path);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ small element-wise kernels
__global__ void k_logspace_add(const double* __restrict__ a, const double* __restrict__ b, long long n,
                               double* __restrict__ out)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = sgb_logspace_add(a[i], b[i], g_logtab, g_exptab);
}

cudaError_t sgbk_logspace_add(const double* d_a, const double* d_b, long long n, double* d_out, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_logspace_add<<<(unsigned)blocks, 256, 0, st>>>(d_a, d_b, n, d_out);
    return cudaGetLastError();
}

// test hook: evaluate one deterministic primitive on the device (fn codes as in oracle orc_math_eval)
__global__ void k_math_eval(int fn, const double* __restrict__ in, const double* __restrict__ in2, long long n,
                            double* __restrict__ out)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double x = in[i], r;
        switch (fn) {
        case 0: r = sgb_exp(x, g_exptab); break;
        case 1: r = sgb_log(x, g_logtab); break;
        case 2: r = sgb_log1p_01(x, g_logtab); break;
        case 3: r = sgb_log1p(x, g_logtab); break;
        case 4: r = sgb_lgamma(x, g_logtab); break;
        case 5: r = sgb_logspace_add(x, in2[i], g_logtab, g_exptab); break;
        case 6: r = sgb_logfact((int64_t)x, g_lftab, g_logtab); break;
        default: r = sgb_nan();
        }
        out[i] = r;
    }
}

cudaError_t sgbk_math_eval(int fn, const double* d_in, const double* d_in2, long long n, double* d_out,
                           cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_math_eval<<<(unsigned)blocks, 256, 0, st>>>(fn, d_in, d_in2, n, d_out);
    return cudaGetLastError();
}

// Post-decode table columns (R/callHSCN.R:282-290 clamp block; numeric part of add_states R/util.R:1078-1097;
// the ASCN block R/callASCN.R:416-447 reduces to the same arithmetic).  Codes:
//   phase 0 Balanced / 1 A / 2 B ; state_phase 0 Balanced 1 A-Gained 2 B-Gained 3 A-Hom 4 B-Hom ; LOH 1/0
__global__ void k_postprocess(const void* __restrict__ cn, int cn_itemsize, const uint8_t* __restrict__ path,
                              long long N, int32_t* __restrict__ A, int32_t* __restrict__ B,
                              int32_t* __restrict__ state_min, uint8_t* __restrict__ loh,
                              uint8_t* __restrict__ phase, uint8_t* __restrict__ state_phase,
                              double* __restrict__ state_baf)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        int st = (cn_itemsize == 1) ? (int)((const uint8_t*)cn)[i] : ((const int32_t*)cn)[i];
        int sm = path[i];
        int a = st - sm, b = sm;
        if (b < 0) b = 0;
        if (a < 0) a = 0;
        if (b > st) b = st;
        if (a > st) a = st;
        A[i] = a; B[i] = b;
        int mn = a < b ? a : b;
        state_min[i] = mn;
        loh[i] = (uint8_t)(mn == 0);
        phase[i] = (uint8_t)((b < a) ? 1 : (b > a) ? 2 : 0);
        state_phase[i] = (uint8_t)(1 * ((b < a) && (b != 0)) + 2 * ((b > a) && (a != 0)) +
                                   3 * ((b < a) && (b == 0)) + 4 * ((b > a) && (a == 0)));
        double q = sgb_div(sgb_div((double)b, (double)st), 0.1);
        double baf = sgb_mul(rint(q), 0.1);
        if (baf != baf) baf = 0.5;
        state_baf[i] = baf;
    }
}

cudaError_t sgbk_postprocess(const void* d_cn, int cn_itemsize, const uint8_t* d_path, long long N, int32_t* d_A,
                             int32_t* d_B, int32_t* d_state_min, uint8_t* d_loh, uint8_t* d_phase,
                             uint8_t* d_state_phase, double* d_state_baf, cudaStream_t st)
{
    if (N <= 0) return cudaSuccess;
    long long blocks = (N + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_postprocess<<<(unsigned)blocks, 256, 0, st>>>(d_cn, cn_itemsize, d_path, N, d_A, d_B, d_state_min, d_loh,
                                                    d_phase, d_state_phase, d_state_baf);
    return cudaGetLastError();
}

// Between-pass statistics: everything the drivers reduce over the whole decoded table before
// the second HMM pass, as exact integer sums.
//   out[0], out[1], out[2] : sum(alleleB), sum(totalcounts), row count over the "A-Hom" rows (B == 0 < A after
//                            the clamp block) -> infloherror = weighted.mean(BAF, totalcounts) = out[0] / out[1]
//                            (R/callHSCN.R:898-902, R/callASCN.R:334-343)
//   out[3 + A*S + B]       : number of rows with that (A, B), B > 0 -> the modal state fitBB starts from
//                            (R/callHSCN.R:620-621)
__global__ void __launch_bounds__(256) k_pass_stats(const int32_t* __restrict__ n, const int32_t* __restrict__ x,
                                                    const void* __restrict__ cn, int cn_itemsize,
                                                    const uint8_t* __restrict__ path, long long N, int S,
                                                    unsigned long long* __restrict__ out)
{
    __shared__ unsigned int hist[SGB_MAX_STATES * SGB_MAX_STATES];
    for (int i = threadIdx.x; i < S * S; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    unsigned long long sx = 0, sn = 0, cnt = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        int st = (cn_itemsize == 1) ? (int)((const uint8_t*)cn)[i] : ((const int32_t*)cn)[i];
        int sm = path[i];
        int a = st - sm, b = sm;
        if (a < 0) a = 0;
        if (b > st) b = st;
        if (a > st) a = st;
        if (b == 0 && a > 0) { sx += (unsigned long long)x[i]; sn += (unsigned long long)n[i]; cnt++; }
        if (b > 0 && a < S && b < S) atomicAdd(&hist[a * S + b], 1u);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        sx += __shfl_xor_sync(SGB_FULL, sx, off);
        sn += __shfl_xor_sync(SGB_FULL, sn, off);
        cnt += __shfl_xor_sync(SGB_FULL, cnt, off);
    }
    if ((threadIdx.x & 31) == 0) {
        if (sx) atomicAdd(&out[0], sx);
        if (sn) atomicAdd(&out[1], sn);
        if (cnt) atomicAdd(&out[2], cnt);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < S * S; i += blockDim.x)
        if (hist[i]) atomicAdd(&out[3 + i], (unsigned long long)hist[i]);
}

cudaError_t sgbk_pass_stats(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize,
                            const uint8_t* d_path, long long N, int S, unsigned long long* d_out, cudaStream_t st)
{
    if (N <= 0) return cudaSuccess;
    long long blocks = (N + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_pass_stats<<<(unsigned)blocks, 256, 0, st>>>(d_n, d_x, d_cn, cn_itemsize, d_path, N, S, d_out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ FP64 roofline denominator
// Dependent-chain-free DFMA loop: 8 accumulators per thread, 16 warps per SM resident.
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int n_iter)
{
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1e-3, a2 = a0 + 2e-3, a3 = a0 + 3e-3;
    double a4 = a0 + 4e-3, a5 = a0 + 5e-3, a6 = a0 + 6e-3, a7 = a0 + 7e-3;
    const double b = 0.9999999, c = 1e-9;
    for (int i = 0; i < n_iter; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
            a0 = __fma_rn(a0, b, c); a1 = __fma_rn(a1, b, c); a2 = __fma_rn(a2, b, c); a3 = __fma_rn(a3, b, c);
            a4 = __fma_rn(a4, b, c); a5 = __fma_rn(a5, b, c); a6 = __fma_rn(a6, b, c); a7 = __fma_rn(a7, b, c);
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

cudaError_t sgbk_fp64_peak(int iters, int sm_count, cudaStream_t st, double* tflops)
{
    const int blocks = sm_count * 8, threads = 256, n_iter = 4096;
    double* d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_out, (size_t)blocks * threads * sizeof(double));
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_fp64_peak<<<blocks, threads, 0, st>>>(d_out, n_iter);
    k_fp64_peak<<<blocks, threads, 0, st>>>(d_out, n_iter);
    cudaEventRecord(e0, st);
    if (iters < 1) iters = 1;
    for (int i = 0; i < iters; i++) k_fp64_peak<<<blocks, threads, 0, st>>>(d_out, n_iter);
    cudaEventRecord(e1, st);
    e = cudaEventSynchronize(e1);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d_out);
    if (e != cudaSuccess) return e;
    const double flops = (double)blocks * threads * (double)n_iter * 32.0 * 2.0 * (double)iters;
    *tflops = flops / ((double)ms * 1e-3) / 1e12;
    return cudaGetLastError();
}
