// split_kernels.cu -- the SPLIT schedule of the HSCN / ASCN decode (the default for decode = reference).
//
// The fused k_decode (hmm_kernels.cu) gives every (cell, chromosome) sequence ONE lane, so a batch of n sequences offers
// n/32 warps and every warp walks a dependent chain of emission + recursion steps: batches below ~75,000 sequences
// cannot fill a B200 (148 SMs x 16+ resident warps).  The emission log-likelihoods (R/callHSCN.R:48-80,
// R/callASCN.R:50-81) are independent per bin; only the max-plus recursion (src/code.cpp:27-33) is sequential.  So:
//
//   k_split_class    one warp per sequence: largest total copy number -> number of 4-state "planes" the sequence needs
//   k_split_scan / k_split_scatter   stable partition of the sequences by plane count (descending), so that the 32
//                    sequences of a recursion warp need the same number of planes (and warps stay homogeneous)
//   k_split_groups / k_split_gscan   per group of 32 sequences: plane count, longest sequence, tile offset
//   k_emit           one warp per sequence, LANES OVER BINS: coalesced loads of the caller's columns (any of the
//                    accepted element types), all live states of a bin evaluated by the lane that owns the bin
//                    (copy number is piecewise constant along a chromosome, so a warp is branch-uniform), emissions
//                    written as 32-byte sectors into the group's tile
//   k_recur          one warp per group, one lane per sequence: the tile is streamed through a shared-memory ring by
//                    1-D TMA bulk copies (cp.async.bulk + mbarrier, 4 bins x 32 sequences per copy, several copies
//                    ahead), the forward column lives in registers, per bin the O(S) step of hmm_core.cuh and the
//                    reference's decode rule
//
// Tile of a group with P planes: chunks of 4 bins, 4 P KB each, contiguous (one TMA copy); inside a chunk the 4 P sectors
// of one sequence (4 bins x P planes) are contiguous -- a k_emit warp therefore writes whole 128-byte lines -- and are
// rotated by the lane number so that the 32 lanes of k_recur read them from different shared-memory banks:
//     sector(chunk c, lane l, bin b, plane p) = c * 128 P + l * 4 P + (b * P + p + l) mod 4 P
//
// Both kernels read the exp / log tables from lane-private copies in shared memory (entry j of lane l at column l mod 16,
// resp. pair i at column l mod 8), so a table look-up with 32 unrelated indices is free of bank conflicts; the values,
// and so the bits of every result, are those of the packed tables.  Same arithmetic functions as the fused kernel
// (sgb_emission_fast*, sgb_emit_item_slow, the max-plus step), hence the same bits.
#define SGB_ETI(j) ((j) * 16)             // exp table: 16 lane-private columns of doubles
#define SGB_LTI(i) ((i) * 16)             // log table: 8 lane-private columns of {1/c, log c} pairs
#include <cstdlib>
#include "kernels.h"
#include "sgb_async.cuh"

static __device__ const double gs_logtab[SGB_LOGTAB_LEN] = SGB_LOGTAB_INIT;
static __device__ const double gs_exptab[SGB_EXPTAB_LEN] = SGB_EXPTAB_INIT;
static __device__ const double gs_lftab[SGB_LFTAB_LEN] = SGB_LFTAB_INIT;

#define SGB_FULL 0xffffffffu
// Programmatic dependent launch: the kernels of one decode are launched back to back with the stream-serialization attribute;
// each lets its successor start launching right away (its blocks become resident as this kernel's blocks retire and run
// their prologue -- table loads, barrier set-up) and waits here for its predecessor to have completed before touching its
// results.  Launched without the attribute both calls are no-ops.
#define SGB_PDL_ENTER() do { cudaTriggerProgrammaticLaunchCompletion(); cudaGridDependencySynchronize(); } while (0)


#define SGB_NCLS 9                        // plane classes 0 (empty sequence) .. 8
#define SGB_PART_TILE 1024                // sequences per partition tile

__device__ __forceinline__ int sgb_load_cn(const void* cn, int kind, long long row)
{
    if (kind == SGB_IN_I32_I32) return __ldg((const int32_t*)cn + row);
    if (kind == SGB_IN_F64) {
        const double v = __ldg((const double*)cn + row);
        return (v >= 0.0 && v <= 255.0) ? (int)v : -1;
    }
    return (int)__ldg((const uint8_t*)cn + row);
}

// ------------------------------------------------------------------ k_split_class
__global__ void __launch_bounds__(256) k_split_class(SgbPlanDev plan, long long nseq, const void* __restrict__ cn, int kind,
                                                     int model, int S, SgbSplitDev W)
{
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < nseq; i += nwarps) {
        const int len = plan.len[i];
        const long long o = plan.ostart[i];
        int mx = 0;
#pragma unroll 4
        for (int t = lane; t < len; t += 32) {          // (unrolled: the loads of four strides are in flight together)
            const int c = sgb_load_cn(cn, kind, o + t);
            mx = c > mx ? c : mx;
        }
        mx = __reduce_max_sync(SGB_FULL, mx);
        if (lane == 0) {
            mx = mx > 255 ? 255 : mx;              // out-of-range values are reported by k_emit; clamp for the class only
            const int L = sgb_live_states(model, sgb_cprime(model, mx), S);
            const int c = len > 0 ? sgb_planes_for(L) : 0;
            W.cls[i] = (uint8_t)c;
            atomicAdd(&W.part[(i / SGB_PART_TILE) * SGB_NCLS + c], 1u);
        }
    }
}

// ------------------------------------------------------------------ k_split_scan (one block)
// Classes are laid out in DESCENDING plane count (heavy groups start first), every class on a group boundary.
__global__ void __launch_bounds__(256) k_split_scan(SgbSplitDev W)
{
    __shared__ unsigned int tot[SGB_NCLS], start[SGB_NCLS];
    SGB_PDL_ENTER();
    const int c = threadIdx.x;
    if (c < SGB_NCLS) {
        unsigned int s = 0;
        for (int b = 0; b < W.ntile; b++) s += W.part[b * SGB_NCLS + c];
        tot[c] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int pos = 0;
        for (int k = SGB_NCLS - 1; k >= 1; k--) {
            start[k] = pos;
            pos = (pos + tot[k] + 31u) & ~31u;
        }
        start[0] = pos;            // empty sequences are not placed at all
        *W.ngroups = (int)(pos >> 5);
        W.counters[0] = 0; W.counters[1] = 0;
    }
    __syncthreads();
    if (c < SGB_NCLS) {
        unsigned int run = start[c];
        for (int b = 0; b < W.ntile; b++) {
            const unsigned int h = W.part[b * SGB_NCLS + c];
            W.part[b * SGB_NCLS + c] = run;
            run += h;
        }
    }
}

// ------------------------------------------------------------------ k_split_scatter (one block per tile, stable)
__global__ void __launch_bounds__(SGB_PART_TILE) k_split_scatter(long long nseq, SgbSplitDev W)
{
    __shared__ unsigned int wcnt[32][SGB_NCLS];
    SGB_PDL_ENTER();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long i = (long long)blockIdx.x * SGB_PART_TILE + threadIdx.x;
    const int c = i < nseq ? (int)W.cls[i] : 0;
    unsigned int rank = 0;
#pragma unroll
    for (int k = 1; k < SGB_NCLS; k++) {
        const unsigned int m = __ballot_sync(SGB_FULL, c == k);
        if (c == k) rank = __popc(m & ((1u << lane) - 1u));
        if (lane == 0) wcnt[wid][k] = __popc(m);
    }
    __syncthreads();
    if (c > 0) {
        unsigned int before = 0;
        for (int w = 0; w < wid; w++) before += wcnt[w][c];
        W.perm[W.part[blockIdx.x * SGB_NCLS + c] + before + rank] = (int)i;
    }
}

// ------------------------------------------------------------------ k_split_groups (one warp per group)
__global__ void __launch_bounds__(256) k_split_groups(SgbPlanDev plan, SgbSplitDev W)
{
    SGB_PDL_ENTER();
    const int lane = threadIdx.x & 31;
    const int g = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (g >= *W.ngroups) return;
    const int idx = W.perm[(long long)g * 32 + lane];
    int npl = 0, len = 0;
    if (idx >= 0) { npl = W.cls[idx]; len = plan.len[idx]; }
    npl = __reduce_max_sync(SGB_FULL, npl);
    len = __reduce_max_sync(SGB_FULL, len);
    if (lane == 0) {
        W.g_npl[g] = npl;
        W.g_tmax[g] = len;
        W.g_size[g] = (unsigned long long)((len + 7) & ~7) * (unsigned long long)npl * 32ull;     // whole ring stages stay inside the tile
    }
}

// exclusive scan of the tile sizes (one block); a tile that would not fit the emission buffer raises an error bit
__global__ void __launch_bounds__(1024) k_split_gscan(SgbSplitDev W, int* err)
{
    __shared__ unsigned long long wsum[32];
    __shared__ unsigned long long carry;
    SGB_PDL_ENTER();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ng = *W.ngroups;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int g0 = 0; g0 < ng; g0 += 1024) {
        const int g = g0 + threadIdx.x;
        const unsigned long long v = g < ng ? W.g_size[g] : 0ull;
        unsigned long long incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long u = __shfl_up_sync(SGB_FULL, incl, d);
            if (lane >= d) incl += u;
        }
        if (lane == 31) wsum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            unsigned long long s = wsum[lane], si = s;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long u = __shfl_up_sync(SGB_FULL, si, d);
                if (lane >= d) si += u;
            }
            wsum[lane] = si - s;         // exclusive over warps
        }
        __syncthreads();
        const unsigned long long excl = carry + wsum[wid] + incl - v;
        if (g < ng) {
            W.g_eoff[g] = (long long)excl;
            if (excl + v > (unsigned long long)W.e_cap_sectors) atomicOr(err, SGB_ERR_INTERNAL);
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}

// ------------------------------------------------------------------ k_emit
struct SgbEmitArgs {
    SgbPlanDev plan;
    SgbSplitDev W;
    SgbColsIn in;
    const SgbDevTables* tabs;
    int S, lik;
    double loherror, rho;
};

#define SGB_EMIT_THREADS 256
#ifndef SGB_EMIT_BIN_MINBLOCKS
#define SGB_EMIT_BIN_MINBLOCKS 3
#endif
// resident CTAs per SM the register allocation is bounded for (measured, 57 M cell-bins: binomial 3 -> 1.154 ms, 4 -> 1.163, 2 -> 1.379;
// beta-binomial 3 -> 2.91 ms, 2 -> 3.12, 4 -> 2.97; evaluating two beta-binomial items at a time: 2.96 at 2 CTAs, 3.25 at 3)
#ifndef SGB_EMIT_BB_MINBLOCKS
#define SGB_EMIT_BB_MINBLOCKS 3
#endif
#define SGB_EMIT_LF 1024
#define SGB_LTP_LEN (((SGB_LOGTAB_LEN + 1) / 2) * 16)      // lane-private log table: pairs x 8 columns x 2 doubles
#define SGB_ETP_LEN (SGB_EXPTAB_LEN * 16)                   // lane-private exp table: entries x 16 columns
#ifndef SGB_EMIT_BB_PAIR
#define SGB_EMIT_BB_PAIR 0
#endif
#ifndef SGB_PAIR_OK
#define SGB_PAIR_OK(likbb) (!(likbb) || SGB_EMIT_BB_PAIR)      // beta-binomial: one item at a time (instruction-cache footprint, see hmm_kernels.cu)
#endif
static inline size_t emit_smem_bytes(int S, int lik_bb)
{
    size_t d = SGB_LTP_LEN + SGB_ETP_LEN + SGB_EMIT_LF + (size_t)SGB_CN_TABLE * S * (lik_bb ? 5 : 2);
    return d * sizeof(double) + ((size_t)SGB_CN_TABLE * S + SGB_CN_TABLE) * sizeof(int);
}

// One row of the caller's columns, whatever their element type, in two halves: sgb_row_fetch only issues the loads (the
// row of the NEXT 32-bin tile is fetched while the current one is evaluated), sgb_row_check converts and validates the
// values when they are consumed -- a check right behind the loads would wait for them and undo the prefetch (it did: a
// quarter of k_emit's stall samples sat on the first instruction after the loads).  KIND >= 0: the element types are
// known at compile time (no branches); KIND < 0: taken from in.kind.
struct SgbRawRow {
    int n, x, cn;            // integer kinds
    double dn, dx, dc;       // SGB_IN_F64
};
template <int KIND>
__device__ __forceinline__ void sgb_row_fetch(const SgbColsIn& in, long long row, SgbRawRow& r)
{
    const int kind = KIND >= 0 ? KIND : in.kind;
    if (kind == SGB_IN_U16_U8) {
        r.n = (int)__ldg((const uint16_t*)in.n + row); r.x = (int)__ldg((const uint16_t*)in.x + row);
        r.cn = (int)__ldg((const uint8_t*)in.cn + row);
    } else if (kind == SGB_IN_F64) {
        r.dn = __ldg((const double*)in.n + row); r.dx = __ldg((const double*)in.x + row);
        r.dc = __ldg((const double*)in.cn + row);
    } else {
        r.n = __ldg((const int32_t*)in.n + row); r.x = __ldg((const int32_t*)in.x + row);
        r.cn = kind == SGB_IN_I32_I32 ? __ldg((const int32_t*)in.cn + row) : (int)__ldg((const uint8_t*)in.cn + row);
    }
}
// returns validation bits; a bad bin is neutralised (n = x = cn = 0), the call fails afterwards -- the checks of the fused path
template <int KIND>
__device__ __forceinline__ int sgb_row_check(const SgbColsIn& in, const SgbRawRow& r, int& n, int& x, int& cnv)
{
    int bad = 0;
    const int kind = KIND >= 0 ? KIND : in.kind;
    if (kind == SGB_IN_F64) {
        const double dn = r.dn, dx = r.dx, dc = r.dc;
        const double lim = (double)(1 << SGB_COUNT_BITS);
        if (dn != dn || dx != dx || dc != dc) bad |= SGB_ERR_NOT_INTEGER;
        else if (dn < 0.0 || dx < 0.0) bad |= SGB_ERR_NEGATIVE_COUNT;
        else if (dn >= lim || dx >= lim) bad |= SGB_ERR_COUNT_TOO_LARGE;
        else if (dc < 0.0 || dc > 255.0) bad |= SGB_ERR_BAD_CN;
        else if (dn != rint(dn) || dx != rint(dx) || dc != rint(dc)) bad |= SGB_ERR_NOT_INTEGER;
        n = bad ? 0 : (int)dn; x = bad ? 0 : (int)dx; cnv = bad ? 0 : (int)dc;
    } else {
        n = r.n; x = r.x; cnv = r.cn;
    }
    if ((unsigned)(n | x) >> SGB_COUNT_BITS) {
        bad |= (n | x) < 0 ? SGB_ERR_NEGATIVE_COUNT : SGB_ERR_COUNT_TOO_LARGE;
        n = 0; x = 0; cnv = 0;
    }
    if ((unsigned)cnv > 255u) { bad |= SGB_ERR_BAD_CN; n = 0; x = 0; cnv = 0; }
    return bad;
}

template <int MODEL, int LIKBB, int KIND>
__global__ void __launch_bounds__(SGB_EMIT_THREADS, LIKBB ? SGB_EMIT_BB_MINBLOCKS : SGB_EMIT_BIN_MINBLOCKS) k_emit(SgbEmitArgs A)
{
    extern __shared__ __align__(16) double smem[];
    const int S = A.S;
    const int CTS = SGB_CN_TABLE * S;
    double* sLTp = smem;                                   // lane-private log table
    double* sETp = sLTp + SGB_LTP_LEN;                     // lane-private exp table
    double* sLF = sETp + SGB_ETP_LEN;
    double* sLP = sLF + SGB_EMIT_LF;
    double* sA = sLP + 2 * CTS;
    double* sB = sA + (LIKBB ? CTS : 0);
    double* sL0 = sB + (LIKBB ? CTS : 0);
    int* sKind = (int*)(sL0 + (LIKBB ? CTS : 0));
    int* sKM = sKind + CTS;                 // per total-CN bit mask of the states on the lean path
    if (threadIdx.x < SGB_CN_TABLE) {
        // 1 when every live state of total copy number c is an ordinary one (0 < p < 1) and the call's eps / Abias allow the
        // lean emission: then a bin with this copy number and a count inside the support takes the lean path
        const int Lc_ = sgb_live_states(MODEL, (int)threadIdx.x, S);
        int ok = A.tabs->sp.fast_ok;
        for (int st = 0; st < Lc_; st++)
            if (A.tabs->kind[threadIdx.x * SGB_MAX_STATES + st] != SGB_KIND_NORMAL) ok = 0;
        sKM[threadIdx.x] = ok;
    }
    for (int i = threadIdx.x; i < SGB_LTP_LEN; i += blockDim.x) {
        const int pair = i >> 4, w = i & 1;                // (pair, column, word): column = (i >> 1) & 7
        const int src = 2 * pair + w;
        sLTp[i] = src < SGB_LOGTAB_LEN ? gs_logtab[src] : 0.0;
    }
    for (int i = threadIdx.x; i < SGB_ETP_LEN; i += blockDim.x) sETp[i] = gs_exptab[i >> 4];
    for (int i = threadIdx.x; i < SGB_EMIT_LF; i += blockDim.x) sLF[i] = gs_lftab[i];
    for (int i = threadIdx.x; i < CTS; i += blockDim.x) {
        const int c = i / S, s = i - c * S, src = c * SGB_MAX_STATES + s;
        sLP[2 * i] = A.tabs->lp[src];
        sLP[2 * i + 1] = A.tabs->lq[src];
        sKind[i] = A.tabs->kind[src];
        if (LIKBB) { sA[i] = A.tabs->a[src]; sB[i] = A.tabs->b[src]; sL0[i] = A.tabs->l0[src]; }
    }
    __shared__ SgbScalars s_sp;
    if (threadIdx.x == 0) s_sp = A.tabs->sp;
    __syncthreads();
    SGB_PDL_ENTER();            // (the tables above come from k_setup, which completed before the partition kernels started)
    const SgbScalars& sp = s_sp;
    const int lane = threadIdx.x & 31;
    const double* LT = sLTp + 2 * (lane & 7);              // this lane's columns of the two tables
    const double* ET = sETp + (lane & 15);
    SgbTables TB;                                          // per thread: the table pointers are lane-private
    TB.LT = LT; TB.ET = ET; TB.LP = sLP; TB.LQ = sLP + 1; TB.pstride = 2; TB.A = sA; TB.B = sB; TB.L0 = sL0;
    TB.LGK = A.tabs->lgk; TB.KIND = sKind; TB.S = S; TB.lik = A.lik; TB.loherror = A.loherror; TB.rho = A.rho;
    const double NEG_INF = -sgb_inf();
    const long long nslots = (long long)(*A.W.ngroups) * 32;
    int bad = 0;
    for (;;) {
        long long r = 0;
        if (lane == 0) r = (long long)atomicAdd(&A.W.counters[0], 1u);
        r = __shfl_sync(SGB_FULL, r, 0);
        if (r >= nslots) break;
        const int idx = A.W.perm[r];
        if (idx < 0) continue;
        const int len = A.plan.len[idx];
        const long long o = A.plan.ostart[idx];
        const int g = (int)(r >> 5), lr = (int)(r & 31);
        const int npl = A.W.g_npl[g];
        const int run = 4 * npl;                                          // sectors of one sequence inside a chunk
        const int lrm = lr % run;
        double* tile = A.W.E + (A.W.g_eoff[g] + (long long)lr * run) * 4;  // + chunk * 128 npl sectors + rotated position
        SgbRawRow raw;
        raw.n = raw.x = raw.cn = 0; raw.dn = raw.dx = raw.dc = 0.0;
        if (lane < len) sgb_row_fetch<KIND>(A.in, o + lane, raw);
        for (int t0 = 0; t0 < len; t0 += 32) {
            const int t = t0 + lane;
            const bool act = t < len;
            int nc = 0, xc = 0, cc = 0;
            if (act) bad |= sgb_row_check<KIND>(A.in, raw, nc, xc, cc);
            // next tile's row: in flight while this one is evaluated
            if (t + 32 < len) sgb_row_fetch<KIND>(A.in, o + t + 32, raw);
            const int cprime = sgb_cprime(MODEL, cc);
            const int ins = xc <= nc;
            const int L = act ? sgb_live_states(MODEL, cprime, S) : 0;
            double lcx = 0.0, lcy = 0.0, lgs = 0.0;
            if (act && ins) {
                double fn, fx, fy;
                if (nc < SGB_EMIT_LF) { fn = sLF[nc]; fx = sLF[xc]; fy = sLF[nc - xc]; }
                else if (nc < SGB_NLF) { fn = gs_lftab[nc]; fx = gs_lftab[xc]; fy = gs_lftab[nc - xc]; }
                else { fn = sgb_logfact(nc, gs_lftab, LT); fx = sgb_logfact(xc, gs_lftab, LT); fy = sgb_logfact(nc - xc, gs_lftab, LT); }
                lcx = sgb_sub(sgb_sub(fn, fx), fy);
                if (MODEL == SGB_MODEL_ASCN) lcy = sgb_sub(sgb_sub(fn, fy), fx);
                if (LIKBB) lgs = sgb_lgamma_pos(sgb_add(sp.bb_s, (double)nc), LT);
            }
            const double kd = (double)xc, nkd = (double)(nc - xc);
            double* chunk = tile + (long long)(t >> 2) * (128 * npl) * 4;
            const int rot = (t & 3) * npl + lrm;                           // + plane, mod run (lrm = lr mod run)
            const int Lw = __reduce_max_sync(SGB_FULL, L);
            // lean path for the whole warp: every live state of every lane's bin is an ordinary one (tabulated total CN,
            // 0 < p < 1, count inside the support, default eps / no Abias) -- the usual case.  Idle lanes ride along on
            // harmless operands.  Otherwise the general per-item function (same bits where both apply).
            const bool lean = !act || (ins && cprime < SGB_CN_TABLE && sKM[cprime < SGB_CN_TABLE ? cprime : 0]);
            const bool warp_lean = __all_sync(SGB_FULL, lean);
            const bool uniformL = __all_sync(SGB_FULL, L == Lw || !act);
            const int trow = (act && cprime < SGB_CN_TABLE ? cprime : 1) * S;      // idle lanes: any valid row
            const int Lc = L > 0 ? L - 1 : 0;                                       // states beyond L are computed on row Lc and dropped
            for (int p = 0; p < npl; p++) {
                double e0 = NEG_INF, e1 = NEG_INF, e2 = NEG_INF, e3 = NEG_INF;
                if (4 * p < Lw) {
                    if (warp_lean) {
#define SGB_TI(m) (trow + ((m) < Lc ? (m) : Lc))
#define SGB_LEAN2(ma, mb, ra, rb) do {                                                                                  \
                            const int ta_ = SGB_TI(ma), tb_ = SGB_TI(mb);                                                  \
                            const double2 pa_ = ((const double2*)sLP)[ta_], pb_ = ((const double2*)sLP)[tb_];            \
                            sgb_emission_fast2<MODEL, LIKBB>(kd, nkd, lcx, lcy, lgs, pa_.x, pa_.y, LIKBB ? sA[ta_] : 0.0,  \
                                LIKBB ? sB[ta_] : 0.0, LIKBB ? sL0[ta_] : 0.0, kd, nkd, lcx, lcy, lgs, pb_.x, pb_.y,        \
                                LIKBB ? sA[tb_] : 0.0, LIKBB ? sB[tb_] : 0.0, LIKBB ? sL0[tb_] : 0.0,                      \
                                LIKBB ? TB.LGK + (long)ta_ * 2 * SGB_LGK : LT, LIKBB ? TB.LGK + (long)tb_ * 2 * SGB_LGK : LT, \
                                sp, LT, ET, ra, rb); } while (0)
#define SGB_LEAN1(ma, ra) do {                                                                                          \
                            const int ta_ = SGB_TI(ma);                                                                    \
                            const double2 pa_ = ((const double2*)sLP)[ta_];                                               \
                            ra = sgb_emission_fast<MODEL, LIKBB>(kd, nkd, lcx, lcy, lgs, pa_.x, pa_.y, LIKBB ? sA[ta_] : 0.0, \
                                LIKBB ? sB[ta_] : 0.0, LIKBB ? sL0[ta_] : 0.0, LIKBB ? TB.LGK + (long)ta_ * 2 * SGB_LGK : LT, \
                                sp, LT, ET); } while (0)
                        if (SGB_PAIR_OK(LIKBB)) {
                            if (4 * p + 1 < Lw) SGB_LEAN2(4 * p, 4 * p + 1, e0, e1); else SGB_LEAN1(4 * p, e0);
                            if (4 * p + 2 < Lw) { if (4 * p + 3 < Lw) SGB_LEAN2(4 * p + 2, 4 * p + 3, e2, e3); else SGB_LEAN1(4 * p + 2, e2); }
                        } else {
                            SGB_LEAN1(4 * p, e0);
                            if (4 * p + 1 < Lw) SGB_LEAN1(4 * p + 1, e1);
                            if (4 * p + 2 < Lw) SGB_LEAN1(4 * p + 2, e2);
                            if (4 * p + 3 < Lw) SGB_LEAN1(4 * p + 3, e3);
                        }
#undef SGB_LEAN1
#undef SGB_LEAN2
#undef SGB_TI
                        if (!uniformL) {        // lanes with fewer live states than the warp's maximum drop the surplus
                            e0 = 4 * p + 0 < L ? e0 : NEG_INF; e1 = 4 * p + 1 < L ? e1 : NEG_INF;
                            e2 = 4 * p + 2 < L ? e2 : NEG_INF; e3 = 4 * p + 3 < L ? e3 : NEG_INF;
                        }
                    } else {
                        if (4 * p + 0 < L) e0 = sgb_emit_item_slow<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 0, TB, sp);
                        if (4 * p + 1 < L) e1 = sgb_emit_item_slow<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 1, TB, sp);
                        if (4 * p + 2 < L) e2 = sgb_emit_item_slow<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 2, TB, sp);
                        if (4 * p + 3 < L) e3 = sgb_emit_item_slow<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 3, TB, sp);
                    }
                }
                if (act) {
                    int pos = rot + p;                  // < 2 run
                    pos = pos >= run ? pos - run : pos;
                    double* dst = chunk + pos * 4;
                    __stcg((double2*)dst, make_double2(e0, e1));
                    __stcg((double2*)(dst + 2), make_double2(e2, e3));
                }
            }
        }
    }
    if (bad) atomicOr(A.in.err, bad);
}

// ------------------------------------------------------------------ k_recur
struct SgbRecurArgs {
    SgbPlanDev plan;
    SgbSplitDev W;
    const SgbDevTables* tabs;
    uint8_t* state_out;
    int S;
};

#define SGB_RECUR_WARPS 4               // per CTA when at most 3 planes (maxCN <= 11); 2 for the 8-plane kernel
#define SGB_RECUR_NWARPS(maxnpl) ((maxnpl) <= 3 ? SGB_RECUR_WARPS : 2)
#define SGB_RECUR_MAXDEPTH 16            // ring slots per warp: one mbarrier each, phases kept in a bit mask
// Decoded states are staged per warp as [bin mod 32][sequence] (row stride 33 bytes: both the per-step column write and the
// per-sequence row read of the flush are spread over the banks) and written out 32 bins of ONE sequence per store
// instruction -- one or two 32-byte sectors instead of 32 scattered ones (a lane-per-sequence byte store costs the L1 ~64
// cycles per warp, which with 8 warps per SM was 2/3 of the step).
#ifndef SGB_FAST_MAXNPL
#define SGB_FAST_MAXNPL 2
#endif
#define SGB_STAGE_STRIDE 33
#define SGB_STAGE_BYTES (32 * SGB_STAGE_STRIDE + 16)

// One group of 32 sequences with NPL planes.  Chunk c of its tile (4 bins, 4 NPL KB, contiguous) is one 1-D TMA bulk copy
// into the next slot of the warp's ring, `depth` chunks ahead of the recursion.
// DD = 1: diagonal-dominant transition (every realistic selftransitionprob), forward column in registers, O(S) step.
// DD = 0: selftransitionprob below 1/S -- the two-maxima variant, kept out of line so that it does not set the hot
// path's register count.
template <int NPL, int DD, bool STAGED>
__device__ __forceinline__ void recur_group_t(const SgbRecurArgs& A, const SgbScalars& sp, unsigned char* ring, unsigned long long* bar,
                                              unsigned& phases, int depth, int lane, int Tmax, long long eoff, int len,
                                              long long ostart, unsigned char* stage)
{
    constexpr int NL = 4 * NPL;
    constexpr int RUN = 4 * NPL;                       // sectors of one sequence inside a chunk
    constexpr unsigned SLOT = NPL * 4096;
    // the four-step block pays in the throughput regime (measured 0.49 -> 0.44 ms at 10,000 cells); on a latency-bound batch and for
    // wider groups the one-step-at-a-time body is the faster one (0.164 vs 0.179 ms at 1,000 cells)
    constexpr bool FAST = DD != 0 && STAGED && NPL <= SGB_FAST_MAXNPL;
    const unsigned char* src = (const unsigned char*)A.W.E + eoff * 32;
    const int nchunks = (Tmax + 3) >> 2;
    auto issue = [&](int c, int s) {
        if (lane == 0) {
            sgb_mbar_expect_tx(&bar[s], SLOT);
            sgb_tma_load_1d(ring + s * SLOT, src + (long long)c * SLOT, SLOT, &bar[s]);
        }
    };
    for (int c = 0; c < depth && c < nchunks; c++) issue(c, c);
    double T1[DD ? NL : 1];
    double m1 = 0.0;
    SgbSeqState<DD ? 1 : NL> st;
    const int Seff = A.S < NL ? A.S : NL;
    const double tr_self = sp.tr_self, tr_off = sp.tr_off, init = sp.init;
    uint8_t* outp = A.state_out + ostart;
    const int lrm = lane % RUN;
    const unsigned char* mine0 = ring + lane * (RUN * 32);         // this lane's run inside slot 0
    int lastback = 0;                                              // backpointer found at the sequence's last bin
    int s = 0;
    for (int c = 0; c < nchunks; c++) {
        sgb_mbar_wait(&bar[s], (phases >> s) & 1u);
        phases ^= 1u << s;
        const unsigned char* mine = mine0 + s * SLOT;
        auto load_e = [&](int b, double (&e)[NL]) {
#pragma unroll
            for (int p = 0; p < NPL; p++) {
                int pos = b * NPL + p + lrm;        // < 2 RUN
                pos = pos >= RUN ? pos - RUN : pos;
                const double2 lo = *(const double2*)(mine + pos * 32), hi = *(const double2*)(mine + pos * 32 + 16);
                e[4 * p] = lo.x; e[4 * p + 1] = lo.y; e[4 * p + 2] = hi.x; e[4 * p + 3] = hi.y;
            }
        };
        if (FAST && c >= 1) {
            if constexpr (FAST) {
            // Four ordinary steps as ONE basic block: the branch-free step, lanes past their sequence's end riding along
            // (stores predicated off).  Steps whose backpointer needs the tie analysis are flagged; the (rare) chunk that has
            // one is redone from the column it started with by the full step.
            double T1s[NL];
#pragma unroll
            for (int i = 0; i < NL; i++) T1s[i] = T1[i];
            const double m1s = m1;
            int back[4];
            unsigned ties = 0;
#pragma unroll
            for (int b = 0; b < 4; b++) {
                double e[NL];
                load_e(b, e);
                bool tie;
                back[b] = sgb_seq_step_fast<NL>(*(double(*)[NL])T1, e, m1, tr_self, tr_off, tie);
                ties |= (tie && 4 * c + b < len) ? 1u << b : 0u;
            }
            if (__any_sync(SGB_FULL, ties != 0)) {
                if (ties) {
                    // this lane redoes its steps up to the last flagged one: the full step where flagged, else the lean one
                    double m1r = m1s;
#pragma unroll 1
                    for (int b = 0; (ties >> b) != 0; b++) {
                        double e[NL];
                        load_e(b, e);
                        if ((ties >> b) & 1u) {
                            const int bk = sgb_seq_step_r<NL>(T1s, e, m1r, tr_self, tr_off);
                            back[0] = (b == 0) ? bk : back[0]; back[1] = (b == 1) ? bk : back[1];
                            back[2] = (b == 2) ? bk : back[2]; back[3] = (b == 3) ? bk : back[3];
                        } else {
                            bool tie_;
                            sgb_seq_step_fast<NL>(T1s, e, m1r, tr_self, tr_off, tie_);
                        }
                    }
                }
                __syncwarp();
            }
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int t = 4 * c + b;
                if (STAGED) stage[(t & 31) * SGB_STAGE_STRIDE + lane] = (unsigned char)back[b];
                else if (t < len) outp[t - 1] = (uint8_t)back[b];
                lastback = (t == len - 1) ? back[b] : lastback;
            }
            }
        } else {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int t = 4 * c + b;
                if (t < len) {
                    double e[NL];
                    load_e(b, e);
                    int back = 0;
                    if (DD) {
                        if (t == 0) sgb_seq_init_r<NL>(*(double(*)[NL])T1, e, init, m1);
                        else back = sgb_seq_step_r<NL>(*(double(*)[NL])T1, e, m1, tr_self, tr_off);
                    } else {
                        if (t == 0) sgb_seq_init<NL>(*(SgbSeqState<NL>*)&st, e, Seff, init);
                        else back = sgb_seq_step_general<NL>(*(SgbSeqState<NL>*)&st, e, Seff, tr_self, tr_off);
                    }
                    if (STAGED) stage[(t & 31) * SGB_STAGE_STRIDE + lane] = (unsigned char)back;
                    else if (t > 0) outp[t - 1] = (uint8_t)back;
                    lastback = (t == len - 1) ? back : lastback;
                }
            }
        }
        __syncwarp();                       // every lane has read slot s: it can be refilled
        if (c + depth < nchunks) issue(c + depth, s);
        if (STAGED && ((c & 7) == 7 || c == nchunks - 1)) {
            // flush: bins t = 32 k + lane of sequence j hold path[t - 1] (the decode rule writes one bin behind)
            const int t = (c >> 3) * 32 + lane;
            uint8_t* const outt = A.state_out + (t - 1);
#pragma unroll
            for (int j0 = 0; j0 < 32; j0 += 8) {
                int len_j[8];
                long long o_j[8];
                unsigned char v[8];
#pragma unroll
                for (int j = 0; j < 8; j++) {             // (eight sequences' shuffles and reads in flight together)
                    len_j[j] = __shfl_sync(SGB_FULL, len, j0 + j);
                    o_j[j] = __shfl_sync(SGB_FULL, ostart, j0 + j);
                    v[j] = stage[lane * SGB_STAGE_STRIDE + j0 + j];
                }
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (t >= 1 && t < len_j[j]) outt[o_j[j]] = v[j];
            }
            __syncwarp();
        }
        s = (s + 1 == depth) ? 0 : s + 1;
    }
    if (len > 0) outp[len - 1] = (uint8_t)lastback;     // path[T-1] = path[T-2]; T == 1: 0 (src/code.cpp:37)
}

template <int NPL, bool STAGED>
__device__ __noinline__ void recur_group_general(const SgbRecurArgs& A, const SgbScalars& sp, unsigned char* ring, unsigned long long* bar,
                                                 unsigned& phases, int depth, int lane, int Tmax, long long eoff, int len,
                                                 long long ostart, unsigned char* stage)
{
    recur_group_t<NPL, 0, STAGED>(A, sp, ring, bar, phases, depth, lane, Tmax, eoff, len, ostart, stage);
}

template <int NPL, bool STAGED>
__device__ __forceinline__ void recur_group(const SgbRecurArgs& A, const SgbScalars& sp, unsigned char* ring, unsigned long long* bar,
                                            unsigned& phases, int ring_bytes, int lane, int Tmax, long long eoff, int len,
                                            long long ostart, unsigned char* stage)
{
    int depth = ring_bytes / (NPL * 4096);
    depth = depth > SGB_RECUR_MAXDEPTH ? SGB_RECUR_MAXDEPTH : depth;
    if (sp.diag_dominant) recur_group_t<NPL, 1, STAGED>(A, sp, ring, bar, phases, depth, lane, Tmax, eoff, len, ostart, stage);
    else recur_group_general<NPL, STAGED>(A, sp, ring, bar, phases, depth, lane, Tmax, eoff, len, ostart, stage);
}

#ifdef SGB_RECUR_TIMING
static __device__ unsigned long long gs_dbg[4 * 8192];
extern "C" __attribute__((visibility("default"))) int sgb_debug_recur_read(unsigned long long* dst, int n)
{
    return (int)cudaMemcpyFromSymbol(dst, gs_dbg, (size_t)n * 8);
}
__device__ __forceinline__ unsigned long long sgb_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif
template <int MAXNPL, bool STAGED>
__global__ void __launch_bounds__(SGB_RECUR_NWARPS(MAXNPL) * 32, MAXNPL <= 3 ? 2 : 1) k_recur(SgbRecurArgs A, int ring_bytes)
{
    extern __shared__ __align__(128) unsigned char rsm[];
    __shared__ __align__(8) unsigned long long bars[SGB_RECUR_WARPS][SGB_RECUR_MAXDEPTH];
    __shared__ __align__(16) unsigned char stages[SGB_RECUR_WARPS][STAGED ? SGB_STAGE_BYTES : 16];
    __shared__ SgbScalars s_sp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned char* ring = rsm + (size_t)wid * ring_bytes;
    unsigned long long* bar = bars[wid];
    if (threadIdx.x == 0) s_sp = A.tabs->sp;
    if (lane < SGB_RECUR_MAXDEPTH) sgb_mbar_init(&bar[lane], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncthreads();
    SGB_PDL_ENTER();
    const SgbScalars& sp = s_sp;
    unsigned phases = 0;              // bit s: parity of the next completed phase of slot s's barrier
    const int ng = *A.W.ngroups;
    // Groups are ordered heaviest first (plane count, then length).  The first one of every warp is assigned by position, strided
    // over the CTAs and rotated over the warp slots, so that the heavy groups land on different SMs and different
    // schedulers (handed out by arrival order they ended up eight to an SM); further groups come from a counter.
    const int nw = (int)(blockDim.x >> 5);
    bool first = true;
    for (;;) {
        int g = 0;
        if (first) g = ((wid + (int)blockIdx.x) % nw) * (int)gridDim.x + (int)blockIdx.x;
        else {
            if (lane == 0) g = (int)atomicAdd(&A.W.counters[1], 1u) + nw * (int)gridDim.x;
            g = __shfl_sync(SGB_FULL, g, 0);
        }
        first = false;
        if (g >= ng) break;
        const int npl = A.W.g_npl[g];
        const int Tmax = A.W.g_tmax[g];
        const long long eoff = A.W.g_eoff[g];
        const int idx = A.W.perm[(long long)g * 32 + lane];
        int len = 0;
        long long ostart = 0;
        if (idx >= 0) { len = A.plan.len[idx]; ostart = A.plan.ostart[idx]; }
        if (Tmax <= 0 || npl <= 0) continue;
#ifdef SGB_RECUR_TIMING
        const unsigned long long dbg_t0 = sgb_gtime();
        const long long dbg_c0 = clock64();
#endif
#define SGB_RG(N) recur_group<N, STAGED>(A, sp, ring, bar, phases, ring_bytes, lane, Tmax, eoff, len, ostart, stages[wid])
        if (npl == 1) SGB_RG(1);
        else if (npl == 2) SGB_RG(2);
        else if (npl == 3) SGB_RG(3);
        else if (MAXNPL > 3) {
            if (npl == 4) SGB_RG(4);
            else if (npl == 6) SGB_RG(6);
            else SGB_RG(8);
        }
#undef SGB_RG
        __syncwarp();
#ifdef SGB_RECUR_TIMING
        if (lane == 0 && g < 8192) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            gs_dbg[4 * g] = (unsigned long long)npl | ((unsigned long long)(Tmax & 0xffff) << 8) | ((unsigned long long)(smid & 0xff) << 24);
            gs_dbg[4 * g + 1] = dbg_t0;
            gs_dbg[4 * g + 2] = sgb_gtime();
            gs_dbg[4 * g + 3] = (unsigned long long)(clock64() - dbg_c0);
        }
#endif
    }
}

// ------------------------------------------------------------------ host side
static bool pdl_enabled()
{
    const char* e = getenv("SGB_PDL");
    return !(e && *e == '0');
}
template <typename... KArgs, typename... Args>
static cudaError_t launch_dep(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

SgbSplitSizes sgbk_split_sizes(long long nseq, long long rows, int maxlen, int S)
{
    SgbSplitSizes z;
    z.gcap = (int)((nseq + 31) / 32) + SGB_NCLS;
    z.ntile = (int)((nseq + SGB_PART_TILE - 1) / SGB_PART_TILE);
    if (z.ntile < 1) z.ntile = 1;
    z.cls = (size_t)nseq + 16;
    z.perm = (size_t)z.gcap * 32 * sizeof(int);
    z.part = (size_t)z.ntile * SGB_NCLS * sizeof(unsigned int);
    z.g_i32 = (size_t)z.gcap * sizeof(int);
    z.g_i64 = (size_t)z.gcap * sizeof(long long);
    // Within a class the sequences keep the plan's length-descending order, so 32 x (longest of a group) is at most
    // the bins of the previous group of that class: sum over groups of 32 x roundup8(Tmax) <= rows + per-class and
    // per-group slack.  Every tile has at most planes_for(S) planes of 32-byte sectors.
    const long long slack = 32ll * (((long long)maxlen + 7) & ~7ll) * SGB_NCLS + 7ll * 32ll * z.gcap;
    z.e_bytes = (size_t)(rows + slack) * (size_t)sgb_planes_for(S) * 32;
    return z;
}

// smem_max: the largest dynamic shared-memory size any call of this kernel uses (the attribute is a per-device LIMIT:
// it is set once, to the maximum, so that calls with different table sizes never lower it under one another)
template <typename K>
static int split_occupancy(K kern, int threads, size_t smem, size_t smem_max, int* cache)
{
    int dev = 0;
    cudaGetDevice(&dev);
    int& slot = cache[dev & 63];
    if (slot > 0) return slot;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max) != cudaSuccess) { cudaGetLastError(); return 0; }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    slot = occ;
    return occ;
}

template <int MODEL, int LIKBB, int KIND>
static cudaError_t launch_emit_k(const SgbEmitArgs& A, long long nseq, int sm_count, cudaStream_t st)
{
    static int cache[64][SGB_MAX_STATES + 1];            // per device and S (the table size depends on S)
    static int flat[64];
    int dev = 0;
    cudaGetDevice(&dev);
    const size_t smem = emit_smem_bytes(A.S, LIKBB);
    flat[dev & 63] = cache[dev & 63][A.S];
    const int occ = split_occupancy(k_emit<MODEL, LIKBB, KIND>, SGB_EMIT_THREADS, smem, emit_smem_bytes(SGB_MAX_STATES, LIKBB), flat);
    cache[dev & 63][A.S] = occ;
    if (occ < 1) return cudaErrorInvalidConfiguration;
    long long blocks = (nseq + 7) / 8;
    if (blocks > (long long)sm_count * occ) blocks = (long long)sm_count * occ;
    if (blocks < 1) blocks = 1;
    return launch_dep(k_emit<MODEL, LIKBB, KIND>, (unsigned)blocks, SGB_EMIT_THREADS, smem, st, A);
}
template <int MODEL, int LIKBB>
static cudaError_t launch_emit(const SgbEmitArgs& A, long long nseq, int sm_count, cudaStream_t st)
{
    switch (A.in.kind) {
    case SGB_IN_I32_U8: return launch_emit_k<MODEL, LIKBB, SGB_IN_I32_U8>(A, nseq, sm_count, st);
    case SGB_IN_I32_I32: return launch_emit_k<MODEL, LIKBB, SGB_IN_I32_I32>(A, nseq, sm_count, st);
    case SGB_IN_U16_U8: return launch_emit_k<MODEL, LIKBB, SGB_IN_U16_U8>(A, nseq, sm_count, st);
    default: return launch_emit_k<MODEL, LIKBB, -1>(A, nseq, sm_count, st);
    }
}

template <int MAXNPL, bool STAGED>
static cudaError_t launch_recur_k(const SgbRecurArgs& A, int gcap, int sm_count, cudaStream_t st)
{
    // Ring size per warp: the TMA copies must run far enough ahead of the recursion to cover the memory latency (~1 us)
    // at ~100-250 cycles per bin, i.e. >= 8 bins whatever the plane count (a chunk is 4 bins x planes KB).  Small
    // batches (few warps per SM; the chain latency is all that matters) get deep rings, large ones trade depth for
    // resident warps (HBM-bound: what counts is the number of bytes in flight per SM).
    static int cache[2][64];
    constexpr int NW = SGB_RECUR_NWARPS(MAXNPL);
    const int big = MAXNPL <= 3 ? 48 : 96, small = MAXNPL <= 3 ? 24 : 64;
    const int ci = gcap <= sm_count * NW ? 0 : 1;
    const int ring_kb = ci == 0 ? big : small;
    const size_t smem = (size_t)NW * ring_kb * 1024;
    const int occ = split_occupancy(k_recur<MAXNPL, STAGED>, NW * 32, smem, (size_t)NW * big * 1024, cache[ci]);
    if (occ < 1) return cudaErrorInvalidConfiguration;
    long long blocks = ((long long)gcap + NW - 1) / NW;
    if (blocks > (long long)sm_count * occ) blocks = (long long)sm_count * occ;
    if (blocks < 1) blocks = 1;
    return launch_dep(k_recur<MAXNPL, STAGED>, (unsigned)blocks, NW * 32, smem, st, A, ring_kb * 1024);
}

// Batches that fill the device several times over are bound by the L1's store path unless the decoded states are staged
// (one byte per lane per bin = 32 sectors per store instruction); a batch of at most one wave of groups is bound by the
// latency of its longest chain, where the flush of the staging buffer would only add to it.
template <int MAXNPL>
static cudaError_t launch_recur(const SgbRecurArgs& A, int gcap, int sm_count, cudaStream_t st)
{
    const char* e = getenv("SGB_RECUR_STAGED");          // 0 / 1 forces a variant (tests)
    const bool staged = (e && *e) ? atoi(e) != 0 : gcap > 2 * sm_count * SGB_RECUR_NWARPS(MAXNPL);
    return staged ? launch_recur_k<MAXNPL, true>(A, gcap, sm_count, st) : launch_recur_k<MAXNPL, false>(A, gcap, sm_count, st);
}

cudaError_t sgbk_split_decode(const SgbPlanDev& plan, long long nseq, const SgbColsIn& in, const SgbDevTables* d_tabs,
                              const SgbModelArgs& m, const SgbSplitDev& W, uint8_t* d_state_out, int sm_count,
                              cudaStream_t st, cudaEvent_t ev_prep, cudaEvent_t ev_mid)
{
    if (nseq <= 0) return cudaSuccess;
    cudaError_t e;
    if ((e = cudaMemsetAsync(W.part, 0, (size_t)W.ntile * SGB_NCLS * sizeof(unsigned int), st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(W.perm, 0xff, (size_t)W.gcap * 32 * sizeof(int), st)) != cudaSuccess) return e;
    long long cb = (nseq + 7) / 8;
    if (cb > (long long)sm_count * 8) cb = (long long)sm_count * 8;
    k_split_class<<<(unsigned)cb, 256, 0, st>>>(plan, nseq, in.cn, in.kind, m.model, m.S, W);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if ((e = launch_dep(k_split_scan, 1u, 256u, 0, st, W)) != cudaSuccess) return e;
    if ((e = launch_dep(k_split_scatter, (unsigned)W.ntile, (unsigned)SGB_PART_TILE, 0, st, nseq, W)) != cudaSuccess) return e;
    if ((e = launch_dep(k_split_groups, (unsigned)((W.gcap + 7) / 8), 256u, 0, st, plan, W)) != cudaSuccess) return e;
    if ((e = launch_dep(k_split_gscan, 1u, 1024u, 0, st, W, in.err)) != cudaSuccess) return e;
    if (ev_prep && (e = cudaEventRecord(ev_prep, st)) != cudaSuccess) return e;
    SgbEmitArgs EA;
    EA.plan = plan; EA.W = W; EA.in = in; EA.tabs = d_tabs; EA.S = m.S;
    EA.lik = m.lik_bb ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL; EA.loherror = m.loherror; EA.rho = m.rho;
    if (m.model == SGB_MODEL_HSCN && !m.lik_bb) e = launch_emit<0, 0>(EA, nseq, sm_count, st);
    else if (m.model == SGB_MODEL_HSCN) e = launch_emit<0, 1>(EA, nseq, sm_count, st);
    else if (!m.lik_bb) e = launch_emit<1, 0>(EA, nseq, sm_count, st);
    else e = launch_emit<1, 1>(EA, nseq, sm_count, st);
    if (e != cudaSuccess) return e;
    if (ev_mid && (e = cudaEventRecord(ev_mid, st)) != cudaSuccess) return e;
    SgbRecurArgs RA;
    RA.plan = plan; RA.W = W; RA.tabs = d_tabs; RA.state_out = d_state_out; RA.S = m.S;
    return m.S <= 12 ? launch_recur<3>(RA, W.gcap, sm_count, st) : launch_recur<8>(RA, W.gcap, sm_count, st);
}
