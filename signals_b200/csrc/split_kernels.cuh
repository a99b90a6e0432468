// split_kernels.cuh -- the SPLIT schedule of the HSCN / ASCN decode (included by hmm_kernels.cu).
//
// The fused k_decode gives every (cell, chromosome) sequence ONE lane, so a batch of n sequences offers n/32 warps and
// every warp walks a dependent chain of emission + recursion steps: batches below ~75,000 sequences cannot fill a
// B200 (148 SMs x 16+ resident warps).  The emission log-likelihoods (R/callHSCN.R:48-80, R/callASCN.R:50-81) are
// independent per bin; only the max-plus recursion (src/code.cpp:27-33) is sequential.  So:
//
//   k_split_class    one warp per sequence: largest total copy number -> number of 4-state "planes" the sequence needs
//   k_split_scan / k_split_scatter   stable partition of the sequences by plane count (descending), so that the 32
//                    sequences of a recursion warp need the same number of planes (and warps stay homogeneous)
//   k_split_groups / k_split_gscan   per group of 32 sequences: plane count, longest sequence, tile offset
//   k_emit           one warp per sequence, LANES OVER BINS: coalesced loads of the caller's columns (any of the
//                    accepted element types), all live states of a bin evaluated by the lane that owns the bin
//                    (copy number is piecewise constant along a chromosome, so a warp is branch-uniform), emissions
//                    written as 32-byte sectors into the group's tile  E[bin][plane][lane][4]
//   k_recur          one warp per group, one lane per sequence: the tile is streamed through a shared-memory ring by
//                    1-D TMA bulk copies (cp.async.bulk + mbarrier, several bins ahead), the forward column lives
//                    in registers, per bin the O(S) step of hmm_core.cuh and the reference's decode rule
//
// k_emit is FP64 / issue bound and has N x L-way parallelism; k_recur is HBM-bound (reads 32 B per cell-bin and
// plane) and its dependent chain per bin is ~50 instructions instead of a whole emission.  Same arithmetic functions
// as the fused kernel (sgb_emit_item, the max-plus step), hence the same bits.
#pragma once

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
        for (int t = lane; t < len; t += 32) {
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
#define SGB_EMIT_LF 1024
static inline size_t emit_smem_bytes(int S, int lik_bb)
{
    size_t d = SGB_LOGTAB_LEN + SGB_EXPTAB_LEN + SGB_EMIT_LF + (size_t)SGB_CN_TABLE * S * (lik_bb ? 5 : 2);
    return d * sizeof(double) + (size_t)SGB_CN_TABLE * S * sizeof(int);
}

template <int MODEL, int LIKBB>
__global__ void __launch_bounds__(SGB_EMIT_THREADS, LIKBB ? 2 : 3) k_emit(SgbEmitArgs A)
{
    extern __shared__ __align__(16) double smem[];
    const int S = A.S;
    const int CTS = SGB_CN_TABLE * S;
    double* sLT = smem;
    double* sET = sLT + SGB_LOGTAB_LEN;
    double* sLF = sET + SGB_EXPTAB_LEN;
    double* sLP = sLF + SGB_EMIT_LF;
    double* sA = sLP + 2 * CTS;
    double* sB = sA + (LIKBB ? CTS : 0);
    double* sL0 = sB + (LIKBB ? CTS : 0);
    int* sKind = (int*)(sL0 + (LIKBB ? CTS : 0));
    for (int i = threadIdx.x; i < SGB_LOGTAB_LEN; i += blockDim.x) sLT[i] = g_logtab[i];
    for (int i = threadIdx.x; i < SGB_EXPTAB_LEN; i += blockDim.x) sET[i] = g_exptab[i];
    for (int i = threadIdx.x; i < SGB_EMIT_LF; i += blockDim.x) sLF[i] = g_lftab[i];
    for (int i = threadIdx.x; i < CTS; i += blockDim.x) {
        const int c = i / S, s = i - c * S, src = c * SGB_MAX_STATES + s;
        sLP[2 * i] = A.tabs->lp[src];
        sLP[2 * i + 1] = A.tabs->lq[src];
        sKind[i] = A.tabs->kind[src];
        if (LIKBB) { sA[i] = A.tabs->a[src]; sB[i] = A.tabs->b[src]; sL0[i] = A.tabs->l0[src]; }
    }
    __shared__ SgbScalars s_sp;
    __shared__ SgbTables s_TB;
    if (threadIdx.x == 0) {
        s_sp = A.tabs->sp;
        s_TB.LT = sLT; s_TB.ET = sET; s_TB.LP = sLP; s_TB.LQ = sLP + 1; s_TB.pstride = 2; s_TB.A = sA; s_TB.B = sB; s_TB.L0 = sL0;
        s_TB.LGK = A.tabs->lgk; s_TB.KIND = sKind; s_TB.S = S; s_TB.lik = A.lik; s_TB.loherror = A.loherror; s_TB.rho = A.rho;
    }
    __syncthreads();
    const SgbScalars& sp = s_sp;
    const SgbTables& TB = s_TB;
    const int lane = threadIdx.x & 31;
    const double NEG_INF = -sgb_inf();
    const int kind = A.in.kind;
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
        double* tile = A.W.E + (A.W.g_eoff[g] * 4 + lr * 4);          // + (t * npl + p) * 128 doubles
        for (int t0 = 0; t0 < len; t0 += 32) {
            const int t = t0 + lane;
            const bool act = t < len;
            int n = 0, x = 0, cnv = 0;
            if (act) {
                const long long row = o + t;
                if (kind == SGB_IN_U16_U8) {
                    n = (int)__ldg((const uint16_t*)A.in.n + row); x = (int)__ldg((const uint16_t*)A.in.x + row);
                    cnv = (int)__ldg((const uint8_t*)A.in.cn + row);
                } else if (kind == SGB_IN_F64) {
                    const double dn = __ldg((const double*)A.in.n + row), dx = __ldg((const double*)A.in.x + row);
                    const double dc = __ldg((const double*)A.in.cn + row);
                    const double lim = (double)(1 << SGB_COUNT_BITS);
                    if (!(dn >= 0.0) || !(dx >= 0.0)) bad |= (dn != dn || dx != dx) ? SGB_ERR_NOT_INTEGER : SGB_ERR_NEGATIVE_COUNT;
                    else if (dn >= lim || dx >= lim) bad |= SGB_ERR_COUNT_TOO_LARGE;
                    else if (dn != rint(dn) || dx != rint(dx)) bad |= SGB_ERR_NOT_INTEGER;
                    if (!(dc >= 0.0 && dc <= 255.0)) bad |= (dc != dc) ? SGB_ERR_NOT_INTEGER : SGB_ERR_BAD_CN;
                    else if (dc != rint(dc)) bad |= SGB_ERR_NOT_INTEGER;
                    const bool ok = dn >= 0.0 && dx >= 0.0 && dn < lim && dx < lim && dn == rint(dn) && dx == rint(dx) &&
                                    dc >= 0.0 && dc <= 255.0 && dc == rint(dc);
                    n = ok ? (int)dn : 0; x = ok ? (int)dx : 0; cnv = ok ? (int)dc : 0;
                } else {
                    n = __ldg((const int32_t*)A.in.n + row); x = __ldg((const int32_t*)A.in.x + row);
                    cnv = kind == SGB_IN_I32_I32 ? __ldg((const int32_t*)A.in.cn + row) : (int)__ldg((const uint8_t*)A.in.cn + row);
                }
                // the checks of the fused path: a bad bin is neutralised, the call fails afterwards
                if ((unsigned)(n | x) >> SGB_COUNT_BITS) {
                    bad |= (n | x) < 0 ? SGB_ERR_NEGATIVE_COUNT : SGB_ERR_COUNT_TOO_LARGE;
                    n = 0; x = 0; cnv = 0;
                }
                if ((unsigned)cnv > 255u) { bad |= SGB_ERR_BAD_CN; n = 0; x = 0; cnv = 0; }
            }
            const int cprime = sgb_cprime(MODEL, cnv);
            const int ins = x <= n;
            const int L = act ? sgb_live_states(MODEL, cprime, S) : 0;
            double lcx = 0.0, lcy = 0.0, lgs = 0.0;
            if (act && ins) {
                double fn, fx, fy;
                if (n < SGB_EMIT_LF) { fn = sLF[n]; fx = sLF[x]; fy = sLF[n - x]; }
                else if (n < SGB_NLF) { fn = g_lftab[n]; fx = g_lftab[x]; fy = g_lftab[n - x]; }
                else { fn = sgb_logfact(n, g_lftab, sLT); fx = sgb_logfact(x, g_lftab, sLT); fy = sgb_logfact(n - x, g_lftab, sLT); }
                lcx = sgb_sub(sgb_sub(fn, fx), fy);
                if (MODEL == SGB_MODEL_ASCN) lcy = sgb_sub(sgb_sub(fn, fy), fx);
                if (LIKBB) lgs = sgb_lgamma_pos(sgb_add(sp.bb_s, (double)n), sLT);
            }
            const double kd = (double)x, nkd = (double)(n - x);
            double* dst = tile + (long long)t * npl * 128;
            const int Lw = __reduce_max_sync(SGB_FULL, L);
            for (int p = 0; p < npl; p++) {
                double e0 = NEG_INF, e1 = NEG_INF, e2 = NEG_INF, e3 = NEG_INF;
                if (4 * p < Lw) {
                    if (4 * p + 0 < L) e0 = sgb_emit_item<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 0, TB, sp);
                    if (4 * p + 1 < L) e1 = sgb_emit_item<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 1, TB, sp);
                    if (4 * p + 2 < L) e2 = sgb_emit_item<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 2, TB, sp);
                    if (4 * p + 3 < L) e3 = sgb_emit_item<MODEL, LIKBB>(cprime, ins, kd, nkd, lcx, lcy, lgs, 4 * p + 3, TB, sp);
                }
                if (act) {
                    __stcg((double2*)(dst + p * 128), make_double2(e0, e1));
                    __stcg((double2*)(dst + p * 128 + 2), make_double2(e2, e3));
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

#define SGB_RECUR_WARPS 4
#define SGB_RECUR_STAGES 4

// DD = 1: diagonal-dominant transition (every realistic selftransitionprob), forward column in registers, O(S) step.
// DD = 0: selftransitionprob below 1/S -- the two-maxima variant, kept out of line so that it does not set the hot
// path's register count.
template <int NPL, int STAGE_BYTES, int DD>
__device__ __forceinline__ void recur_group_t(const SgbRecurArgs& A, const SgbScalars& sp, unsigned char* ring, unsigned long long* bar,
                                              unsigned& phases, int lane, int Tmax, long long eoff, int len, long long ostart)
{
    constexpr int NL = 4 * NPL;
    constexpr int BPS = (STAGE_BYTES / (NPL * 1024)) > 0 ? (STAGE_BYTES / (NPL * 1024)) : 1;   // bins per stage
    constexpr unsigned CHUNK = BPS * NPL * 1024;
    const unsigned char* src = (const unsigned char*)A.W.E + eoff * 32;
    const int nchunks = (Tmax + BPS - 1) / BPS;
    auto issue = [&](int c) {
        if (lane == 0) {
            const int s = c % SGB_RECUR_STAGES;
            sgb_mbar_expect_tx(&bar[s], CHUNK);
            sgb_tma_load_1d(ring + s * STAGE_BYTES, src + (long long)c * CHUNK, CHUNK, &bar[s]);
        }
    };
    for (int c = 0; c < SGB_RECUR_STAGES && c < nchunks; c++) issue(c);
    double T1[DD ? NL : 1];
    double m1 = 0.0;
    SgbSeqState<DD ? 1 : NL> st;
    const int Seff = A.S < NL ? A.S : NL;
    const double tr_self = sp.tr_self, tr_off = sp.tr_off, init = sp.init;
    for (int c = 0; c < nchunks; c++) {
        const int s = c % SGB_RECUR_STAGES;
        sgb_mbar_wait(&bar[s], (phases >> s) & 1u);
        phases ^= 1u << s;
#pragma unroll
        for (int b = 0; b < BPS; b++) {
            const int t = c * BPS + b;
            if (t < len) {
                double e[NL];
                const double2* row = (const double2*)(ring + s * STAGE_BYTES + b * NPL * 1024 + lane * 32);
#pragma unroll
                for (int p = 0; p < NPL; p++) {
                    const double2 v0 = row[p * 64], v1 = row[p * 64 + 1];
                    e[4 * p] = v0.x; e[4 * p + 1] = v0.y; e[4 * p + 2] = v1.x; e[4 * p + 3] = v1.y;
                }
                int back = 0;
                if (DD) {
                    if (t == 0) sgb_seq_init_r<NL>(*(double(*)[NL])T1, e, init, m1);
                    else back = sgb_seq_step_r<NL>(*(double(*)[NL])T1, e, m1, tr_self, tr_off);
                } else {
                    if (t == 0) sgb_seq_init<NL>(*(SgbSeqState<NL>*)&st, e, Seff, init);
                    else back = sgb_seq_step_general<NL>(*(SgbSeqState<NL>*)&st, e, Seff, tr_self, tr_off);
                }
                if (t == 0) {
                    if (len == 1) A.state_out[ostart] = 0;
                } else {
                    A.state_out[ostart + t - 1] = (uint8_t)back;
                    if (t == len - 1) A.state_out[ostart + t] = (uint8_t)back;
                }
            }
        }
        __syncwarp();
        if (c + SGB_RECUR_STAGES < nchunks) issue(c + SGB_RECUR_STAGES);
    }
}

template <int NPL, int STAGE_BYTES>
__device__ __noinline__ void recur_group_general(const SgbRecurArgs& A, const SgbScalars& sp, unsigned char* ring, unsigned long long* bar,
                                                 unsigned& phases, int lane, int Tmax, long long eoff, int len, long long ostart)
{
    recur_group_t<NPL, STAGE_BYTES, 0>(A, sp, ring, bar, phases, lane, Tmax, eoff, len, ostart);
}

template <int NPL, int STAGE_BYTES>
__device__ __forceinline__ void recur_group(const SgbRecurArgs& A, const SgbScalars& sp, unsigned char* ring, unsigned long long* bar,
                                            unsigned& phases, int lane, int Tmax, long long eoff, int len, long long ostart)
{
    if (sp.diag_dominant) recur_group_t<NPL, STAGE_BYTES, 1>(A, sp, ring, bar, phases, lane, Tmax, eoff, len, ostart);
    else recur_group_general<NPL, STAGE_BYTES>(A, sp, ring, bar, phases, lane, Tmax, eoff, len, ostart);
}

template <int MAXNPL>
__global__ void __launch_bounds__(SGB_RECUR_WARPS * 32, MAXNPL <= 3 ? 3 : 1) k_recur(SgbRecurArgs A)
{
    constexpr int STAGE_BYTES = MAXNPL <= 3 ? 4096 : 8192;
    extern __shared__ __align__(128) unsigned char rsm[];
    __shared__ __align__(8) unsigned long long bars[SGB_RECUR_WARPS][SGB_RECUR_STAGES];
    __shared__ SgbScalars s_sp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned char* ring = rsm + (size_t)wid * SGB_RECUR_STAGES * STAGE_BYTES;
    unsigned long long* bar = bars[wid];
    if (threadIdx.x == 0) s_sp = A.tabs->sp;
    if (lane == 0) {
        for (int s = 0; s < SGB_RECUR_STAGES; s++) sgb_mbar_init(&bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncthreads();
    const SgbScalars& sp = s_sp;
    unsigned phases = 0;              // bit s: parity of the next completed phase of stage s's barrier
    const int ng = *A.W.ngroups;
    for (;;) {
        int g = 0;
        if (lane == 0) g = (int)atomicAdd(&A.W.counters[1], 1u);
        g = __shfl_sync(SGB_FULL, g, 0);
        if (g >= ng) break;
        const int npl = A.W.g_npl[g];
        const int Tmax = A.W.g_tmax[g];
        const long long eoff = A.W.g_eoff[g];
        const int idx = A.W.perm[(long long)g * 32 + lane];
        int len = 0;
        long long ostart = 0;
        if (idx >= 0) { len = A.plan.len[idx]; ostart = A.plan.ostart[idx]; }
        if (Tmax <= 0 || npl <= 0) continue;
#define SGB_RG(N) recur_group<N, STAGE_BYTES>(A, sp, ring, bar, phases, lane, Tmax, eoff, len, ostart)
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
    }
}

// ------------------------------------------------------------------ host side
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

template <typename K>
static int split_occupancy(K kern, int threads, size_t smem, int* cache)
{
    int dev = 0;
    cudaGetDevice(&dev);
    int& slot = cache[dev & 63];
    if (slot > 0) return slot;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    slot = occ;
    return occ;
}

template <int MODEL, int LIKBB>
static cudaError_t launch_emit(const SgbEmitArgs& A, long long nseq, int sm_count, cudaStream_t st)
{
    static int cache[64][SGB_MAX_STATES + 1];            // per device and S (the table size depends on S)
    static int flat[64];
    int dev = 0;
    cudaGetDevice(&dev);
    const size_t smem = emit_smem_bytes(A.S, LIKBB);
    flat[dev & 63] = cache[dev & 63][A.S];
    const int occ = split_occupancy(k_emit<MODEL, LIKBB>, SGB_EMIT_THREADS, smem, flat);
    cache[dev & 63][A.S] = occ;
    if (occ < 1) return cudaErrorInvalidConfiguration;
    long long blocks = (nseq + 7) / 8;
    if (blocks > (long long)sm_count * occ) blocks = (long long)sm_count * occ;
    if (blocks < 1) blocks = 1;
    k_emit<MODEL, LIKBB><<<(unsigned)blocks, SGB_EMIT_THREADS, smem, st>>>(A);
    return cudaGetLastError();
}

template <int MAXNPL>
static cudaError_t launch_recur(const SgbRecurArgs& A, int gcap, int sm_count, cudaStream_t st)
{
    static int cache[64];
    const size_t smem = (size_t)SGB_RECUR_WARPS * SGB_RECUR_STAGES * (MAXNPL <= 3 ? 4096 : 8192);
    const int occ = split_occupancy(k_recur<MAXNPL>, SGB_RECUR_WARPS * 32, smem, cache);
    if (occ < 1) return cudaErrorInvalidConfiguration;
    long long blocks = ((long long)gcap + SGB_RECUR_WARPS - 1) / SGB_RECUR_WARPS;
    if (blocks > (long long)sm_count * occ) blocks = (long long)sm_count * occ;
    if (blocks < 1) blocks = 1;
    k_recur<MAXNPL><<<(unsigned)blocks, SGB_RECUR_WARPS * 32, smem, st>>>(A);
    return cudaGetLastError();
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
    k_split_scan<<<1, 256, 0, st>>>(W);
    k_split_scatter<<<(unsigned)W.ntile, SGB_PART_TILE, 0, st>>>(nseq, W);
    k_split_groups<<<(unsigned)((W.gcap + 7) / 8), 256, 0, st>>>(plan, W);
    k_split_gscan<<<1, 1024, 0, st>>>(W, in.err);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
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
