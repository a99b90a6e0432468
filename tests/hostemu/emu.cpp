// Host emulation of the per-lane work of the fused decode kernel, built from the SAME
// hmm_core.cuh the CUDA kernels include (compiled here with g++, -ffp-contract=off).
// Test-only: lets the CPU test-suite check the kernel arithmetic / control flow against
// the oracle without a GPU.  Never linked into the product.
#include "../../signals_b200/csrc/hmm_core.cuh"
#include <vector>

static const double LT[SGB_LOGTAB_LEN] = SGB_LOGTAB_INIT;
static const double ET[SGB_EXPTAB_LEN] = SGB_EXPTAB_INIT;
static const double LF[SGB_LFTAB_LEN] = SGB_LFTAB_INIT;

template <int SMAX, int MODEL, int LIKBB>
static void run(int S, int decode, double loherror, double eps, double rho, double Abias, double stp, long T,
                const int* n, const int* x, const int* cn, unsigned char* out, double* l_out)
{
    const int lik = LIKBB ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL;
    SgbScalars sp = sgb_scalars(S, eps, Abias, stp, LIKBB ? rho : 0.0, LT);
    // the per-(CN, state) tables exactly as k_setup fills them
    std::vector<double> tLP(SGB_CN_TABLE * S), tLQ(SGB_CN_TABLE * S), tA(SGB_CN_TABLE * S), tB(SGB_CN_TABLE * S),
        tL0(SGB_CN_TABLE * S);
    std::vector<int> tK(SGB_CN_TABLE * S);
    std::vector<double> tG((size_t)SGB_CN_TABLE * S * 2 * SGB_LGK, 0.0);
    for (int c = 0; c < SGB_CN_TABLE; c++)
        for (int m = 0; m < S; m++) {
            SgbStateConst sc = sgb_state_const(c, m, loherror, lik, rho, sp.bb_s, LT);
            int i = c * S + m;
            tLP[i] = sc.lp; tLQ[i] = sc.lq; tA[i] = sc.a; tB[i] = sc.b; tL0[i] = sc.l0; tK[i] = sc.kind;
            if (LIKBB && sc.kind == SGB_KIND_NORMAL)
                for (int k = 0; k < SGB_LGK; k++) {
                    tG[(size_t)i * 2 * SGB_LGK + k] = sgb_lgamma_pos(sgb_add(sc.a, (double)k), LT);
                    tG[(size_t)i * 2 * SGB_LGK + SGB_LGK + k] = sgb_lgamma_pos(sgb_add(sc.b, (double)k), LT);
                }
        }
    SgbTables TB;
    TB.LT = LT; TB.ET = ET; TB.LP = tLP.data(); TB.LQ = tLQ.data(); TB.pstride = 1; TB.A = tA.data(); TB.B = tB.data();
    TB.L0 = tL0.data(); TB.LGK = tG.data(); TB.KIND = tK.data(); TB.S = S; TB.lik = lik; TB.loherror = loherror; TB.rho = rho;
    SgbSeqState<SMAX> st;
    double T1m[SMAX], m1m = 0.0;
    int Lkeep = SMAX;
    std::vector<unsigned char> bp((size_t)T * S);
    for (long t = 0; t < T; t++) {
        SgbBin b = sgb_bin_prepare(MODEL, LIKBB, n[t], x[t], cn[t], sp.bb_s, LF, LT);
        int L = sgb_live_states(MODEL, b.cprime, S);
        double e[SMAX];
        for (int i = 0; i < SMAX; i++) e[i] = -sgb_inf();
        auto lean = [&](int m) {
            return b.cprime < SGB_CN_TABLE && b.in_support && sp.fast_ok && tK[b.cprime * S + m] == SGB_KIND_NORMAL;
        };
        {
            // like the kernel: pairs of lean items go through the two-wide function, the rest one by one
            for (int i = 0; i < L; i += 2) {
                const int j = (i + 1 < L) ? i + 1 : i;
                if (lean(i) && lean(j)) {
                    const int t0 = b.cprime * S + i, t1 = b.cprime * S + j;
                    double r0, r1;
                    sgb_emission_fast2<MODEL, LIKBB>(b.kd, b.nkd, b.lc_x, b.lc_y, b.lg_sn, tLP[t0], tLQ[t0], tA[t0], tB[t0], tL0[t0],
                                                     b.kd, b.nkd, b.lc_x, b.lc_y, b.lg_sn, tLP[t1], tLQ[t1], tA[t1], tB[t1], tL0[t1],
                                                     tG.data() + (size_t)t0 * 2 * SGB_LGK, tG.data() + (size_t)t1 * 2 * SGB_LGK,
                                                     sp, LT, ET, r0, r1);
                    e[i] = r0; e[j] = r1;
                } else {
                    e[i] = sgb_emit_item<MODEL, LIKBB>(b.cprime, b.in_support, b.kd, b.nkd, b.lc_x, b.lc_y, b.lg_sn, i, TB, sp);
                    e[j] = sgb_emit_item<MODEL, LIKBB>(b.cprime, b.in_support, b.kd, b.nkd, b.lc_x, b.lc_y, b.lg_sn, j, TB, sp);
                }
            }
        }
        if (l_out) for (int i = 0; i < S; i++) l_out[t + T * i] = e[i];
        const int Lprev = Lkeep; Lkeep = L;
        const bool mem = (decode == SGB_DECODE_REFERENCE) && sp.diag_dominant;      // the kernel's fast path
        if (t == 0) {
            if (mem) sgb_seq_init_mem<SMAX>(T1m, 1, e, 1, S, sp.init, m1m);
            else sgb_seq_init<SMAX>(st, e, S, sp.init);
            continue;
        }
        if (mem) {
            const int nlive = (L > Lprev) ? L : Lprev;       // the kernel uses the warp-wide maxima of the same two counts
            out[t - 1] = (unsigned char)sgb_seq_step_mem<SMAX>(T1m, 1, e, 1, m1m, sp.tr_self, sp.tr_off, nlive);
        } else if (decode == SGB_DECODE_REFERENCE) {
            int back = sgb_seq_step<SMAX>(st, e, e, 1, S, sp.tr_self, sp.tr_off, sp.diag_dominant);
            out[t - 1] = (unsigned char)back;
        } else {
            const int nlive = (L > Lprev) ? L : Lprev;       // (the kernel: warp-wide maxima of the two counts)
            sgb_seq_step_chained<SMAX>(st, e, S, sp.tr_self, sp.tr_off, &bp[(size_t)t * S], nlive);
        }
    }
    if (decode == SGB_DECODE_REFERENCE) {
        out[T - 1] = (T > 1) ? out[T - 2] : 0;
    } else {
        int cur = st.i1;
        out[T - 1] = (unsigned char)cur;
        for (long t = T - 1; t > 0; t--) { cur = bp[(size_t)t * S + cur]; out[t - 1] = (unsigned char)cur; }
    }
}

template <int SMAX>
static void run_s(int model, int lik_bb, int S, int decode, double loherror, double eps, double rho, double Abias,
                  double stp, long T, const int* n, const int* x, const int* cn, unsigned char* out, double* l_out)
{
    if (model == 0 && !lik_bb) run<SMAX, 0, 0>(S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else if (model == 0) run<SMAX, 0, 1>(S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else if (!lik_bb) run<SMAX, 1, 0>(S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else run<SMAX, 1, 1>(S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
}

extern "C" int emu_decode_sequence(int model, int lik, int S, int decode, double loherror, double eps, double rho,
                                   double Abias, double stp, long T, const int* n, const int* x, const int* cn,
                                   unsigned char* out, double* l_out)
{
    if (T <= 0) return 0;
    const int lik_bb = (lik == SGB_LIK_BETABINOMIAL) && rho > 0.0;
    if (S <= 12) run_s<12>(model, lik_bb, S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else if (S <= 16) run_s<16>(model, lik_bb, S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else if (S <= 32) run_s<32>(model, lik_bb, S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else return -1;
    return 0;
}


// ---- the split schedule's per-lane work: emissions of the live states as k_emit writes them (planes of 4 states, -Inf
// beyond the bin's live states), then the register-resident recursion of k_recur with NL = 4 x planes_for(largest live
// count of the sequence) states.
template <int NPL, int MODEL, int LIKBB>
static void run_split(int S, double loherror, double eps, double rho, double Abias, double stp, long T, const int* n,
                      const int* x, const int* cn, unsigned char* out, int block)
{
    constexpr int NL = 4 * NPL;
    const int lik = LIKBB ? SGB_LIK_BETABINOMIAL : SGB_LIK_BINOMIAL;
    SgbScalars sp = sgb_scalars(S, eps, Abias, stp, LIKBB ? rho : 0.0, LT);
    std::vector<double> tLP(SGB_CN_TABLE * S), tLQ(SGB_CN_TABLE * S), tA(SGB_CN_TABLE * S), tB(SGB_CN_TABLE * S), tL0(SGB_CN_TABLE * S);
    std::vector<int> tK(SGB_CN_TABLE * S);
    std::vector<double> tG((size_t)SGB_CN_TABLE * S * 2 * SGB_LGK, 0.0);
    for (int c = 0; c < SGB_CN_TABLE; c++)
        for (int m = 0; m < S; m++) {
            SgbStateConst sc = sgb_state_const(c, m, loherror, lik, rho, sp.bb_s, LT);
            int i = c * S + m;
            tLP[i] = sc.lp; tLQ[i] = sc.lq; tA[i] = sc.a; tB[i] = sc.b; tL0[i] = sc.l0; tK[i] = sc.kind;
            if (LIKBB && sc.kind == SGB_KIND_NORMAL)
                for (int k = 0; k < SGB_LGK; k++) {
                    tG[(size_t)i * 2 * SGB_LGK + k] = sgb_lgamma_pos(sgb_add(sc.a, (double)k), LT);
                    tG[(size_t)i * 2 * SGB_LGK + SGB_LGK + k] = sgb_lgamma_pos(sgb_add(sc.b, (double)k), LT);
                }
        }
    SgbTables TB;
    TB.LT = LT; TB.ET = ET; TB.LP = tLP.data(); TB.LQ = tLQ.data(); TB.pstride = 1; TB.A = tA.data(); TB.B = tB.data();
    TB.L0 = tL0.data(); TB.LGK = tG.data(); TB.KIND = tK.data(); TB.S = S; TB.lik = lik; TB.loherror = loherror; TB.rho = rho;
    double T1[NL], m1 = 0.0;
    SgbSeqState<NL> st;
    const int Seff = S < NL ? S : NL;
    auto emit = [&](long t, double (&e)[NL]) {
        SgbBin b = sgb_bin_prepare(MODEL, LIKBB, n[t], x[t], cn[t], sp.bb_s, LF, LT);
        const int L = sgb_live_states(MODEL, b.cprime, S);
        for (int m = 0; m < NL; m++)
            e[m] = m < L ? sgb_emit_item<MODEL, LIKBB>(b.cprime, b.in_support, b.kd, b.nkd, b.lc_x, b.lc_y, b.lg_sn, m, TB, sp)
                         : -sgb_inf();
    };
    if (block && sp.diag_dominant) {
        // k_recur's variant for large batches: chunk 0 one step at a time, every later chunk of four bins with the
        // branch-free step; a chunk in which some step flags a possible off-diagonal predecessor is redone from the column
        // it started with -- the full step where flagged, the lean one elsewhere, up to the last flagged step.  Bins past
        // the end of the sequence ride along on garbage exactly as the lanes of a shorter sequence do (here: clamped reads).
        int lastback = 0;
        for (long c = 0; 4 * c < T; c++) {
            if (c == 0) {
                for (long t = 0; t < 4 && t < T; t++) {
                    double e[NL];
                    emit(t, e);
                    int back = 0;
                    if (t == 0) sgb_seq_init_r<NL>(T1, e, sp.init, m1);
                    else back = sgb_seq_step_r<NL>(T1, e, m1, sp.tr_self, sp.tr_off);
                    if (t > 0) out[t - 1] = (unsigned char)back;
                    if (t == T - 1) lastback = back;
                }
                continue;
            }
            double T1s[NL];
            for (int i = 0; i < NL; i++) T1s[i] = T1[i];
            const double m1s = m1;
            int back[4];
            unsigned ties = 0;
            for (int b = 0; b < 4; b++) {
                const long t = 4 * c + b;
                double e[NL];
                emit(t < T ? t : T - 1, e);
                bool tie;
                back[b] = sgb_seq_step_fast<NL>(T1, e, m1, sp.tr_self, sp.tr_off, tie);
                if (tie && t < T) ties |= 1u << b;
            }
            if (ties) {
                double m1r = m1s;
                for (int b = 0; (ties >> b) != 0; b++) {
                    double e[NL];
                    emit(4 * c + b, e);
                    if ((ties >> b) & 1u) back[b] = sgb_seq_step_r<NL>(T1s, e, m1r, sp.tr_self, sp.tr_off);
                    else { bool t_; sgb_seq_step_fast<NL>(T1s, e, m1r, sp.tr_self, sp.tr_off, t_); }
                }
            }
            for (int b = 0; b < 4; b++) {
                const long t = 4 * c + b;
                if (t < T) out[t - 1] = (unsigned char)back[b];
                if (t == T - 1) lastback = back[b];
            }
        }
        out[T - 1] = (unsigned char)lastback;
        return;
    }
    for (long t = 0; t < T; t++) {
        double e[NL];
        emit(t, e);
        int back = 0;
        if (sp.diag_dominant) {
            if (t == 0) sgb_seq_init_r<NL>(T1, e, sp.init, m1);
            else back = sgb_seq_step_r<NL>(T1, e, m1, sp.tr_self, sp.tr_off);
        } else {
            if (t == 0) sgb_seq_init<NL>(st, e, Seff, sp.init);
            else back = sgb_seq_step_general<NL>(st, e, Seff, sp.tr_self, sp.tr_off);
        }
        if (t > 0) out[t - 1] = (unsigned char)back;
    }
    out[T - 1] = (T > 1) ? out[T - 2] : 0;
}

template <int NPL>
static void run_split_s(int model, int lik_bb, int S, double loherror, double eps, double rho, double Abias, double stp, long T,
                        const int* n, const int* x, const int* cn, unsigned char* out, int block)
{
    if (model == 0 && !lik_bb) run_split<NPL, 0, 0>(S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block);
    else if (model == 0) run_split<NPL, 0, 1>(S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block);
    else if (!lik_bb) run_split<NPL, 1, 0>(S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block);
    else run_split<NPL, 1, 1>(S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block);
}

static int decode_split(int model, int lik, int S, double loherror, double eps, double rho, double Abias, double stp, long T,
                        const int* n, const int* x, const int* cn, unsigned char* out, int block)
{
    if (T <= 0) return 0;
    if (S < 1 || S > 32) return -1;
    const int lik_bb = (lik == SGB_LIK_BETABINOMIAL) && rho > 0.0;
    int mx = 0;
    for (long t = 0; t < T; t++) { int c = cn[t] < 0 ? 0 : (cn[t] > 255 ? 255 : cn[t]); if (c > mx) mx = c; }
    const int npl = sgb_planes_for(sgb_live_states(model, sgb_cprime(model, mx), S));
    switch (npl) {
    case 1: run_split_s<1>(model, lik_bb, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block); break;
    case 2: run_split_s<2>(model, lik_bb, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block); break;
    case 3: run_split_s<3>(model, lik_bb, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block); break;
    case 4: run_split_s<4>(model, lik_bb, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block); break;
    case 6: run_split_s<6>(model, lik_bb, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block); break;
    default: run_split_s<8>(model, lik_bb, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, block); break;
    }
    return 0;
}

extern "C" int emu_decode_sequence_split(int model, int lik, int S, double loherror, double eps, double rho, double Abias,
                                         double stp, long T, const int* n, const int* x, const int* cn, unsigned char* out)
{
    return decode_split(model, lik, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, 0);
}

// the four-bin block variant of the recursion (k_recur for large batches; every plane count, the kernel uses it up to two planes)
extern "C" int emu_decode_sequence_split_block(int model, int lik, int S, double loherror, double eps, double rho, double Abias,
                                               double stp, long T, const int* n, const int* x, const int* cn, unsigned char* out)
{
    return decode_split(model, lik, S, loherror, eps, rho, Abias, stp, T, n, x, cn, out, 1);
}
