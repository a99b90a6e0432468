// Host emulation of the per-lane work of the fused decode kernel, built from the SAME
// hmm_core.cuh the CUDA kernels include (compiled here with g++, -ffp-contract=off).
// Test-only: lets the CPU test-suite check the kernel arithmetic / control flow against
// the oracle without a GPU.  Never linked into the product.
#include "../../signals_b200/csrc/hmm_core.cuh"
#include <vector>

static const double LT[SGB_LOGTAB_LEN] = SGB_LOGTAB_INIT;
static const double ET[SGB_EXPTAB_LEN] = SGB_EXPTAB_INIT;
static const double LF[SGB_LFTAB_LEN] = SGB_LFTAB_INIT;

template <int SMAX>
static void run(int model, int lik, int S, int decode, double loherror, double eps, double rho, double Abias,
                double stp, long T, const int* n, const int* x, const int* cn, unsigned char* out, double* l_out)
{
    const int lik_bb = (lik == SGB_LIK_BETABINOMIAL) && rho > 0.0;
    SgbScalars sp = sgb_scalars(S, eps, Abias, stp, rho, LT);
    SgbSeqState<SMAX> st;
    std::vector<unsigned char> bp((size_t)T * S);
    for (long t = 0; t < T; t++) {
        SgbBin b = sgb_bin_prepare(model, lik_bb, n[t], x[t], cn[t], sp.bb_s, LF, LT);
        int L = sgb_live_states(model, b.cprime, S);
        double e[SMAX];
        for (int i = 0; i < SMAX; i++) e[i] = -sgb_inf();
        for (int i = 0; i < L; i++) {
            SgbStateConst sc = sgb_state_const(b.cprime, i, loherror, lik, rho, sp.bb_s, LT);
            e[i] = sgb_emission(model, lik_bb, b, i, sc, sp, LT, ET);
        }
        if (l_out) for (int i = 0; i < S; i++) l_out[t + T * i] = e[i];
        if (t == 0) { sgb_seq_init<SMAX>(st, e, S, sp.init); continue; }
        if (decode == SGB_DECODE_REFERENCE) {
            int back = sgb_seq_step<SMAX>(st, e, S, sp.tr_self, sp.tr_off);
            out[t - 1] = (unsigned char)back;
        } else {
            sgb_seq_step_chained<SMAX>(st, e, S, sp.tr_self, sp.tr_off, &bp[(size_t)t * S]);
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

extern "C" int emu_decode_sequence(int model, int lik, int S, int decode, double loherror, double eps, double rho,
                                   double Abias, double stp, long T, const int* n, const int* x, const int* cn,
                                   unsigned char* out, double* l_out)
{
    if (T <= 0) return 0;
    if (S <= 12) run<12>(model, lik, S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else if (S <= 32) run<32>(model, lik, S, decode, loherror, eps, rho, Abias, stp, T, n, x, cn, out, l_out);
    else return -1;
    return 0;
}
