// ref_driver.cpp -- plain-C entry points around the reference's own C++ routines (src/code.cpp:5-45), which
// oracle/Makefile compiles verbatim from /root/reference against oracle/rcpp_mock/Rcpp.h into oracle/_ref/.
// Test infrastructure: only tests/ load the result, to pin oracle.c's viterbi / logspace_add against it.
#include <Rcpp.h>

double logspace_addcpp(double logx, double logy);
Rcpp::NumericVector viterbi(Rcpp::NumericMatrix emission, Rcpp::NumericMatrix transition, Rcpp::NumericVector observations);

extern "C" {

double ref_logspace_addcpp(double a, double b) { return logspace_addcpp(a, b); }

// emission: nrow_e x S column-major (an R matrix), transition: S x S column-major, observations: T 1-based doubles
int ref_viterbi(const double* emission, int nrow_e, int S, const double* transition, const double* observations, int T,
                double* path_out)
{
    if (T < 1 || S < 1 || nrow_e < 1) return 1;
    Rcpp::NumericVector r = viterbi(Rcpp::NumericMatrix(nrow_e, S, emission), Rcpp::NumericMatrix(S, S, transition),
                                    Rcpp::NumericVector(observations, T));
    if (r.length() != T) return 2;
    for (int i = 0; i < T; i++) path_out[i] = r[i];
    return 0;
}

}
