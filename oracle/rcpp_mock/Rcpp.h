// Rcpp.h -- MOCK of the handful of Rcpp classes and sugar functions that the reference's src/code.cpp uses.
//
// TEST INFRASTRUCTURE ONLY.  Rcpp is not installed where this repository is built, so the reference's own
// viterbi / logspace_addcpp cannot be compiled against the real library.  With this header they compile
// VERBATIM from /root/reference/src/code.cpp (oracle/Makefile, target _ref): the loop structure, the order of
// the floating-point additions and the decode rule in oracle/_ref are then the reference's own text, and only
// the container / sugar semantics below are restated here (from Rcpp's documented behaviour, not its sources):
//   * NumericVector / NumericMatrix: double storage, matrices column-major, `m(_, j)` is column j;
//   * `a + b`, `a + scalar`, `a - scalar`: element-wise, evaluated left to right, so `x + y + s` is (x[i]+y[i])+s;
//   * max(v): largest element;  which_max(v): 0-based index of the FIRST largest element (strict '>').
// Values are eager (Rcpp's are lazy expression templates); element for element the arithmetic is the same.
#ifndef SGB_RCPP_MOCK_H
#define SGB_RCPP_MOCK_H
#include <math.h>
#include <cstddef>
#include <vector>

namespace Rcpp {

struct Placeholder {};
static const Placeholder _ = Placeholder();

class NumericVector {
public:
    NumericVector() {}
    explicit NumericVector(int n) : v_((size_t)n, 0.0) {}
    NumericVector(int n, double fill) : v_((size_t)n, fill) {}
    NumericVector(const double* p, int n) : v_(p, p + n) {}
    int length() const { return (int)v_.size(); }
    int size() const { return (int)v_.size(); }
    double& operator()(int i) { return v_[(size_t)i]; }
    double operator()(int i) const { return v_[(size_t)i]; }
    double& operator[](int i) { return v_[(size_t)i]; }
    double operator[](int i) const { return v_[(size_t)i]; }
private:
    std::vector<double> v_;
};

class NumericMatrix {
public:
    class Column {
    public:
        Column(NumericMatrix* m, int j) : m_(m), j_(j) {}
        Column& operator=(const NumericVector& v)
        {
            for (int i = 0; i < m_->nrow(); i++) (*m_)(i, j_) = v[i];
            return *this;
        }
        operator NumericVector() const
        {
            NumericVector out(m_->nrow());
            for (int i = 0; i < m_->nrow(); i++) out[i] = (*m_)(i, j_);
            return out;
        }
    private:
        NumericMatrix* m_;
        int j_;
    };
    NumericMatrix() : nr_(0), nc_(0) {}
    NumericMatrix(int nrow, int ncol) : nr_(nrow), nc_(ncol), v_((size_t)nrow * (size_t)ncol, 0.0) {}
    NumericMatrix(int nrow, int ncol, const double* colmajor) : nr_(nrow), nc_(ncol), v_(colmajor, colmajor + (size_t)nrow * (size_t)ncol) {}
    int nrow() const { return nr_; }
    int ncol() const { return nc_; }
    double& operator()(int i, int j) { return v_[(size_t)i + (size_t)nr_ * (size_t)j]; }
    double operator()(int i, int j) const { return v_[(size_t)i + (size_t)nr_ * (size_t)j]; }
    Column operator()(Placeholder, int j) { return Column(this, j); }
private:
    int nr_, nc_;
    std::vector<double> v_;
};

inline NumericMatrix transpose(const NumericMatrix& m)
{
    NumericMatrix t(m.ncol(), m.nrow());
    for (int j = 0; j < m.ncol(); j++)
        for (int i = 0; i < m.nrow(); i++) t(j, i) = m(i, j);
    return t;
}

inline NumericVector operator+(const NumericVector& a, const NumericVector& b)
{
    NumericVector out(a.length());
    for (int i = 0; i < a.length(); i++) out[i] = a[i] + b[i];
    return out;
}
inline NumericVector operator+(const NumericVector& a, double s)
{
    NumericVector out(a.length());
    for (int i = 0; i < a.length(); i++) out[i] = a[i] + s;
    return out;
}
inline NumericVector operator-(const NumericVector& a, double s)
{
    NumericVector out(a.length());
    for (int i = 0; i < a.length(); i++) out[i] = a[i] - s;
    return out;
}

// largest element; a NaN element is returned as soon as it is met (Rcpp propagates NA)
inline double max(const NumericVector& v)
{
    double best = v[0];
    if (best != best) return best;
    for (int i = 1; i < v.length(); i++) {
        const double cur = v[i];
        if (cur != cur) return cur;
        if (cur > best) best = cur;
    }
    return best;
}
// 0-based index of the first largest element
inline int which_max(const NumericVector& v)
{
    double best = v[0];
    int idx = 0;
    for (int i = 1; i < v.length(); i++) {
        const double cur = v[i];
        if (cur > best) { best = cur; idx = i; }
    }
    return idx;
}

}  // namespace Rcpp
#endif
