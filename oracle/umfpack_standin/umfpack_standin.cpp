// Dense-LU implementation of the UMFPACK API subset declared in ./umfpack.h.
// TEST INFRASTRUCTURE ONLY (see the header): lets the reference's InnerState be compiled and
// executed at checker sizes.  Complex arrays are packed (re, im) when Az == NULL, which is the
// only form the reference uses (umfpack_interface.h:50-51, 93-94, 141-144).
#include "umfpack.h"

#include <chrono>
#include <cmath>
#include <complex>
#include <cstdio>
#include <vector>

namespace {

struct Sym {
    int64_t n;
};

template <class T>
struct Num {
    int64_t n;
    std::vector<T> lu;        // row-major n x n, L (unit) below and U on/above the diagonal
    std::vector<int64_t> piv; // row exchanged with row k at step k
};

template <class T>
double mag(const T& v) { return std::abs(v); }

template <class T, class I, class Get>
int factor(const I* Ap, const I* Ai, Get get, const Sym* s, void** out) {
    if (!Ap || !Ai || !s || !out) return UMFPACK_ERROR_argument_missing;
    const int64_t n = s->n;
    auto* N = new Num<T>{n, std::vector<T>((size_t)n * n, T(0)), std::vector<int64_t>(n)};
    for (int64_t j = 0; j < n; ++j)
        for (I p = Ap[j]; p < Ap[j + 1]; ++p) {
            if (Ai[p] < 0 || Ai[p] >= n) { delete N; return UMFPACK_ERROR_invalid_matrix; }
            N->lu[(size_t)Ai[p] * n + j] += get(p);
        }
    int status = UMFPACK_OK;
    T* a = N->lu.data();
    for (int64_t k = 0; k < n; ++k) {
        int64_t r = k;
        double best = mag(a[k * n + k]);
        for (int64_t i = k + 1; i < n; ++i)
            if (mag(a[i * n + k]) > best) { best = mag(a[i * n + k]); r = i; }
        N->piv[k] = r;
        if (!(best > 0.0) || !std::isfinite(best)) { status = UMFPACK_WARNING_singular_matrix; continue; }
        if (r != k)
            for (int64_t j = 0; j < n; ++j) std::swap(a[k * n + j], a[r * n + j]);
        const T d = T(1) / a[k * n + k];
        for (int64_t i = k + 1; i < n; ++i) {
            const T l = a[i * n + k] * d;
            if (l == T(0)) continue;
            a[i * n + k] = l;
            for (int64_t j = k + 1; j < n; ++j) a[i * n + j] -= l * a[k * n + j];
        }
    }
    *out = N;
    return status;
}

// P A = L U.  sys = UMFPACK_A: x = U^-1 L^-1 P b;  UMFPACK_Aat: A^T = U^T L^T P  =>
// x = P^T L^-T U^-T b;  UMFPACK_At: the same with conjugated factors.
template <class T, class Conj>
int solve_with(int sys, const Num<T>* N, T* x, const T* b, Conj cj) {
    if (!N || !x || !b) return UMFPACK_ERROR_argument_missing;
    const int64_t n = N->n;
    const T* a = N->lu.data();
    std::vector<T> y(b, b + n);
    if (sys == UMFPACK_A) {
        for (int64_t k = 0; k < n; ++k) std::swap(y[k], y[N->piv[k]]);
        for (int64_t i = 0; i < n; ++i) {
            T s = y[i];
            for (int64_t j = 0; j < i; ++j) s -= a[i * n + j] * y[j];
            y[i] = s;
        }
        for (int64_t i = n - 1; i >= 0; --i) {
            T s = y[i];
            for (int64_t j = i + 1; j < n; ++j) s -= a[i * n + j] * y[j];
            y[i] = s / a[i * n + i];
        }
    } else if (sys == UMFPACK_Aat || sys == UMFPACK_At) {
        const bool c = sys == UMFPACK_At;
        auto el = [&](int64_t i, int64_t j) { return c ? cj(a[i * n + j]) : a[i * n + j]; };
        for (int64_t i = 0; i < n; ++i) {          // U^T y = b
            T s = y[i];
            for (int64_t j = 0; j < i; ++j) s -= el(j, i) * y[j];
            y[i] = s / el(i, i);
        }
        for (int64_t i = n - 1; i >= 0; --i) {     // L^T z = y
            T s = y[i];
            for (int64_t j = i + 1; j < n; ++j) s -= el(j, i) * y[j];
            y[i] = s;
        }
        for (int64_t k = n - 1; k >= 0; --k) std::swap(y[k], y[N->piv[k]]);
    } else {
        return UMFPACK_ERROR_argument_missing;
    }
    for (int64_t i = 0; i < n; ++i) x[i] = y[i];
    return UMFPACK_OK;
}

template <class I>
int symbolic(I n_row, I n_col, const I* Ap, const I* Ai, void** out) {
    if (!Ap || !Ai || !out) return UMFPACK_ERROR_argument_missing;
    if (n_row != n_col || n_row <= 0) return UMFPACK_ERROR_invalid_matrix;
    *out = new Sym{(int64_t)n_row};
    return UMFPACK_OK;
}

using cplx = std::complex<double>;

}  // namespace

extern "C" {

double umfpack_timer(void) {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

#define JPB_IMPL(P, I)                                                                              \
    int umfpack_d##P##_symbolic(I n_row, I n_col, const I Ap[], const I Ai[], const double*,        \
                                void** S, const double*, double*) {                                 \
        return symbolic<I>(n_row, n_col, Ap, Ai, S);                                                \
    }                                                                                               \
    int umfpack_d##P##_numeric(const I Ap[], const I Ai[], const double Ax[], void* S, void** Nn,   \
                               const double*, double*) {                                            \
        if (!Ax) return UMFPACK_ERROR_argument_missing;                                             \
        return factor<double, I>(Ap, Ai, [&](I p) { return Ax[p]; }, (const Sym*)S, Nn);            \
    }                                                                                               \
    int umfpack_d##P##_solve(int sys, const I*, const I*, const double*, double X[],                \
                             const double B[], void* Nn, const double*, double*) {                  \
        return solve_with<double>(sys, (const Num<double>*)Nn, X, B, [](double v) { return v; });   \
    }                                                                                               \
    void umfpack_d##P##_free_symbolic(void** S) { if (S && *S) { delete (Sym*)*S; *S = nullptr; } } \
    void umfpack_d##P##_free_numeric(void** Nn) {                                                   \
        if (Nn && *Nn) { delete (Num<double>*)*Nn; *Nn = nullptr; }                                 \
    }                                                                                               \
    void umfpack_d##P##_report_symbolic(void* S, const double*) {                                   \
        if (S) printf("umfpack stand-in: dense LU, n = %lld\n", (long long)((Sym*)S)->n);           \
    }                                                                                               \
    int umfpack_z##P##_symbolic(I n_row, I n_col, const I Ap[], const I Ai[], const double*,        \
                                const double*, void** S, const double*, double*) {                  \
        return symbolic<I>(n_row, n_col, Ap, Ai, S);                                                \
    }                                                                                               \
    int umfpack_z##P##_numeric(const I Ap[], const I Ai[], const double Ax[], const double Az[],    \
                               void* S, void** Nn, const double*, double*) {                        \
        if (!Ax) return UMFPACK_ERROR_argument_missing;                                             \
        return factor<cplx, I>(Ap, Ai, [&](I p) {                                                   \
            return Az ? cplx(Ax[p], Az[p]) : cplx(Ax[2 * p], Ax[2 * p + 1]); }, (const Sym*)S, Nn); \
    }                                                                                               \
    int umfpack_z##P##_solve(int sys, const I*, const I*, const double*, const double*,             \
                             double Xx[], double Xz[], const double Bx[], const double Bz[],        \
                             void* Nn, const double*, double*) {                                    \
        const Num<cplx>* N = (const Num<cplx>*)Nn;                                                  \
        if (!N || !Xx || !Bx) return UMFPACK_ERROR_argument_missing;                                \
        std::vector<cplx> b(N->n), x(N->n);                                                         \
        for (int64_t i = 0; i < N->n; ++i)                                                          \
            b[i] = Bz ? cplx(Bx[i], Bz[i]) : cplx(Bx[2 * i], Bx[2 * i + 1]);                        \
        int s = solve_with<cplx>(sys, N, x.data(), b.data(), [](cplx v) { return std::conj(v); });  \
        for (int64_t i = 0; i < N->n; ++i) {                                                        \
            if (Xz) { Xx[i] = x[i].real(); Xz[i] = x[i].imag(); }                                   \
            else { Xx[2 * i] = x[i].real(); Xx[2 * i + 1] = x[i].imag(); }                          \
        }                                                                                           \
        return s;                                                                                   \
    }                                                                                               \
    void umfpack_z##P##_free_symbolic(void** S) { if (S && *S) { delete (Sym*)*S; *S = nullptr; } } \
    void umfpack_z##P##_free_numeric(void** Nn) {                                                   \
        if (Nn && *Nn) { delete (Num<cplx>*)*Nn; *Nn = nullptr; }                                   \
    }                                                                                               \
    void umfpack_z##P##_report_symbolic(void* S, const double*) {                                   \
        if (S) printf("umfpack stand-in: dense LU, n = %lld\n", (long long)((Sym*)S)->n);           \
    }

JPB_IMPL(i, int32_t)
JPB_IMPL(l, int64_t)
#undef JPB_IMPL

}  // extern "C"
