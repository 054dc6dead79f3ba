/* Stand-in for <umfpack.h> (SuiteSparse UMFPACK v5.x) -- TEST INFRASTRUCTURE ONLY.
 *
 * UMFPACK is not vendored by the reference and is absent from this image, so the reference's
 * native module (/root/reference/source/jax_plate_lib/src/main.cpp, include/InnerState.h,
 * include/umfpack_interface.h) cannot be built as shipped.  This header declares exactly the
 * part of UMFPACK's published C API that umfpack_interface.h:9-197 calls, with UMFPACK's
 * published constant values; umfpack_standin.cpp implements it with a DENSE LU with partial
 * pivoting (O(n^3): checker sizes only).  With it the reference's own InnerState class --
 * add_mat, the 5-mode shape tables of solve / matvec, the OpenMP loops, csc_matvec.h -- is
 * compiled from where it lies into oracle/_ref/ and used by tests/ as the checker.
 * It is NOT UMFPACK: fill-reducing ordering, threshold pivoting, scaling and iterative
 * refinement of the real library are not reproduced (the answers agree to rounding). */
#ifndef JPB_UMFPACK_STANDIN_H
#define JPB_UMFPACK_STANDIN_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define UMFPACK_OK 0
#define UMFPACK_WARNING_singular_matrix 1
#define UMFPACK_ERROR_invalid_matrix (-8)
#define UMFPACK_ERROR_argument_missing (-5)
#define UMFPACK_CONTROL 20
#define UMFPACK_INFO 90
#define UMFPACK_A 0    /* A x = b */
#define UMFPACK_At 1   /* A^H x = b */
#define UMFPACK_Aat 2  /* A^T x = b (array transpose, no conjugation) */

typedef int64_t SuiteSparse_long;

double umfpack_timer(void);

#define JPB_DECL(P, I)                                                                             \
    int umfpack_d##P##_symbolic(I n_row, I n_col, const I Ap[], const I Ai[], const double Ax[],   \
                                void** Symbolic, const double Control[], double Info[]);          \
    int umfpack_d##P##_numeric(const I Ap[], const I Ai[], const double Ax[], void* Symbolic,      \
                               void** Numeric, const double Control[], double Info[]);            \
    int umfpack_d##P##_solve(int sys, const I Ap[], const I Ai[], const double Ax[], double X[],   \
                             const double B[], void* Numeric, const double Control[],             \
                             double Info[]);                                                      \
    void umfpack_d##P##_free_symbolic(void** Symbolic);                                           \
    void umfpack_d##P##_free_numeric(void** Numeric);                                             \
    void umfpack_d##P##_report_symbolic(void* Symbolic, const double Control[]);                  \
    int umfpack_z##P##_symbolic(I n_row, I n_col, const I Ap[], const I Ai[], const double Ax[],   \
                                const double Az[], void** Symbolic, const double Control[],       \
                                double Info[]);                                                   \
    int umfpack_z##P##_numeric(const I Ap[], const I Ai[], const double Ax[], const double Az[],   \
                               void* Symbolic, void** Numeric, const double Control[],            \
                               double Info[]);                                                    \
    int umfpack_z##P##_solve(int sys, const I Ap[], const I Ai[], const double Ax[],               \
                             const double Az[], double Xx[], double Xz[], const double Bx[],      \
                             const double Bz[], void* Numeric, const double Control[],            \
                             double Info[]);                                                      \
    void umfpack_z##P##_free_symbolic(void** Symbolic);                                           \
    void umfpack_z##P##_free_numeric(void** Numeric);                                             \
    void umfpack_z##P##_report_symbolic(void* Symbolic, const double Control[]);

JPB_DECL(i, int32_t)
JPB_DECL(l, int64_t)
#undef JPB_DECL

#ifdef __cplusplus
}
#endif
#endif
