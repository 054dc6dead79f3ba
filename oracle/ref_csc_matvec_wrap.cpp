// TEST INFRASTRUCTURE ONLY.  extern "C" instantiations of the reference's own templates
// csc_matvec / csc_matvec_transpose (source/jax_plate_lib/include/csc_matvec.h:31-66),
// #included from where the header lies under /root/reference (never copied into this
// repository).  Built by oracle/Makefile into oracle/_ref/libref_csc_matvec.so and used by
// tests/ to validate oracle.jax_plate_oracle.csc_matvec{,_transpose} and the CUDA SpMV.
#include <complex>
#include <cstdint>
#include "csc_matvec.h"

typedef std::complex<double> zd;

extern "C" {
void ref_csc_matvec_z32(int32_t n_col, const int32_t* Ap, const int32_t* Ai, const zd* Ax,
                        const zd* Xx, zd* Yx) {
    csc_matvec<int32_t, zd>(n_col, Ap, Ai, Ax, Xx, Yx);
}
void ref_csc_matvec_transpose_z32(int32_t n_col, const int32_t* Ap, const int32_t* Ai, const zd* Ax,
                                  const zd* Xx, const int32_t* perm, zd* Yx) {
    csc_matvec_transpose<int32_t, zd>(n_col, Ap, Ai, Ax, Xx, perm, Yx);
}
void ref_csc_matvec_d64(int64_t n_col, const int64_t* Ap, const int64_t* Ai, const double* Ax,
                        const double* Xx, double* Yx) {
    csc_matvec<int64_t, double>(n_col, Ap, Ai, Ax, Xx, Yx);
}
}
