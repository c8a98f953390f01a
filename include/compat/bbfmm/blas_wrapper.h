// aniso_b200 drop-in header: the BLAS-1 wrappers callers use on host Vectors (reference bbfmm/blas_wrapper.h; the
// level-2/3 wrappers are internal to the reference's CPU FMM and have no role here: those contractions run on the GPU).
#ifndef ANISO_B200_COMPAT_BLAS_WRAPPER_H
#define ANISO_B200_COMPAT_BLAS_WRAPPER_H

#include "linalg.h"

namespace bbfmm {

inline void dscal(scalar_t alpha, Vector& v) {
    for (index_t i = 0; i < v.row(); ++i) v(i) *= alpha;
}
inline void daxpy(scalar_t a, Vector& x, Vector& y) {  // y += a x
    assert(x.row() == y.row());
    for (index_t i = 0; i < x.row(); ++i) y(i) += a * x(i);
}
inline scalar_t ddot(Vector& x, Vector& y) {
    scalar_t s = 0;
    for (index_t i = 0; i < x.row(); ++i) s += x(i) * y(i);
    return s;
}
inline scalar_t nrm2(Vector& x) { return std::sqrt(ddot(x, x)); }

}  // namespace bbfmm

#endif
