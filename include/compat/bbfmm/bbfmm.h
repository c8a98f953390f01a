// aniso_b200 drop-in header: the parts of bbfmm/bbfmm.h visible to callers of Aniso: bbfmm::point (bbfmm.h:54-75) and
// a `kernel` shell carrying the kernel contract (eval callback, Chebyshev order, rank, depth; bbfmm.h:451-505).  The FMM
// itself (tree, lists, passes, operator cache) lives on the device behind libaniso_b200.so.
#ifndef ANISO_B200_COMPAT_BBFMM_H
#define ANISO_B200_COMPAT_BBFMM_H

#include "blas_wrapper.h"
#include "linalg.h"
#include "utils.h"

namespace bbfmm {

class basePoint {
public:
    scalar_t x, y;
    basePoint() : x(0.), y(0.) {}
    basePoint(scalar_t _x, scalar_t _y) : x(_x), y(_y) {}
    virtual ~basePoint() {}
    bool operator==(const basePoint& r) const { return std::fabs(x - r.x) < EPS && std::fabs(y - r.y) < EPS; }
};

class point : public basePoint {
public:
    point() : basePoint() {}
    point(scalar_t _x, scalar_t _y) : basePoint(_x, _y) {}
    ~point() {}
};

class kernel {
public:
    // kernel contract (bbfmm.h:457): eval(source, target).  Populated by KernelFactory::makeKernels for host-side
    // spot checks; the device path evaluates the same two families in fill_blocks_kernel.
    std::function<scalar_t(point&, point&)> eval;
    index_t nChebyshev = 0, rank = 0, maxLevel = 0;
    size_t cacheBytes = 0;  // bytes of device-resident operator blocks (Cache of bbfmm.h:466)
};

}  // namespace bbfmm

#endif
