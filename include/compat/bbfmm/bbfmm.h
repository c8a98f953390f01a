// aniso_b200 drop-in header: the parts of bbfmm/bbfmm.h visible to callers: bbfmm::point (bbfmm.h:54-75) and bbfmm::kernel with
// its contract (eval callback, initialize, run / runCache / runFast; bbfmm.h:451-592).  The FMM itself (tree, lists, passes,
// operator cache) lives behind libaniso_b200.so.
#ifndef ANISO_B200_COMPAT_BBFMM_H
#define ANISO_B200_COMPAT_BBFMM_H

#include <memory>

#include "aniso_b200.h"
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

// bbfmm::kernel (bbfmm.h:451-592): eval callback, initialize, run / runCache / runFast.  The tree, the lists, the transfer
// operators and the operator cache live behind libaniso_b200.so (aniso_fmm_create / aniso_fmm_cache / aniso_fmm_apply): eval is
// a host std::function, so the operators are tabulated on the host (OpenMP, once per runCache) and applied on the device.
// Sources and targets must be one point set, as in every call the reference itself makes (KernelFactory.cpp:284-286).
class kernel {
public:
    // kernel contract (bbfmm.h:457): eval(source, target).  KernelFactory::makeKernels populates it with the two aniso
    // families for host-side spot checks (the device evaluates those itself, fill_blocks_kernel); a caller may install any
    // kernel of its own and use initialize / run* below.
    std::function<scalar_t(point&, point&)> eval;
    index_t nChebyshev = 0, rank = 0, maxLevel = 0;
    Vector chargeTree;
    size_t cacheBytes = 0;  // bytes of device-resident operator blocks (the reference's Cache, bbfmm.h:466, lives on the device)
    // bbfmm.h:466.  Kept so that code which names or walks `Cache` compiles; it stays EMPTY here: the operator blocks are
    // device memory (use cacheBytes / Aniso::displayKernelCacheSize for their size).
    vector<vector<Matrix>> Cache;

    void initialize(index_t _nChebyshev, vector<point>& _source, vector<point>& _target, Vector& _charge, index_t _nSource,
                    index_t _nTarget, index_t _rank, index_t _maxLevel) {
        if (_nSource != _nTarget) throw std::invalid_argument("bbfmm::kernel (aniso_b200): sources and targets must be one point set");
        std::vector<double> x((size_t)_nSource), y((size_t)_nSource);
        for (index_t i = 0; i < _nSource; ++i) {
            if (!(_source[i] == _target[i])) throw std::invalid_argument("bbfmm::kernel (aniso_b200): sources and targets must be one point set");
            x[i] = _source[i].x;
            y[i] = _source[i].y;
        }
        // the reference rebuilds the tree at every call (KernelFactory.cpp:356, 380 re-initialise before each run); the same
        // points, order and leaf rule give the same tree, so it is kept together with its cached operators
        if (!state || state->np != _nChebyshev || state->cap != _rank || state->ml != _maxLevel || state->x != x || state->y != y) {
            state = std::make_shared<State>();
            if (aniso_fmm_create(_nChebyshev, _nSource, x.data(), y.data(), _rank, _maxLevel, &state->h) != 0)
                throw std::runtime_error(aniso_last_error());
            state->np = _nChebyshev; state->cap = _rank; state->ml = _maxLevel;
            state->x.swap(x); state->y.swap(y);
        }
        nChebyshev = _nChebyshev;
        rank = _nChebyshev * _nChebyshev;  // bbfmm.h:489
        maxLevel = _maxLevel;
        chargeTree = _charge;
    }
    void run(Vector& potentialMatrix) {  // bbfmm.h:507-532: kernel evaluated during the pass, nothing kept
        tabulate();
        apply(potentialMatrix);
        state->cached = false;
    }
    void runCache(Vector& potentialMatrix) {  // bbfmm.h:535-562
        tabulate();
        apply(potentialMatrix);
    }
    void runFast(Vector& potentialMatrix) {  // bbfmm.h:565-592
        if (!state || !state->cached) throw std::logic_error("bbfmm::kernel::runFast before runCache");
        apply(potentialMatrix);
    }

private:
    struct State {
        aniso_t* h = nullptr;
        index_t np = 0, cap = 0, ml = 0;
        std::vector<double> x, y;
        bool cached = false;
        ~State() { aniso_destroy(h); }
    };
    std::shared_ptr<State> state;  // kernel objects are copied around (vector<kernel>): copies share the device state

    static double trampoline(void* ctx, double sx, double sy, double tx, double ty) {
        point a(sx, sy), b(tx, ty);
        return (*static_cast<std::function<scalar_t(point&, point&)>*>(ctx))(a, b);
    }
    void tabulate() {
        if (!state) throw std::logic_error("bbfmm::kernel: initialize first");
        if (!eval) throw std::logic_error("bbfmm::kernel: eval is not set");
        if (aniso_fmm_cache(state->h, &kernel::trampoline, &eval, 0) != 0) throw std::runtime_error(aniso_last_error());
        state->cached = true;
        cacheBytes = aniso_cache_bytes(state->h, 0);
    }
    void apply(Vector& potentialMatrix) {
        const int n = (int)state->x.size();
        potentialMatrix = Vector(n);  // bbfmm.h:579: the output is resized / overwritten
        if (aniso_fmm_apply(state->h, chargeTree.data(), potentialMatrix.data()) != 0) throw std::runtime_error(aniso_last_error());
    }
};

}  // namespace bbfmm

#endif
