// TEST INFRASTRUCTURE ONLY (oracle/): C-ABI around the UNMODIFIED reference classes.
//
// This translation unit is compiled together with the reference's own sources where they
// lie under /root/reference (Aniso.cpp Geometry.cpp KernelFactory.cpp Quadrature.cpp
// bbfmm/blas_wrapper.cpp gmres.cpp) into oracle/_ref/libaniso_ref*.so by oracle/Makefile.
// Nothing of the reference is copied here: every entry point below only *calls* the
// reference's public API in the order its own drivers do
//   - MEX gateway      /root/reference/AnisoWrapper.cpp:10-136  (new/getNodes/setCoeff/cache/mapping)
//   - standalone main  /root/reference/main.cpp:48-141          (cache_mapping/apply_mapping/GMRES)
// The product (aniso_b200/) never links or loads this library; only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs do.
#include <sstream>
#include <string>
#include <vector>
#include <cstdio>
#include <cstring>
#include "Aniso.h"

// gmres.h cannot be included after <iostream> on libstdc++ 13 (its `inline scalar_t abs(scalar_t)`
// clashes with ::abs(double)); forward-declare the entry point instead (gmres.h:23).
int GMRES(const std::function<Vector(Vector &)> A, Vector &x, Vector &b, int m, int max_iter, scalar_t tol);

namespace {
struct RefHandle {
    Aniso *a;
    long applies;
};

// apply_mapping of main.cpp:78-119 == 'mapping' of AnisoWrapper.cpp:92-136, with the
// intermediate vectors exposed for stage-wise parity.
void mapping_stages(Aniso *as, const double *charge, int id, double *out_s_, double *real_fmm, double *k7,
                    double *k8, double *k9, double *out) {
    int n = as->numberOfNodes;
    Vector scaledFunction(n), unscaledFunction(n), output_t, output_s;
    std::vector<Vector> unscaledCoefficient;
    for (int i = 0; i < n; ++i) {
        unscaledFunction(i) = charge[i];
        scaledFunction(i) = charge[i] * as->weights[i];
    }
    as->interpolation(unscaledFunction, unscaledCoefficient);
    as->runKernelsFast(id, scaledFunction, output_s);
    if (out_s_) memcpy(out_s_, output_s.data(), sizeof(double) * n);
    as->runKernelsFastSing(id, scaledFunction, output_t);
    if (real_fmm) memcpy(real_fmm, output_t.data(), sizeof(double) * n);
    as->nearRemoval(id, scaledFunction, output_t);
    if (k7) memcpy(k7, output_t.data(), sizeof(double) * n);
    as->refineAddOnFast(id, scaledFunction, output_t);
    if (k8) memcpy(k8, output_t.data(), sizeof(double) * n);
    as->singularAddFast(id, unscaledCoefficient, output_t);
    if (k9) memcpy(k9, output_t.data(), sizeof(double) * n);
    daxpy(1.0, output_s, output_t);
    dscal(M_1_PI / 2.0, output_t);
    if (out) memcpy(out, output_t.data(), sizeof(double) * n);
}
}  // namespace

extern "C" {

void *ref_create(int sz, int deg, int K, double g, int sr, int np, int maxLevel) {
#ifdef RUN_OMP
    omp_set_num_threads(omp_get_max_threads());
#endif
    RefHandle *h = new RefHandle;
    h->a = new Aniso(sz, deg, K, g, sr, np, maxLevel);
    h->applies = 0;
    return h;
}

void ref_destroy(void *p) {
    RefHandle *h = (RefHandle *)p;
    delete h->a;
    delete h;
}

int ref_num_nodes(void *p) { return ((RefHandle *)p)->a->numberOfNodes; }
int ref_kernel_size(void *p) { return ((RefHandle *)p)->a->kernelSize; }
int ref_num_threads(void) {
#ifdef RUN_OMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ref_get_nodes(void *p, double *x, double *y, double *w) {
    Aniso *as = ((RefHandle *)p)->a;
    for (int i = 0; i < as->numberOfNodes; ++i) {
        if (x) x[i] = as->nodes[i].x;
        if (y) y[i] = as->nodes[i].y;
        if (w) w[i] = as->weights[i];
    }
}

// 'setCoeff' of AnisoWrapper.cpp:46-69
void ref_set_coeff(void *p, const double *sigma_s, const double *sigma_t) {
    Aniso *as = ((RefHandle *)p)->a;
    for (int i = 0; i < as->numberOfNodes; ++i) {
        as->sigma_t(i) = sigma_t[i];
        as->sigma_s(i) = sigma_s[i];
    }
    as->interpolation();
    as->singPrecompute();
    as->makeKernels();
}

// 'cache' of AnisoWrapper.cpp:72-90; times[4] (seconds) optional
void ref_cache(void *p, int id, double *times) {
    Aniso *as = ((RefHandle *)p)->a;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto sec = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration_cast<std::chrono::microseconds>(b - a).count() / 1e6;
    };
    auto t0 = now();
    as->runKernelsCache(id);
    auto t1 = now();
    as->runKernelsCacheSing(id);
    auto t2 = now();
    as->refineAddOnCache(id);
    auto t3 = now();
    as->singularAddCache(id);
    auto t4 = now();
    if (times) {
        times[0] = sec(t0, t1);
        times[1] = sec(t1, t2);
        times[2] = sec(t2, t3);
        times[3] = sec(t3, t4);
    }
}

void ref_mapping(void *p, const double *charge, int id, double *out) {
    RefHandle *h = (RefHandle *)p;
    mapping_stages(h->a, charge, id, nullptr, nullptr, nullptr, nullptr, nullptr, out);
    h->applies++;
}

void ref_mapping_stages(void *p, const double *charge, int id, double *out_s, double *real_fmm, double *k7,
                        double *k8, double *k9, double *out) {
    RefHandle *h = (RefHandle *)p;
    mapping_stages(h->a, charge, id, out_s, real_fmm, k7, k8, k9, out);
    h->applies++;
}

// main.cpp:121-141: A x = x - mapping_0(sigma_s .* x), solved with the reference's own GMRES.
// x holds x0 on entry and the solution on exit.  The reference prints its residual table to
// stdout (gmres.cpp:108-114); it is captured here and parsed into hist[] (the value printed
// at iteration j is the residual after j-1 steps).  Returns GMRES' own return code.
int ref_gmres(void *p, const double *b, double *x, int m, int max_iter, double tol, int *iters, double *hist,
              int hist_cap) {
    RefHandle *h = (RefHandle *)p;
    Aniso *as = h->a;
    int n = as->numberOfNodes;
    Vector rhs(n), xv(n);
    for (int i = 0; i < n; ++i) {
        rhs(i) = b[i];
        xv(i) = x[i];
    }
    auto forwardOperator = [&](Vector &scaledCharge) {
        Vector in = scaledCharge;
        std::vector<double> load(n), scatter(n);
        for (int i = 0; i < n; ++i) load[i] = in(i) * as->sigma_s(i);
        mapping_stages(as, load.data(), 0, nullptr, nullptr, nullptr, nullptr, nullptr, scatter.data());
        h->applies++;
        for (int i = 0; i < n; ++i) in(i) -= scatter[i];
        return in;
    };
    std::ostringstream cap;
    std::streambuf *old = std::cout.rdbuf(cap.rdbuf());
    int rc = GMRES(forwardOperator, xv, rhs, m, max_iter, tol);
    std::cout.rdbuf(old);
    for (int i = 0; i < n; ++i) x[i] = xv(i);
    // parse "<iter> <resid> <seconds>" rows
    std::istringstream in(cap.str());
    std::string line;
    int last = 0;
    while (std::getline(in, line)) {
        int it;
        double r, t;
        if (sscanf(line.c_str(), "%d %lf %lf", &it, &r, &t) == 3) {
            if (hist && it - 1 < hist_cap && it >= 1) hist[it - 1] = r;
            last = it;
        }
    }
    if (iters) *iters = last;
    return rc;
}

long ref_applies(void *p) { return ((RefHandle *)p)->applies; }

double ref_line_integral(void *p, double x0, double y0, double x1, double y1) {
    return ((RefHandle *)p)->a->lineIntegral(x0, y0, x1, y1);
}

double ref_evaluate(void *p, double x, double y) {
    point q(x, y);
    return ((RefHandle *)p)->a->evaluate(q);
}

// kernel family of KernelFactory.cpp:240-267; part 0 = realParts ("Sing"), 1 = imagParts. a = source, b = target.
double ref_kernel_eval(void *p, int id, int part, double ax, double ay, double bx, double by) {
    Aniso *as = ((RefHandle *)p)->a;
    point a(ax, ay), b(bx, by);
    return part ? as->imagParts[id].eval(a, b) : as->realParts[id].eval(a, b);
}

// tree statistics of the kernel object last initialised for `id` (imag part): SURVEY 8 table.
// out[0..8] = nodes, leaves, maxLevel, sumU, sumV, sumW, sumX, p2p pairs, max points per leaf
void ref_tree_stats(void *p, int id, long *out) {
    Aniso *as = ((RefHandle *)p)->a;
    bbfmm::tree &t = as->imagParts[id].t;
    long nodes = (long)t.dict.size(), leaves = 0, su = 0, sv = 0, sw = 0, sx = 0, pairs = 0, mx = 0;
    for (auto &n : t.dict) {
        if (n.isLeaf) {
            leaves++;
            if (n.nSource > mx) mx = n.nSource;
            for (int u : n.uList) pairs += (long)t.dict[u].nSource * n.nTarget;
            for (int u : n.wList) pairs += (long)t.dict[u].nSource * n.nTarget;
        }
        su += n.nUList;
        sv += n.nVList;
        sw += n.nWList;
        sx += n.nXList;
    }
    out[0] = nodes; out[1] = leaves; out[2] = t.maxLevel; out[3] = su; out[4] = sv;
    out[5] = sw; out[6] = sx; out[7] = pairs; out[8] = mx;
}

// dump the tree of kernel object `id`: per node center/radius (4 doubles), level, isLeaf, nSource, parent.
int ref_tree_nodes(void *p, int id, int cap, double *geom, int *info) {
    Aniso *as = ((RefHandle *)p)->a;
    bbfmm::tree &t = as->imagParts[id].t;
    int n = (int)t.dict.size();
    for (int i = 0; i < n && i < cap; ++i) {
        geom[4 * i + 0] = t.dict[i].center.x; geom[4 * i + 1] = t.dict[i].center.y;
        geom[4 * i + 2] = t.dict[i].radius.x; geom[4 * i + 3] = t.dict[i].radius.y;
        info[4 * i + 0] = t.dict[i].nLevel; info[4 * i + 1] = t.dict[i].isLeaf;
        info[4 * i + 2] = t.dict[i].nSource; info[4 * i + 3] = t.dict[i].parent;
    }
    return n;
}

// raw access to one cached operator block (Cache[node][slot], bbfmm.h:466) for tile-level parity.
// returns rows*cols (0 if absent); data copied column-major into out (cap doubles).
int ref_cache_block(void *p, int id, int part, int node, int slot, int *rows, int *cols, double *out, int cap) {
    Aniso *as = ((RefHandle *)p)->a;
    bbfmm::kernel &k = part ? as->imagParts[id] : as->realParts[id];
    if (node < 0 || node >= (int)k.Cache.size()) return 0;
    if (slot < 0 || slot >= (int)k.Cache[node].size()) return 0;
    Matrix &M = k.Cache[node][slot];
    *rows = M.row(); *cols = M.col();
    int n = M.row() * M.col();
    for (int i = 0; i < n && i < cap; ++i) out[i] = M.data()[i];
    return n;
}

// list of node `node` in the *iteration order the reference uses* (unordered_set order, Q10):
// V then X then U then W, matching the Cache slot order of bbfmm.h:959-1011. kinds[i] in {0:V,1:X,2:U,3:W}.
int ref_node_lists(void *p, int id, int node, int cap, int *ids, int *kinds) {
    Aniso *as = ((RefHandle *)p)->a;
    bbfmm::tree &t = as->imagParts[id].t;
    bbfmm::node &n = t.dict[node];
    int c = 0;
    if (n.parent != -1) {
        for (int i : n.vList) { if (c < cap) { ids[c] = i; kinds[c] = 0; } c++; }
        for (int i : n.xList) { if (c < cap) { ids[c] = i; kinds[c] = 1; } c++; }
    }
    if (n.isLeaf && n.nTarget != 0) {
        for (int i : n.uList) { if (c < cap) { ids[c] = i; kinds[c] = 2; } c++; }
        for (int i : n.wList) { if (c < cap) { ids[c] = i; kinds[c] = 3; } c++; }
    }
    return c;
}

int ref_node_points(void *p, int id, int node, int cap, int *idx) {
    Aniso *as = ((RefHandle *)p)->a;
    bbfmm::node &n = as->imagParts[id].t.dict[node];
    for (int i = 0; i < n.nSource && i < cap; ++i) idx[i] = n.sourceIndex[i];
    return n.nSource;
}

void ref_gl_table(int deg, double *x, double *w) {
    Quadrature q;
    get_legendre_data((size_t)deg, q);
    for (size_t i = 0; i < q.points_x.size(); ++i) {
        x[i] = q.points_x[i];
        w[i] = q.weights[i];
    }
}

// small geometry tables for host-side parity: interpolate (deg^2 x deg^2, col-major), nearMapping
// (16deg^2 x deg^2, col-major), legendreNorms (deg^2), refine_quad_{x,y}, refine_weight (16deg^2).
void ref_geometry_tables(void *p, double *interpolate, double *nearMapping, double *legendreNorms, double *rqx,
                         double *rqy, double *rw) {
    Aniso *as = ((RefHandle *)p)->a;
    int c = as->coarseQuadratureSize, r = as->refineQuadratureSize;
    if (interpolate) memcpy(interpolate, as->interpolate.data(), sizeof(double) * c * c);
    if (nearMapping) memcpy(nearMapping, as->nearMapping.data(), sizeof(double) * r * c);
    if (legendreNorms) memcpy(legendreNorms, as->legendreNorms.data(), sizeof(double) * c);
    for (int i = 0; i < r; ++i) {
        if (rqx) rqx[i] = as->refine_quad_x[i];
        if (rqy) rqy[i] = as->refine_quad_y[i];
        if (rw) rw[i] = as->refine_weight[i];
    }
}

// Duffy tables of singPrecompute (KernelFactory.cpp:929-986): singX/Y/W[deg^2][8 sr^2]
void ref_sing_tables(void *p, double *sx, double *sy, double *sw) {
    Aniso *as = ((RefHandle *)p)->a;
    size_t c = as->singX.size();
    for (size_t t = 0; t < c; ++t) {
        size_t m = as->singX[t].size();
        for (size_t q = 0; q < m; ++q) {
            sx[t * m + q] = as->singX[t][q];
            sy[t * m + q] = as->singY[t][q];
            sw[t * m + q] = as->singW[t][q];
        }
    }
}

// sigma_t Legendre coefficients per cell (KernelFactory.cpp:212-227): [sz^2][deg^2]
void ref_sigma_coeff(void *p, double *ct, double *cs) {
    Aniso *as = ((RefHandle *)p)->a;
    int c = as->coarseQuadratureSize;
    for (int i = 0; i < as->numberOfSquares; ++i)
        for (int j = 0; j < c; ++j) {
            if (ct) ct[i * c + j] = as->sigma_t_Coeff[i](j);
            if (cs) cs[i * c + j] = as->sigma_s_Coeff[i](j);
        }
}

// The black-box engine on its own (bbfmm.h:451-592) with kernels that are NOT the two families of makeKernels, on an
// arbitrary point set: eval kind 0 = log(r) (0 on the diagonal), kind 1 = an unsymmetric smooth kernel
// (sx - tx + 0.3) / (r^2 + 0.01), kind 2 = exp(-3 r) / (0.05 + r).  mode 0 = run (uncached), 1 = runCache, 2 = runCache
// then runFast.  stats[9] as ref_tree_stats.  Returns the number of tree nodes.
static double ref_bbfmm_eval(int kind, double sx, double sy, double tx, double ty) {
    const double dx = sx - tx, dy = sy - ty, r2 = dx * dx + dy * dy;
    if (kind == 0) return r2 == 0.0 ? 0.0 : 0.5 * std::log(r2);
    if (kind == 1) return (dx + 0.3) / (r2 + 0.01);
    const double r = std::sqrt(r2);
    return std::exp(-3.0 * r) / (0.05 + r);
}

int ref_bbfmm_run(int np, int n, const double *x, const double *y, int rank, int maxLevel, int kind, int mode,
                  const double *charge, double *potential, long *stats) {
#ifdef RUN_OMP
    omp_set_num_threads(omp_get_max_threads());
#endif
    std::vector<point> pts((size_t)n);
    Vector q(n);
    for (int i = 0; i < n; ++i) {
        pts[i].x = x[i];
        pts[i].y = y[i];
        q(i) = charge[i];
    }
    bbfmm::kernel k;
    k.eval = [kind](point &a, point &b) { return ref_bbfmm_eval(kind, a.x, a.y, b.x, b.y); };
    k.initialize(np, pts, pts, q, n, n, rank, maxLevel);
    Vector pot;
    if (mode == 0) k.run(pot);
    else k.runCache(pot);
    if (mode == 2) {
        k.initialize(np, pts, pts, q, n, n, rank, maxLevel);  // KernelFactory.cpp:356: re-initialised before every runFast
        k.runFast(pot);
    }
    for (int i = 0; i < n; ++i) potential[i] = pot(i);
    if (stats) {
        bbfmm::tree &t = k.t;
        long nodes = (long)t.dict.size(), leaves = 0, su = 0, sv = 0, sw = 0, sx = 0, pairs = 0, mx = 0;
        for (auto &nd : t.dict) {
            if (nd.isLeaf) {
                leaves++;
                if (nd.nSource > mx) mx = nd.nSource;
                for (int u : nd.uList) pairs += (long)t.dict[u].nSource * nd.nTarget;
                for (int u : nd.wList) pairs += (long)t.dict[u].nSource * nd.nTarget;
            }
            su += nd.nUList; sv += nd.nVList; sw += nd.nWList; sx += nd.nXList;
        }
        stats[0] = nodes; stats[1] = leaves; stats[2] = t.maxLevel; stats[3] = su; stats[4] = sv;
        stats[5] = sw; stats[6] = sx; stats[7] = pairs; stats[8] = mx;
    }
    return (int)k.t.dict.size();
}

}  // extern "C"
