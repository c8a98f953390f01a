// aniso_b200 drop-in header: class KernelFactory with the public interface of the reference (KernelFactory.h:14-92).
// Every method forwards to libaniso_b200.so (include/aniso_b200.h); nothing is computed on the host.  Differences in
// behaviour are limited to error handling: a failing device call throws std::runtime_error(aniso_last_error()).
#ifndef ANISO_B200_COMPAT_KERNELFACTORY_H
#define ANISO_B200_COMPAT_KERNELFACTORY_H

#include "Geometry.h"
#include "Quadrature.h"
#include "aniso_b200.h"
#include "bbfmm/bbfmm.h"

using namespace std::placeholders;

class KernelFactory : public Geometry {
public:
    KernelFactory() {}
    // KernelFactory.cpp:7-54.  np and maxLevel are members the Aniso constructor sets afterwards in the reference
    // (Aniso.cpp:7-13); the device handle is therefore created lazily, at the first call that needs it.
    KernelFactory(int geometry_size, int geometry_degree, int kernel_size, double kernel_g, int kernel_degree)
        : Geometry(geometry_size, geometry_degree), kernelSize(2 * kernel_size - 1), kernelArg_(kernel_size),
          g_(kernel_g), singDegree_(kernel_degree) {
        realParts.resize(kernelSize);
        imagParts.resize(kernelSize);
        anisotropy = Vector(kernelSize);
        for (int i = 0; i < kernelSize; ++i) anisotropy(i) = std::pow(kernel_g, i);  // never read (quirk Q12)
        get_legendre_data(kernel_degree, singQuadratureRule);
        affine(singQuadratureRule);
    }
    virtual ~KernelFactory() {
        if (handle_) aniso_destroy(handle_);
    }
    KernelFactory(const KernelFactory&) = delete;
    KernelFactory& operator=(const KernelFactory&) = delete;

    vector<kernel> realParts;
    vector<kernel> imagParts;
    int kernelSize = 0;
    Vector anisotropy;
    Quadrature singQuadratureRule;
    Vector sigma_t;
    Vector sigma_s;
    index_t np = 0;
    index_t maxLevel = 0;
    vector<vector<scalar_t>> singX, singY, singW;
    vector<Vector> sigma_t_Coeff;
    vector<Vector> sigma_s_Coeff;

    aniso_t* handle() {  // the C-ABI handle behind this object
        ensure();
        return handle_;
    }

    scalar_t lineIntegral(point& p, point& q) { return lineIntegral(p.x, p.y, q.x, q.y); }
    scalar_t lineIntegral(double x0, double y0, double x1, double y1) {
        double pts[4] = {x0, y0, x1, y1}, out = 0;
        ck(aniso_line_integral(handle(), 1, pts, &out));
        return out;
    }
    scalar_t evaluate(point& p) {  // sigma_t interpolant at p: the diagonal value of imagParts[0] (quirk Q2)
        double pts[4] = {p.x, p.y, p.x, p.y}, out = 0;
        ck(aniso_kernel_eval(handle(), 0, ANISO_PART_IMAG, 1, pts, &out));
        return out;
    }
    int getRow(double y) { return int(std::floor(y * sz)); }
    int getCol(double x) { return int(std::floor(x * sz)); }

    // interpolation(): sigma_s / sigma_t -> per-cell Legendre coefficients; uploads the coefficient grid (KernelFactory.cpp:212-227)
    void interpolation() {
        ensure();
        ck(aniso_set_coeff(handle_, sigma_s.data(), sigma_t.data()));
        vector<scalar_t> ct(numberOfNodes), cs(numberOfNodes);
        ck(aniso_sigma_coeff(handle_, ct.data(), cs.data()));
        unpack(ct, sigma_t_Coeff);
        unpack(cs, sigma_s_Coeff);
    }
    void interpolation(Vector& h, vector<Vector>& h_coeff) {  // KernelFactory.cpp:988-1005
        vector<scalar_t> c(numberOfNodes);
        ck(aniso_interpolation(handle(), h.data(), c.data()));
        unpack(c, h_coeff);
    }
    void singPrecompute() {  // KernelFactory.cpp:929-986 (tables are built with the handle)
        ensure();
        const int c2 = deg * deg, ns = 8 * singDegree_ * singDegree_;
        vector<scalar_t> X((size_t)c2 * ns), Y((size_t)c2 * ns), W((size_t)c2 * ns);
        ck(aniso_sing_tables(handle_, X.data(), Y.data(), W.data()));
        singX.assign(c2, vector<scalar_t>(ns));
        singY = singW = singX;
        for (int t = 0; t < c2; ++t)
            for (int q = 0; q < ns; ++q) {
                singX[t][q] = X[(size_t)t * ns + q];
                singY[t][q] = Y[(size_t)t * ns + q];
                singW[t][q] = W[(size_t)t * ns + q];
            }
    }
    void makeKernels() {  // KernelFactory.cpp:240-267: eval(source a, target b), evaluated on the device
        ensure();
        for (int k = 0; k < kernelSize; ++k) {
            realParts[k].eval = [this, k](point& a, point& b) { return evalOne(k, ANISO_PART_REAL, a, b); };
            imagParts[k].eval = [this, k](point& a, point& b) { return evalOne(k, ANISO_PART_IMAG, a, b); };
            realParts[k].nChebyshev = imagParts[k].nChebyshev = np;
            realParts[k].rank = imagParts[k].rank = np * np;
            realParts[k].maxLevel = imagParts[k].maxLevel = maxLevel;
        }
    }

    void runKernelsCacheSing(int i) { cachePart(i, ANISO_PART_REAL); }  // KernelFactory.cpp:279-288
    void runKernelsCache(int i) { cachePart(i, ANISO_PART_IMAG); }      // KernelFactory.cpp:323-334
    void refineAddOnCache(int i) { cachePart(i, 0); }                   // KernelFactory.cpp:550-609
    void singularAddCache(int i) { cachePart(i, 0); }                   // KernelFactory.cpp:752-788

    void nearRemoval(int i, Vector& f, Vector& ret) { ck(aniso_near_removal(handle(), i, f.data(), ret.data())); }
    void runKernelsFastSing(int i, Vector& f, Vector& ret) {
        ret = Vector(numberOfNodes);  // the reference overwrites/resizes ret (bbfmm.h:579)
        ck(aniso_run_kernels_fast_sing(handle(), i, f.data(), ret.data()));
    }
    void runKernelsFast(int i, Vector& f, Vector& ret) {
        ret = Vector(numberOfNodes);
        ck(aniso_run_kernels_fast(handle(), i, f.data(), ret.data()));
    }
    void refineAddOnFast(int i, Vector& f, Vector& ret) { ck(aniso_refine_add_on_fast(handle(), i, f.data(), ret.data())); }
    void singularAddFast(int i, vector<Vector>& fCoeff, Vector& ret) {
        const int c2 = deg * deg;
        vector<scalar_t> flat(numberOfNodes);
        for (size_t c = 0; c < fCoeff.size(); ++c)
            for (int r = 0; r < c2; ++r) flat[c * c2 + r] = fCoeff[c](r);
        ck(aniso_singular_add_fast(handle(), i, flat.data(), ret.data()));
    }
    // mapping(charge, id) of AnisoWrapper.cpp:92-136 in one fused device pass (not in the reference class; used by the
    // rebuilt MEX wrapper and by callers that do not need the stages)
    void mapping(int i, Vector& charge, Vector& ret) {
        ret = Vector(numberOfNodes);
        ck(aniso_mapping(handle(), charge.data(), i, ret.data()));
    }

protected:
    aniso_t* handle_ = nullptr;
    int kernelArg_ = 1, singDegree_ = 1;
    double g_ = 0;

    static void ck(int rc) {
        if (rc < 0) throw std::runtime_error(std::string("aniso_b200: ") + aniso_last_error());
    }
    void ensure() {
        if (handle_) return;
        // ANISO_DEVICES=0,1,2,3: the single-process callers of the reference (main.cpp, the MEX gateway) drive several
        // GPUs through one group handle (aniso_create_multi); unset: the current device
        std::vector<int> devs;
        if (const char* env = std::getenv("ANISO_DEVICES"))
            for (const char* q = env; *q;) {
                char* end = nullptr;
                const long d = std::strtol(q, &end, 10);
                if (end == q) break;
                devs.push_back((int)d);
                q = *end ? end + 1 : end;
            }
        if (devs.size() > 1) ck(aniso_create_multi(sz, deg, kernelArg_, g_, singDegree_, np, maxLevel, (int)devs.size(), devs.data(), &handle_));
        else ck(aniso_create(sz, deg, kernelArg_, g_, singDegree_, np, maxLevel, &handle_));
        loadGeometry(handle_);
        if (sigma_t.row() != numberOfNodes) sigma_t = Vector(numberOfNodes);
        if (sigma_s.row() != numberOfNodes) sigma_s = Vector(numberOfNodes);
    }
    void cachePart(int i, int part) {
        ck(aniso_cache_part(handle(), i, part));
        const size_t b = aniso_cache_bytes(handle_, i);
        realParts[i].cacheBytes = b / 2;
        imagParts[i].cacheBytes = b - b / 2;
    }
    scalar_t evalOne(int k, int part, point& a, point& b) {
        double pts[4] = {a.x, a.y, b.x, b.y}, out = 0;
        ck(aniso_kernel_eval(handle_, k, part, 1, pts, &out));
        return out;
    }
    void unpack(const vector<scalar_t>& flat, vector<Vector>& out) {
        const int c2 = deg * deg;
        out.resize(numberOfSquares);
        for (int c = 0; c < numberOfSquares; ++c) {
            out[c] = Vector(c2);
            for (int r = 0; r < c2; ++r) out[c](r) = flat[(size_t)c * c2 + r];
        }
    }
};

#endif
