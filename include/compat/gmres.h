// aniso_b200 drop-in header: GMRES with the reference's entry points (gmres.h:13-24, gmres.cpp:12-169).
//   GMRES(A, x, b, m, max_iter, tol)        any host operator callback, host Vectors: restarted GMRES(m) with modified
//                                           Gram-Schmidt and Givens rotations, same stopping tests and printout.
//   GMRES(aniso, x, b, m, max_iter, tol)    the main.cpp operator x - mapping_0(sigma_s x) with the whole Arnoldi process
//                                           device-resident (aniso_gmres).
#ifndef ANISO_B200_COMPAT_GMRES_H
#define ANISO_B200_COMPAT_GMRES_H

#include "Aniso.h"
#include "bbfmm/blas_wrapper.h"

inline void GeneratePlaneRotation(scalar_t& dx, scalar_t& dy, scalar_t& cs, scalar_t& sn) {
    if (dy == 0.0) {
        cs = 1.0;
        sn = 0.0;
    } else if (std::fabs(dy) > std::fabs(dx)) {
        const scalar_t t = dx / dy;
        sn = 1.0 / std::sqrt(1.0 + t * t);
        cs = t * sn;
    } else {
        const scalar_t t = dy / dx;
        cs = 1.0 / std::sqrt(1.0 + t * t);
        sn = t * cs;
    }
}
inline void ApplyPlaneRotation(scalar_t& dx, scalar_t& dy, scalar_t& cs, scalar_t& sn) {
    const scalar_t t = cs * dx + sn * dy;
    dy = -sn * dx + cs * dy;
    dx = t;
}
inline void Update(Vector& x, int k, Matrix& h, Vector& s, Vector v[]) {
    Vector y(s);
    for (int i = k; i >= 0; i--) {
        y(i) /= h(i, i);
        for (int j = i - 1; j >= 0; j--) y(j) -= h(j, i) * y(i);
    }
    for (int j = 0; j <= k; j++) daxpy(y(j), v[j], x);
}

namespace aniso_b200_detail {
inline void gmres_header() {
    std::cout << "=============== GMRES =================" << std::endl;
    std::cout << "    iter    |  rel error   |   time   " << std::endl;
}
inline void gmres_row(int j, scalar_t resid, double seconds) {
    std::cout << std::setw(6) << j << std::setw(20) << std::scientific << resid << std::setw(12) << std::fixed << seconds
              << std::fixed << std::endl;
}
}  // namespace aniso_b200_detail

inline int GMRES(const std::function<Vector(Vector&)> A, Vector& x, Vector& b, int m, int max_iter, scalar_t tol) {
    using clk = std::chrono::steady_clock;
    const int n = x.row();
    int i = 0, j = 1;
    Vector s(m + 1), cs(m + 1), sn(m + 1);
    scalar_t normb = nrm2(b);
    Vector r = A(x);
    dscal(-1.0, r);
    daxpy(1.0, b, r);
    scalar_t beta = nrm2(r), resid;
    aniso_b200_detail::gmres_header();
    auto begin = clk::now();
    if (normb == 0.0) normb = 1;
    if ((resid = nrm2(r) / normb) <= tol) return 0;
    std::vector<Vector> v(m + 1);
    Matrix H(m + 1, m);
    while (j <= max_iter) {
        v[0] = r;
        dscal(1.0 / beta, v[0]);
        setValue(s, 0.);
        s(0) = beta;
        for (i = 0; i < m && j <= max_iter; i++, j++) {
            aniso_b200_detail::gmres_row(j, resid, std::chrono::duration_cast<std::chrono::microseconds>(clk::now() - begin).count() / 1e6);
            begin = clk::now();
            Vector w = A(v[i]);
            for (int k = 0; k <= i; k++) {
                H(k, i) = ddot(w, v[k]);
                daxpy(-H(k, i), v[k], w);
            }
            H(i + 1, i) = nrm2(w);
            v[i + 1] = w;
            dscal(1.0 / H(i + 1, i), v[i + 1]);
            for (int k = 0; k < i; k++) ApplyPlaneRotation(H(k, i), H(k + 1, i), cs(k), sn(k));
            GeneratePlaneRotation(H(i, i), H(i + 1, i), cs(i), sn(i));
            ApplyPlaneRotation(H(i, i), H(i + 1, i), cs(i), sn(i));
            ApplyPlaneRotation(s(i), s(i + 1), cs(i), sn(i));
            if ((resid = std::fabs(s(i + 1)) / normb) < tol) {
                Update(x, i, H, s, v.data());
                std::cout << "=============== CONVERGED =============" << std::endl;
                return 0;
            }
        }
        Update(x, i - 1, H, s, v.data());
        r = A(x);
        dscal(-1.0, r);
        daxpy(1.0, b, r);
        beta = nrm2(r);
        if ((resid = beta / normb) < tol) {
            std::cout << "=============== CONVERGED =============" << std::endl;
            return 0;
        }
    }
    (void)n;
    std::cout << "============= NOT CONVERGED ===========" << std::endl;
    return 1;
}

// Device-resident solve of (I - K sigma_s) x = b (the forwardOperator of main.cpp:123-136).
inline int GMRES(Aniso& aniso, Vector& x, Vector& b, int m, int max_iter, scalar_t tol) {
    std::vector<double> hist(max_iter + 2, 0.0);
    int iters = 0;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = aniso_gmres(aniso.handle(), b.data(), x.data(), m, max_iter, tol, &iters, hist.data(), (int)hist.size());
    if (rc < 0) throw std::runtime_error(std::string("aniso_b200: ") + aniso_last_error());
    const double dt = std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count() / 1e6;
    aniso_b200_detail::gmres_header();
    for (int j = 1; j <= iters; ++j) aniso_b200_detail::gmres_row(j, hist[j - 1], dt / (iters ? iters : 1));
    std::cout << (rc == 0 ? "=============== CONVERGED =============" : "============= NOT CONVERGED ===========") << std::endl;
    return rc;
}

#endif
