// Standalone solve with the drop-in C++ API, fast path: the same problem the reference's main.cpp sets up (config file,
// Gaussian source, sigma_s = 8(1 - cos 2 pi x), sigma_t = sigma_s + 0.2), but using the fused device matvec
// (Aniso::mapping) for the right-hand side and the device-resident GMRES overload instead of the host callback.
//   usage: solve <config file> [output csv]
#include <iostream>

#include "Aniso.h"
#include "gmres.h"
#include "matlab_io.h"
#include "utility/config.h"

int main(int argc, char* argv[]) {
    if (argc < 2) {
        std::cerr << "usage: solve <config> [result.csv]" << std::endl;
        return 2;
    }
    config cfg(argv[1]);
    cfg.print();
    auto opt = [&](const char* k) { return cfg.options[k]; };
    Aniso aniso(atoi(opt("domainSize").c_str()), atoi(opt("quadRule").c_str()), atoi(opt("kernelSize").c_str()),
                atof(opt("g").c_str()), atoi(opt("singRule").c_str()), atoi(opt("np").c_str()), atoi(opt("maxLevel").c_str()));
    Profiler timer;
    const int n = aniso.numberOfNodes;
    Vector charge(n);
    for (int i = 0; i < n; ++i) {
        const double x = aniso.nodes[i].x, y = aniso.nodes[i].y;
        charge(i) = exp(-25 * (SQR(x - 0.5) + SQR(y - 0.5)));
        aniso.sigma_s(i) = 16 * 0.5 * (1 - cos(2 * M_PI * x));
        aniso.sigma_t(i) = aniso.sigma_s(i) + 0.2;
    }
    timer.tic("interpolate sigma");
    aniso.interpolation();
    timer.toc();
    timer.tic("cache (fused operators + stencils)");
    if (aniso_cache(aniso.handle(), 0) < 0) {
        std::cerr << aniso_last_error() << std::endl;
        return 1;
    }
    timer.toc();
    Vector rhs;
    timer.tic("rhs = mapping(charge, 0)");
    aniso.mapping(0, charge, rhs);
    timer.toc();
    Vector x(n);
    timer.tic("GMRES (device resident)");
    const int rc = GMRES(aniso, x, rhs, 80, 400, 1e-12);
    timer.toc();
    if (argc > 2) write_to_csv(x, argv[2]);
    aniso.displayKernelCacheSize();
    return rc;
}
