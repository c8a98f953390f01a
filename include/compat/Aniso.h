// aniso_b200 drop-in header: class Aniso (Aniso.h:12-22, Aniso.cpp:7-47).
#ifndef ANISO_B200_COMPAT_ANISO_H
#define ANISO_B200_COMPAT_ANISO_H

#include "Geometry.h"
#include "KernelFactory.h"
#include "utility/Profiler.h"

class Aniso : public KernelFactory {
public:
    Aniso(int geometry_size, int geometry_degree, int kernel_size, double kernel_g, int kernel_degree, int np_, int maxLevel_)
        : KernelFactory(geometry_size, geometry_degree, kernel_size, kernel_g, kernel_degree) {
        this->np = np_;
        this->maxLevel = maxLevel_;
        ensure();  // nodes, weights, numberOfNodes, sigma_s, sigma_t are usable right after construction
    }
    ~Aniso() {}

    void displayKernelCacheSize() {  // Aniso.cpp:19-47: one line per kernel id
        for (int i = 0; i < kernelSize; ++i) {
            const double res = double(aniso_cache_bytes(handle(), i)) / 1024.0 / 1024.0;
            if (res < 1024) std::cout << res << " MByte" << std::endl;
            else std::cout << res / 1024.0 << " GByte" << std::endl;
        }
    }
};

#endif
