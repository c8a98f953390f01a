// aniso_b200 drop-in header: Gauss-Legendre rules in the reference's table order (Quadrature.h / Quadrature.cpp:5418-22199).
#ifndef ANISO_B200_COMPAT_QUADRATURE_H
#define ANISO_B200_COMPAT_QUADRATURE_H

#include "aniso_b200.h"
#include "bbfmm/utils.h"

class Quadrature {
public:
    vector<scalar_t> points_x, points_y, weights;  // 1-D rules fill points_x / weights
    Quadrature() {}
    virtual ~Quadrature() {}
};

// get_legendre_data(deg, table): one point and one weight per row; prints the reference's message for deg outside 1..128
inline void get_legendre_data(size_t deg, Quadrature& table) {
    table.points_x.assign(deg, 0.0);
    table.weights.assign(deg, 0.0);
    if (aniso_gl_table((int)deg, table.points_x.data(), table.weights.data()) != 0) {
        std::cout << "The degree is not implemented." << std::endl;
        table.points_x.clear();
        table.weights.clear();
    }
}
// affine(): map the rule from [-1,1] to [0,1] (Quadrature.cpp:22194-22199)
inline void affine(Quadrature& table) {
    for (size_t k = 0; k < table.weights.size(); ++k) {
        table.weights[k] /= 2.0;
        table.points_x[k] += 1.0;
        table.points_x[k] /= 2.0;
    }
}

#endif
