// aniso_b200 drop-in header: class Geometry with the public data members of the reference (Geometry.h:31-63), filled
// from the host-side tables of libaniso_b200 (aniso_get_nodes, aniso_geometry_tables).
#ifndef ANISO_B200_COMPAT_GEOMETRY_H
#define ANISO_B200_COMPAT_GEOMETRY_H

#include "Quadrature.h"
#include "aniso_b200.h"
#include "bbfmm/bbfmm.h"
#include "bbfmm/linalg.h"
#include "bbfmm/utils.h"
#include "utility/Profiler.h"

#define DEFAULT_REFINE_LEVEL 2

using namespace bbfmm;

class Geometry {
public:
    Geometry() {}
    Geometry(int size, int degree) : deg(degree), sz(size) {}
    virtual ~Geometry() {}

    vector<point> nodes;
    vector<scalar_t> weights;
    int deg = 0;
    int sz = 0;
    Quadrature volQuadratureRule;
    vector<scalar_t> refine_quad_x, refine_quad_y, refine_weight;
    int coarseQuadratureSize = 0;
    int refineQuadratureSize = 0;
    Matrix interpolate;
    Matrix nearMapping;
    Vector sqrtWeights;
    Vector legendreNorms;
    int numberOfSquares = 0;
    int numberOfNodes = 0;
    double dx = 0;

protected:
    // fills every public member from an aniso_b200 handle
    void loadGeometry(aniso_t* h) {
        numberOfNodes = aniso_num_nodes(h);
        numberOfSquares = sz * sz;
        dx = 1.0 / sz;
        coarseQuadratureSize = deg * deg;
        refineQuadratureSize = 16 * deg * deg;
        vector<scalar_t> x(numberOfNodes), y(numberOfNodes);
        weights.resize(numberOfNodes);
        aniso_get_nodes(h, x.data(), y.data(), weights.data());
        nodes.resize(numberOfNodes);
        for (int i = 0; i < numberOfNodes; ++i) nodes[i] = point(x[i], y[i]);
        get_legendre_data(deg, volQuadratureRule);
        const int c2 = coarseQuadratureSize, rq = refineQuadratureSize;
        vector<scalar_t> I((size_t)c2 * c2), M((size_t)rq * c2);
        legendreNorms.resize(c2);
        refine_quad_x.resize(rq);
        refine_quad_y.resize(rq);
        refine_weight.resize(rq);
        aniso_geometry_tables(h, I.data(), M.data(), legendreNorms.data(), refine_quad_x.data(), refine_quad_y.data(),
                              refine_weight.data());
        interpolate.resize(c2, c2);
        nearMapping.resize(rq, c2);
        for (int r = 0; r < c2; ++r)
            for (int c = 0; c < c2; ++c) interpolate(r, c) = I[(size_t)r * c2 + c];
        for (int r = 0; r < rq; ++r)
            for (int c = 0; c < c2; ++c) nearMapping(r, c) = M[(size_t)r * c2 + c];
        sqrtWeights.resize(c2);
        for (int a = 0; a < deg; ++a)
            for (int b = 0; b < deg; ++b)
                sqrtWeights(a * deg + b) = std::sqrt(volQuadratureRule.weights[a] * volQuadratureRule.weights[b]);
    }
};

#endif
