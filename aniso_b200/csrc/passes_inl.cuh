// Device helpers shared by the tree-pass kernels (passes.cu, subtree.cu).
#pragma once
#include "device.cuh"

namespace an {

static constexpr int NP_MAX = 16;  // device-side limit on the Chebyshev order (r = np^2 <= 256)

// 1-D interpolation weights of one coordinate: w[m] = S(x, m) = (2 sum_k T_k(x) T_k(t_m) - 1)/np, m < np
// (bbfmm.h:635-656, 742-746).  tn = T_k(t_m) table, m*np + k.
template <int NP>
__device__ __forceinline__ void cheb_weights(const double* __restrict__ tn, int np_rt, double x, double* w) {
    const int np = NP ? NP : np_rt;
    double T[NP ? NP : NP_MAX];
    T[0] = 1.0;
    if (np > 1) T[1] = x;
#pragma unroll
    for (int k = 2; k < (NP ? NP : NP_MAX); ++k)
        if (k < np) T[k] = 2.0 * x * T[k - 1] - T[k - 2];
#pragma unroll
    for (int m = 0; m < (NP ? NP : NP_MAX); ++m) {
        if (m >= np) break;
        double s = -1.0;
#pragma unroll
        for (int k = 0; k < (NP ? NP : NP_MAX); ++k)
            if (k < np) s += 2.0 * T[k] * tn[m * np + k];
        w[m] = s / np;
    }
}

// Sum of the scratch rows of one target (symmetric storage, see stream.cu): the fixed-order reduction that completes
// the cached apply, folded into the kernel that consumes the result.
__device__ __forceinline__ double fold_rows(const DevCSR& m, int g, int t) {
    const int R = m.uniform_block, r1 = m.srow_ptr[g + 1];
    double s0 = 0, s1 = 0;
    int r = m.srow_ptr[g];
    for (; r + 2 <= r1; r += 2) {
        s0 += m.scratch[(size_t)r * R + t];
        s1 += m.scratch[(size_t)(r + 1) * R + t];
    }
    if (r < r1) s0 += m.scratch[(size_t)r * R + t];
    return s0 + s1;
}

}  // namespace an
