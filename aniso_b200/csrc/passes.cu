// K5/K6 Chebyshev passes (upPass bbfmm.h:825-861; L2L / L2P part of downPassFast bbfmm.h:1066-1113) plus the fused
// tail of the matvec.  These move a few tens of MB per matvec: HBM/launch-latency bound, so each kernel makes its
// global loads in one or two coalesced waves, stages the small operands in shared memory and keeps everything else
// in registers.
#include <algorithm>

#include "device.cuh"
#include "passes_inl.cuh"

namespace an {

// q = f .* w with f = x or sigma_s .* x (main.cpp:86-89, 129-131).  f is also scattered into the REFERENCE node order
// (f_ref), where the cell-based correction stencil can index its 3x3 neighbourhood arithmetically.
__global__ void scale_kernel(int n, const double* __restrict__ x, const double* __restrict__ sig_s, const double* __restrict__ w,
                             const int* __restrict__ perm, double* __restrict__ f_ref, double* __restrict__ q) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double v = sig_s ? x[i] * sig_s[i] : x[i];
    if (f_ref) f_ref[perm[i]] = v;
    q[i] = v * w[i];
}
void launch_scale(int n, const double* x, const double* sig_s, const double* w, const int* perm, double* f_ref, double* q,
                  cudaStream_t st) {
    if (n <= 0) return;
    scale_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, x, sig_s, w, perm, f_ref, q);
}

// P2M: nodeCharge[j*np+i] = sum_k Sx(k,i) Sy(k,j) q_k over the leaf's points (bbfmm.h:723-759, 832-844).
// LANES (16 or 32) lanes per leaf.  Phase 1: lane k computes the 2*np weights of point k once (the reference
// recomputes them per matrix entry) from the precomputed box-local coordinate and parks them with q_k in shared
// memory; phase 2: lane m owns mode m and sums over the points.  Covers ALL non-empty leaves: the upward pass is
// replicated on every rank.  Dependent-load depth: leaf table -> points.
template <int NP, int LANES>
__global__ void __launch_bounds__(128) p2m_kernel(DevTree T, DevGeom G, const double* __restrict__ q, double* __restrict__ mcharge) {
    extern __shared__ double sm[];
    __shared__ double tn[NP_MAX * NP_MAX];
    constexpr int UMAX = NP ? (NP * NP + LANES - 1) / LANES : (NP_MAX * NP_MAX + LANES - 1) / LANES;
    const int np = NP ? NP : T.np, r = np * np;
    const int grp = threadIdx.x / LANES, lane = threadIdx.x % LANES;
    const int li = blockIdx.x * (blockDim.x / LANES) + grp;
    const int stride = 2 * np + 1;
    double* __restrict__ ws = sm + (size_t)grp * LANES * stride;
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) tn[i] = T.tnode[i];
    __syncthreads();
    if (li >= T.nleaves) return;
    const unsigned mask = LANES == 32 ? 0xffffffffu : 0xffffu << (16 * ((threadIdx.x >> 4) & 1));
    const int4 info = T.leaf_info[li];  // id, pt_off, npts
    const int id = info.x, off = info.y, n = info.z;
    double acc[UMAX];
#pragma unroll
    for (int u = 0; u < UMAX; ++u) acc[u] = 0;
    for (int k0 = 0; k0 < n; k0 += LANES) {
        const int cnt = min(LANES, n - k0);
        if (lane < cnt) {
            double w[2 * (NP ? NP : NP_MAX)];
            cheb_weights<NP>(tn, np, G.sxn[off + k0 + lane], w);
            cheb_weights<NP>(tn, np, G.syn[off + k0 + lane], w + np);
#pragma unroll
            for (int m = 0; m < 2 * (NP ? NP : NP_MAX); ++m)
                if (m < 2 * np) ws[lane * stride + m] = w[m];
            ws[lane * stride + 2 * np] = q[off + k0 + lane];
        }
        __syncwarp(mask);
#pragma unroll
        for (int u = 0; u < UMAX; ++u) {
            const int m = lane + LANES * u;
            if (m < r) {
                const int i = m % np, j = m / np;
                double a = acc[u];
                for (int k = 0; k < cnt; ++k) a += ws[k * stride + i] * ws[k * stride + np + j] * ws[k * stride + 2 * np];
                acc[u] = a;
            }
        }
        __syncwarp(mask);
    }
#pragma unroll
    for (int u = 0; u < UMAX; ++u) {
        const int m = lane + LANES * u;
        if (m < r) mcharge[(size_t)id * r + m] = acc[u];
    }
}

// M2M: parent[b] = sum_{c non-empty} sum_a R[c][a][b] child_c[a]  (bbfmm.h:855-859).  A CTA takes G parents.  The
// children of a node are numbered consecutively, so its 4r child coefficients are one contiguous run (empty children hold
// zeros); they and the four transfer matrices are fetched into shared memory in one wave of coalesced loads, then thread
// (parent, b) runs down the columns of R.  Dependent-load depth: child0 -> coefficients.
template <bool SMEM_R>
__global__ void __launch_bounds__(256) m2m_kernel(DevTree T, int first, int count, double* __restrict__ mcharge) {
    extern __shared__ double sm[];
    const int r = T.r;
    const int G = max(1, (int)blockDim.x / r);
    const int p0 = blockIdx.x * G;
    double* __restrict__ kids = sm;                // [G][4r]
    double* __restrict__ Rs = sm + (size_t)G * 4 * r;  // [4r][r] when SMEM_R
    __shared__ int c0s[256];
    if (threadIdx.x < G) c0s[threadIdx.x] = p0 + threadIdx.x < count ? T.child0[first + p0 + threadIdx.x] : -1;
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) Rs[i] = __ldg(T.R + i);
    for (int idx = threadIdx.x; idx < G * 4 * r; idx += blockDim.x) {
        const int g = idx / (4 * r), ca = idx - g * 4 * r;
        const int c0 = p0 + g < count ? T.child0[first + p0 + g] : -1;
        kids[idx] = c0 >= 0 ? mcharge[(size_t)c0 * r + ca] : 0.0;
    }
    __syncthreads();
    for (int o = threadIdx.x; o < G * r; o += blockDim.x) {
        const int g = o / r, b = o - g * r;
        if (c0s[g] < 0) continue;  // leaf (written by P2M), empty, or past the end of the level
        const double* __restrict__ src = kids + (size_t)g * 4 * r;
        const double* __restrict__ Rc = SMEM_R ? Rs + b : T.R + b;
        double a0 = 0, a1 = 0;
#pragma unroll 4
        for (int ca = 0; ca < 4 * r; ca += 2) {
            a0 = fma(SMEM_R ? Rc[(size_t)ca * r] : __ldg(Rc + (size_t)ca * r), src[ca], a0);
            a1 = fma(SMEM_R ? Rc[(size_t)(ca + 1) * r] : __ldg(Rc + (size_t)(ca + 1) * r), src[ca + 1], a1);
        }
        mcharge[(size_t)(first + p0 + g) * r + b] = a0 + a1;
    }
}

// L2L: child[a] (+)= sum_b R[index][a][b] parent[b]  (bbfmm.h:1070-1071); RT = R transposed ([c][b][a]) so that the
// threads of a child read consecutive addresses.  FOLD: the child's own M2L sum is still spread over its scratch rows
// (symmetric storage) and is reduced here; otherwise it already sits in mpot.  Dependent-load depth: parent -> coefficients.
template <bool SMEM_R, bool FOLD>
__global__ void __launch_bounds__(256) l2l_kernel(DevTree T, DevCSR m2l, int first, int count, double* __restrict__ mpot) {
    extern __shared__ double sm[];
    const int r = T.r;
    const int G = max(1, (int)blockDim.x / r);
    const int n0 = blockIdx.x * G;
    double* __restrict__ par = sm;                 // [G][r]
    double* __restrict__ Rs = sm + (size_t)G * r;  // [4][r][r] when SMEM_R
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) Rs[i] = __ldg(T.RT + i);
    for (int idx = threadIdx.x; idx < G * r; idx += blockDim.x) {
        const int g = idx / r, b = idx - g * r;
        par[idx] = n0 + g < count ? mpot[(size_t)T.parent[first + n0 + g] * r + b] : 0.0;
    }
    __syncthreads();
    for (int o = threadIdx.x; o < G * r; o += blockDim.x) {
        const int g = o / r, a = o - g * r;
        if (n0 + g >= count) break;
        const int id = first + n0 + g;
        double own;
        if (FOLD) {
            const int grp = m2l.group_of_node[id];
            own = grp >= 0 ? fold_rows(m2l, grp, a) : 0.0;
        } else
            own = mpot[(size_t)id * r + a];
        const size_t roff = (size_t)T.index[id] * r * r + a;
        const double* __restrict__ RT = SMEM_R ? Rs + roff : T.RT + roff;
        const double* __restrict__ src = par + (size_t)g * r;
        double a0 = 0, a1 = 0;
        int b = 0;
#pragma unroll 4
        for (; b + 2 <= r; b += 2) {
            a0 = fma(SMEM_R ? RT[(size_t)b * r] : __ldg(RT + (size_t)b * r), src[b], a0);
            a1 = fma(SMEM_R ? RT[(size_t)(b + 1) * r] : __ldg(RT + (size_t)(b + 1) * r), src[b + 1], a1);
        }
        if (b < r) a0 = fma(SMEM_R ? RT[(size_t)b * r] : __ldg(RT + (size_t)b * r), src[b], a0);
        mpot[(size_t)id * r + a] = own + (a0 + a1);
    }
}

// The top of the tree in one launch each way: the levels with at most 256 nodes are walked by a single CTA with a
// block barrier between levels (saves two launches of ~9 us each way at N = 2^20).
template <bool SMEM_R>
__global__ void __launch_bounds__(1024) m2m_top_kernel(DevTree T, int Lhi, int Llo, double* __restrict__ mcharge) {
    extern __shared__ double Rs[];
    const int r = T.r;
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) Rs[i] = __ldg(T.R + i);
    __syncthreads();
    for (int L = Lhi; L >= Llo; --L) {
        const int first = T.level_ptr[L], count = T.level_ptr[L + 1] - first;
        for (int o = threadIdx.x; o < count * r; o += blockDim.x) {
            const int g = o / r, b = o - g * r;
            const int c0 = T.child0[first + g];
            if (c0 < 0) continue;
            const double* __restrict__ src = mcharge + (size_t)c0 * r;
            const double* __restrict__ Rc = SMEM_R ? Rs + b : T.R + b;
            double a0 = 0, a1 = 0;
#pragma unroll 4
            for (int ca = 0; ca < 4 * r; ca += 2) {
                a0 = fma(SMEM_R ? Rc[(size_t)ca * r] : __ldg(Rc + (size_t)ca * r), src[ca], a0);
                a1 = fma(SMEM_R ? Rc[(size_t)(ca + 1) * r] : __ldg(Rc + (size_t)(ca + 1) * r), src[ca + 1], a1);
            }
            mcharge[(size_t)(first + g) * r + b] = a0 + a1;
        }
        __syncthreads();
    }
}

template <bool SMEM_R, bool FOLD>
__global__ void __launch_bounds__(1024) l2l_top_kernel(DevTree T, DevCSR m2l, int Llo, int Lhi, double* __restrict__ mpot) {
    extern __shared__ double Rs[];
    const int r = T.r;
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) Rs[i] = __ldg(T.RT + i);
    __syncthreads();
    for (int L = Llo; L <= Lhi; ++L) {
        const int first = T.level_ptr[L], count = T.level_ptr[L + 1] - first;
        for (int o = threadIdx.x; o < count * r; o += blockDim.x) {
            const int g = o / r, a = o - g * r;
            const int id = first + g;
            double own;
            if (FOLD) {
                const int grp = m2l.group_of_node[id];
                own = grp >= 0 ? fold_rows(m2l, grp, a) : 0.0;
            } else
                own = mpot[(size_t)id * r + a];
            const double* __restrict__ src = mpot + (size_t)T.parent[id] * r;
            const size_t roff = (size_t)T.index[id] * r * r + a;
            const double* __restrict__ RT = SMEM_R ? Rs + roff : T.RT + roff;
            double a0 = 0, a1 = 0;
            int b = 0;
#pragma unroll 4
            for (; b + 2 <= r; b += 2) {
                a0 = fma(SMEM_R ? RT[(size_t)b * r] : __ldg(RT + (size_t)b * r), src[b], a0);
                a1 = fma(SMEM_R ? RT[(size_t)(b + 1) * r] : __ldg(RT + (size_t)(b + 1) * r), src[b + 1], a1);
            }
            if (b < r) a0 = fma(SMEM_R ? RT[(size_t)b * r] : __ldg(RT + (size_t)b * r), src[b], a0);
            mpot[(size_t)id * r + a] = own + (a0 + a1);
        }
        __syncthreads();
    }
}

// L2P (bbfmm.h:1104-1113) and the tail of `mapping` in one pass over the owned targets:
//   fmm  = pot[p] + sum_{i,j} Sx(p,i) Sy(p,j) nodePotential[leaf][j*np+i]
//   mode 0 (stage): out = fmm
//   mode 1: out = (fmm + stencil . f) / 2pi                       AnisoWrapper.cpp:129-130
//   mode 2: out = x - (fmm + stencil . f) / 2pi                   forwardOperator, main.cpp:129-134
// f_ref is the unscaled source in REFERENCE node order: the 3x3 cell neighbourhood is plain index arithmetic.
template <int NP>
__global__ void __launch_bounds__(128) l2p_kernel(DevTree T, DevGeom G, DevCSR p2p, int fold, const int* __restrict__ leaf_of_slot,
                                                  int slot0, int n,
                                                  const double* __restrict__ mpot, const double* __restrict__ pot,
                                                  const double* __restrict__ C, const double* __restrict__ f_ref,
                                                  const double* __restrict__ x, double* __restrict__ out, int mode) {
    __shared__ double tn[NP_MAX * NP_MAX];
    const int np = NP ? NP : T.np;
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) tn[i] = T.tnode[i];
    __syncthreads();
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int p = slot0 + k;
    const int id = leaf_of_slot[p];
    const double* __restrict__ mp = mpot + (size_t)id * np * np;
    double corr = 0;
    if (mode) {  // issue the stencil loads first: they do not depend on the leaf
        const int c2 = G.c2, sz = G.sz;
        const int tc = G.perm[p] / c2;
        const int ti = tc / sz, tj = tc - ti * sz;
        for (int cs = 0; cs < 9; ++cs) {
            const int si = ti + cs / 3 - 1, sj = tj + cs % 3 - 1;
            if (si < 0 || si >= sz || sj < 0 || sj >= sz) continue;
            const double* __restrict__ fs = f_ref + (size_t)(si * sz + sj) * c2;
            for (int q = 0; q < c2; ++q) corr = fma(C[(size_t)(cs * c2 + q) * n + k], fs[q], corr);
        }
    }
    double wx[NP ? NP : NP_MAX], wy[NP ? NP : NP_MAX];
    cheb_weights<NP>(tn, np, G.sxn[p], wx);
    cheb_weights<NP>(tn, np, G.syn[p], wy);
    double acc = 0;
#pragma unroll
    for (int j = 0; j < (NP ? NP : NP_MAX); ++j) {
        if (j >= np) break;
        double row = 0;
#pragma unroll
        for (int i = 0; i < (NP ? NP : NP_MAX); ++i)
            if (i < np) row = fma(wx[i], mp[j * np + i], row);
        acc = fma(wy[j], row, acc);
    }
    double near;  // P2P sum of this target: already reduced in pot, or still spread over the leaf's scratch rows
    if (fold) {
        const int grp = p2p.group_of_node[id];
        near = grp >= 0 ? fold_rows(p2p, grp, p - p2p.row_off[grp]) : 0.0;
    } else
        near = pot[p];
    double v = near + acc;
    if (mode) {
        v = (v + corr) * (M_1_PI / 2.0);
        if (mode == 2) v = x[k] - v;
    }
    out[k] = v;
}

template <int NP>
static void p2m_launch(const DevTree& T, const DevGeom& G, const double* q, double* mcharge, cudaStream_t st) {
    const int threads = 128;
    if (T.r <= 16 && T.max_leaf <= 16) {
        const int groups = threads / 16;
        p2m_kernel<NP, 16><<<(T.nleaves + groups - 1) / groups, threads, (size_t)groups * 16 * (2 * T.np + 1) * sizeof(double), st>>>(
            T, G, q, mcharge);
    } else {
        const int groups = threads / 32;
        p2m_kernel<NP, 32><<<(T.nleaves + groups - 1) / groups, threads, (size_t)groups * 32 * (2 * T.np + 1) * sizeof(double), st>>>(
            T, G, q, mcharge);
    }
}
void launch_p2m(const DevTree& T, const DevGeom& G, const double* q, double* mcharge, cudaStream_t st) {
    if (T.nleaves == 0) return;
    switch (T.np) {
        case 4: p2m_launch<4>(T, G, q, mcharge, st); break;
        case 6: p2m_launch<6>(T, G, q, mcharge, st); break;
        case 8: p2m_launch<8>(T, G, q, mcharge, st); break;
        default: p2m_launch<0>(T, G, q, mcharge, st);
    }
}
static int pass_threads(int r) { return r <= 128 ? 128 : ((r + 31) / 32) * 32; }
static bool smem_R(int r) { return (size_t)4 * r * r * sizeof(double) <= 36 * 1024; }  // r <= 33 (np <= 5)
void launch_m2m(const DevTree& T, int first, int count, double* mcharge, cudaStream_t st) {
    if (count == 0) return;
    const int threads = pass_threads(T.r), G = std::max(1, threads / T.r);
    const size_t kids = (size_t)G * 4 * T.r * sizeof(double);
    if (smem_R(T.r))
        m2m_kernel<true><<<(count + G - 1) / G, threads, kids + (size_t)4 * T.r * T.r * sizeof(double), st>>>(T, first, count, mcharge);
    else
        m2m_kernel<false><<<(count + G - 1) / G, threads, kids, st>>>(T, first, count, mcharge);
}
void launch_l2l(const DevTree& T, const DevCSR& m2l, bool fold, int first, int count, double* mpot, cudaStream_t st) {
    if (count == 0) return;
    const int threads = pass_threads(T.r), G = std::max(1, threads / T.r);
    const size_t par = (size_t)G * T.r * sizeof(double), rb = (size_t)4 * T.r * T.r * sizeof(double);
    const unsigned blocks = (count + G - 1) / G;
    if (smem_R(T.r)) {
        if (fold) l2l_kernel<true, true><<<blocks, threads, par + rb, st>>>(T, m2l, first, count, mpot);
        else l2l_kernel<true, false><<<blocks, threads, par + rb, st>>>(T, m2l, first, count, mpot);
    } else {
        if (fold) l2l_kernel<false, true><<<blocks, threads, par, st>>>(T, m2l, first, count, mpot);
        else l2l_kernel<false, false><<<blocks, threads, par, st>>>(T, m2l, first, count, mpot);
    }
}
void launch_m2m_top(const DevTree& T, int Lhi, int Llo, double* mcharge, cudaStream_t st) {
    if (Lhi < Llo) return;
    if (smem_R(T.r)) m2m_top_kernel<true><<<1, 1024, (size_t)4 * T.r * T.r * sizeof(double), st>>>(T, Lhi, Llo, mcharge);
    else m2m_top_kernel<false><<<1, 1024, 0, st>>>(T, Lhi, Llo, mcharge);
}
void launch_l2l_top(const DevTree& T, const DevCSR& m2l, bool fold, int Llo, int Lhi, double* mpot, cudaStream_t st) {
    if (Lhi < Llo) return;
    const size_t rb = (size_t)4 * T.r * T.r * sizeof(double);
    if (smem_R(T.r)) {
        if (fold) l2l_top_kernel<true, true><<<1, 1024, rb, st>>>(T, m2l, Llo, Lhi, mpot);
        else l2l_top_kernel<true, false><<<1, 1024, rb, st>>>(T, m2l, Llo, Lhi, mpot);
    } else {
        if (fold) l2l_top_kernel<false, true><<<1, 1024, 0, st>>>(T, m2l, Llo, Lhi, mpot);
        else l2l_top_kernel<false, false><<<1, 1024, 0, st>>>(T, m2l, Llo, Lhi, mpot);
    }
}
void launch_l2p(const DevTree& T, const DevGeom& G, const DevCSR& p2p, bool fold_p2p, const int* leaf_of_slot, int slot0, int n,
                const double* mpot,
                const double* pot, const double* C, const double* f_ref, const double* x, double* out, int mode,
                cudaStream_t st) {
    if (n == 0) return;
    const int blocks = (n + 127) / 128;
    switch (T.np) {
        case 4: l2p_kernel<4><<<blocks, 128, 0, st>>>(T, G, p2p, fold_p2p ? 1 : 0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode); break;
        case 6: l2p_kernel<6><<<blocks, 128, 0, st>>>(T, G, p2p, fold_p2p ? 1 : 0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode); break;
        case 8: l2p_kernel<8><<<blocks, 128, 0, st>>>(T, G, p2p, fold_p2p ? 1 : 0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode); break;
        default: l2p_kernel<0><<<blocks, 128, 0, st>>>(T, G, p2p, fold_p2p ? 1 : 0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode);
    }
}

}  // namespace an
