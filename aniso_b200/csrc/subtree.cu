// Tree passes of the matvec for COMPLETE UNIFORM quadtrees (every BASELINE configuration; SURVEY 8: "all BASELINE
// trees are complete uniform quadtrees"), restructured around sub-trees instead of levels.
//
// The level-by-level passes of passes.cu (upPass bbfmm.h:825-861; L2L/L2P of downPassFast bbfmm.h:1066-1113) make a
// chain of 2 x (depth - 2) dependent launches, each bounded by launch + DRAM latency, not by bytes: at N = 2^20 they
// take 0.31 ms of a 0.82 ms matvec.  Here one CTA owns a sub-tree of up to three levels and walks it in shared memory
// with block barriers, so a pass over the whole tree is ceil(levels / 3) launches and the intermediate levels never
// leave the SM:
//   up_leaf_kernel    points -> multipoles of the leaf level (P2M) and of the up-to-3 levels above it (M2M)
//   up_mid_kernel     multipoles of level Ls+D (global) -> levels Ls+D-1 .. Ls
//   down_mid_kernel   locals of level Ls (global, or the folded M2L sum when Ls is the first M2L level) -> level Ls+D
//   down_leaf_kernel  the same down to the leaf level, then L2P + near-field fold + correction stencil + 1/2pi
//                     (the fused tail of `mapping`, AnisoWrapper.cpp:129-130 / main.cpp:129-134)
// Node numbering (plan.cpp): level-major, Morton order inside a level, and level L of a complete tree starts at node
// (4^L - 1)/3 = lvoff(L); so the descendants of node m of level L at level L+d are the 4^d consecutive nodes starting
// at lvoff(L+d) + 4^d m, leaf k of the DFS order is node lvoff(deepest) + k, and a sub-tree's points are one
// contiguous slot range.
#include <algorithm>
#include <cstdlib>
#include <functional>

#include "device.cuh"
#include "passes_inl.cuh"

namespace an {

__host__ __device__ __forceinline__ int lvoff(int d) { return ((1 << (2 * d)) - 1) / 3; }  // boxes above level d of a sub-tree

// M2M inside shared memory: levels D-1 .. 0 of the sub-tree rooted at node (Ls, m); sm holds level d at lvoff(d)*r.
template <bool SMEM_R>
__device__ __forceinline__ void subtree_m2m(const DevTree& T, int Ls, int m, int D, double* __restrict__ sm,
                                            const double* __restrict__ Rs, double* __restrict__ mcharge) {
    const int r = T.r;
    for (int d = D - 1; d >= 0; --d) {
        const int nb = 1 << (2 * d);
        const double* __restrict__ kids = sm + (size_t)lvoff(d + 1) * r;
        double* __restrict__ mine = sm + (size_t)lvoff(d) * r;
        const size_t gbase = (size_t)mc_node(T, Ls + d, m << (2 * d)) * r;  // a sub-tree's boxes of one level stay contiguous
        for (int o = threadIdx.x; o < nb * r; o += blockDim.x) {
            const int g = o / r, b = o - g * r;
            const double* __restrict__ src = kids + (size_t)g * 4 * r;
            const double* __restrict__ Rc = (SMEM_R ? Rs : T.R) + b;
            double a0 = 0, a1 = 0;
#pragma unroll 4
            for (int ca = 0; ca < 4 * r; ca += 2) {
                a0 = fma(SMEM_R ? Rc[(size_t)ca * r] : __ldg(Rc + (size_t)ca * r), src[ca], a0);
                a1 = fma(SMEM_R ? Rc[(size_t)(ca + 1) * r] : __ldg(Rc + (size_t)(ca + 1) * r), src[ca + 1], a1);
            }
            const double v = a0 + a1;
            mine[o] = v;
            mcharge[gbase + o] = v;
        }
        __syncthreads();
    }
}

// P2M of the 4^D leaves of one sub-tree (same scheme as p2m_kernel: LANES lanes per leaf, weights of a point computed
// once and parked in shared memory), then M2M up to the sub-tree root.
template <int NP, int LANES, bool SMEM_R>
__global__ void __launch_bounds__(256) up_leaf_kernel(DevTree T, DevGeom G, const double* __restrict__ q, double* __restrict__ mcharge,
                                                      int Ls, int D, int cta0) {
    extern __shared__ double sm[];
    __shared__ double tn[NP_MAX * NP_MAX];
    constexpr int UMAX = NP ? (NP * NP + LANES - 1) / LANES : (NP_MAX * NP_MAX + LANES - 1) / LANES;
    const int np = NP ? NP : T.np, r = np * np;
    const int m = blockIdx.x + cta0, nl = 1 << (2 * D);  // cta0: first sub-tree of this rank's share (sharded upward pass)
    const int grp = threadIdx.x / LANES, lane = threadIdx.x % LANES, ngrp = blockDim.x / LANES;
    const int stride = 2 * np + 1;
    double* __restrict__ mult = sm;                                   // [lvoff(D+1)][r]
    double* __restrict__ ws = sm + (size_t)lvoff(D + 1) * r + (size_t)grp * LANES * stride;
    double* __restrict__ Rs = sm + (size_t)lvoff(D + 1) * r + (size_t)ngrp * LANES * stride;
    const int n0 = T.uniform_npts;
    // the point data of the first round is requested before the table loads and the barrier (uniform leaves: its address
    // follows from the leaf index), so the DRAM round trip overlaps them
    double pf_x = 0, pf_y = 0, pf_q = 0;
    if (n0 && grp < nl && lane < n0) {
        const int off = (m * nl + grp) * n0 + lane;
        pf_x = G.sxn[off];
        pf_y = G.syn[off];
        pf_q = q[off];
    }
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) tn[i] = T.tnode[i];
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) Rs[i] = __ldg(T.R + i);
    __syncthreads();
    const unsigned mask = LANES == 32 ? 0xffffffffu : 0xffffu << (16 * ((threadIdx.x >> 4) & 1));
    for (int j0 = 0; j0 < nl; j0 += ngrp) {
        const int j = j0 + grp;  // nl and ngrp are powers of two: a group is either fully inside or fully outside
        if (j >= nl) break;
        int id, off, n;
        if (n0) {
            id = mc_node(T, Ls + D, m * nl + j);
            off = (m * nl + j) * n0;
            n = n0;
        } else {
            const int4 info = T.leaf_info[(size_t)m * nl + j];  // id, pt_off, npts
            id = mc_node(T, Ls + D, info.x - lvoff(Ls + D)); off = info.y; n = info.z;
        }
        double acc[UMAX];
#pragma unroll
        for (int u = 0; u < UMAX; ++u) acc[u] = 0;
        for (int k0 = 0; k0 < n; k0 += LANES) {
            const int cnt = min(LANES, n - k0);
            if (lane < cnt) {
                double w[2 * (NP ? NP : NP_MAX)];
                const bool pf = n0 && j0 == 0 && k0 == 0;
                cheb_weights<NP>(tn, np, pf ? pf_x : G.sxn[off + k0 + lane], w);
                cheb_weights<NP>(tn, np, pf ? pf_y : G.syn[off + k0 + lane], w + np);
#pragma unroll
                for (int mm = 0; mm < 2 * (NP ? NP : NP_MAX); ++mm)
                    if (mm < 2 * np) ws[lane * stride + mm] = w[mm];
                ws[lane * stride + 2 * np] = pf ? pf_q : q[off + k0 + lane];
            }
            __syncwarp(mask);
#pragma unroll
            for (int u = 0; u < UMAX; ++u) {
                const int mm = lane + LANES * u;
                if (mm < r) {
                    const int i = mm % np, jj = mm / np;
                    double a = acc[u];
                    for (int k = 0; k < cnt; ++k) a += ws[k * stride + i] * ws[k * stride + np + jj] * ws[k * stride + 2 * np];
                    acc[u] = a;
                }
            }
            __syncwarp(mask);
        }
#pragma unroll
        for (int u = 0; u < UMAX; ++u) {
            const int mm = lane + LANES * u;
            if (mm < r) {
                mult[(size_t)(lvoff(D) + j) * r + mm] = acc[u];
                mcharge[(size_t)id * r + mm] = acc[u];
            }
        }
    }
    __syncthreads();
    subtree_m2m<SMEM_R>(T, Ls, m, D, mult, Rs, mcharge);
}

template <bool SMEM_R>
__global__ void __launch_bounds__(256) up_mid_kernel(DevTree T, double* __restrict__ mcharge, int Ls, int D, int cta0) {
    extern __shared__ double sm[];
    const int r = T.r, m = blockIdx.x + cta0, nb = 1 << (2 * D);
    double* __restrict__ Rs = sm + (size_t)lvoff(D + 1) * r;
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) Rs[i] = __ldg(T.R + i);
    const double* __restrict__ src = mcharge + (size_t)mc_node(T, Ls + D, m * nb) * r;
    double* __restrict__ dst = sm + (size_t)lvoff(D) * r;
    for (int i = threadIdx.x; i < nb * r; i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    subtree_m2m<SMEM_R>(T, Ls, m, D, sm, Rs, mcharge);
}

// Scratch rows of a tree node (symmetric storage) summed in their fixed order; node_rows[node] = [first, last) row.
__device__ __forceinline__ double fold_rows_at(const double* __restrict__ scratch, int2 rr, int R, int t) {
    const double* __restrict__ base = scratch + t;
    double s = 0;
    for (int r = rr.x; r < rr.y; r += 8) {  // eight independent loads in flight per round
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = r + u < rr.y ? base[(size_t)(r + u) * R] : 0.0;
        s += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
    }
    return s;
}
__device__ __forceinline__ double fold_node(const DevCSR& m, int node, int t) {
    return fold_rows_at(m.scratch, m.node_rows[node], m.uniform_block, t);
}
// block operator: the same fold over the scratch sets of all terms, combined with their coefficients
__device__ __forceinline__ double fold_node_multi(const DevCSR& m, const FoldList& fl, const double* const* scr, int node, int t) {
    const int2 rr = m.node_rows[node];
    double s = 0;
    const int ns = fl.nscr < 0 ? fl.n : fl.nscr;
    for (int j = 0; j < ns; ++j) s = fma(fl.coef[j], fold_rows_at(scr[j], rr, m.uniform_block, t), s);
    return s;
}

// L2L inside shared memory: levels 0 .. D of the sub-tree rooted at node (Ls, m), in two steps.
// subtree_own: every node's own contribution goes to sm[(lvoff(d) + g)*r + a] - for the root the finished local
// expansion written by the chunk above (or, top: its own M2L sum, the first level with M2L lists has no parent term), for
// the others their M2L sum.  FOLD: M2L sums are still spread over scratch rows (symmetric storage), otherwise they sit
// in mpot.  None of these loads depends on another node, so they are all in flight together (no barrier inside).
template <bool FOLD, bool MULTI>
__device__ __forceinline__ void subtree_own(const DevTree& T, const DevCSR& m2l, const FoldList& fl, int Ls, int m, int D, bool top,
                                            double* __restrict__ sm, const double* __restrict__ mpot) {
    const int r = T.r, total = lvoff(D + 1) * r;
    for (int o = threadIdx.x; o < total; o += blockDim.x) {
        const int q = o / r, a = o - q * r;
        int d = 0;
        while (lvoff(d + 1) <= q) ++d;
        const int id = lvoff(Ls + d) + (m << (2 * d)) + (q - lvoff(d));
        if (FOLD && (d > 0 || top)) sm[o] = MULTI ? fold_node_multi(m2l, fl, fl.m2l_scr, id, a) : fold_node(m2l, id, a);
        else sm[o] = mpot[(size_t)id * r + a];
    }
}
// subtree_l2l: child += R[index] * parent level by level (bbfmm.h:1070-1071), one block barrier per level; starts with
// the barrier that publishes subtree_own's values.
template <bool SMEM_R>
__device__ __forceinline__ void subtree_l2l(const DevTree& T, int Ls, int m, int D, bool write_last, double* __restrict__ sm,
                                            const double* __restrict__ RTs, double* __restrict__ mpot) {
    const int r = T.r;
    __syncthreads();
    for (int d = 1; d <= D; ++d) {
        const int nb = 1 << (2 * d);
        const double* __restrict__ pars = sm + (size_t)lvoff(d - 1) * r;
        double* __restrict__ mine = sm + (size_t)lvoff(d) * r;
        const int first = lvoff(Ls + d) + (m << (2 * d));
        for (int o = threadIdx.x; o < nb * r; o += blockDim.x) {
            const int g = o / r, a = o - g * r;
            const size_t roff = (size_t)(g & 3) * r * r + a;  // children are consecutive: sibling position = child index
            const double* __restrict__ RT = (SMEM_R ? RTs : T.RT) + roff;
            const double* __restrict__ src = pars + (size_t)(g >> 2) * r;
            double a0 = 0, a1 = 0;
            int b = 0;
#pragma unroll 4
            for (; b + 2 <= r; b += 2) {
                a0 = fma(SMEM_R ? RT[(size_t)b * r] : __ldg(RT + (size_t)b * r), src[b], a0);
                a1 = fma(SMEM_R ? RT[(size_t)(b + 1) * r] : __ldg(RT + (size_t)(b + 1) * r), src[b + 1], a1);
            }
            if (b < r) a0 = fma(SMEM_R ? RT[(size_t)b * r] : __ldg(RT + (size_t)b * r), src[b], a0);
            const double v = mine[o] + (a0 + a1);
            mine[o] = v;
            if (write_last && d == D) mpot[(size_t)(first + g) * r + a] = v;
        }
        __syncthreads();
    }
}

template <bool SMEM_R, bool FOLD, bool MULTI>
__global__ void __launch_bounds__(256) down_mid_kernel(DevTree T, DevCSR m2l, double* __restrict__ mpot, int Ls, int D, int top,
                                                       const __grid_constant__ FoldList fl) {
    extern __shared__ double sm[];
    const int r = T.r;
    double* __restrict__ RTs = sm + (size_t)lvoff(D + 1) * r;
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) RTs[i] = __ldg(T.RT + i);
    subtree_own<FOLD, MULTI>(T, m2l, fl, Ls, blockIdx.x, D, top != 0, sm, mpot);
    subtree_l2l<SMEM_R>(T, Ls, blockIdx.x, D, true, sm, RTs, mpot);
}

// L2L down to the leaves of one sub-tree, then per point of the sub-tree (one contiguous slot range)
//   fmm = near(p) + sum_{i,j} Sx(p,i) Sy(p,j) local[leaf][j*np+i]                 L2P, bbfmm.h:1104-1113
//   mode 0: out = fmm;  mode 1: out = (fmm + C.f)/2pi;  mode 2: out = x - (fmm + C.f)/2pi
// near(p): P2P sum, already reduced in pot or still spread over the leaf's scratch rows (fold_p2p).
template <int NP, bool SMEM_R, bool FOLD, bool MULTI>
__global__ void __launch_bounds__(256, NP == 4 ? 4 : 2) down_leaf_kernel(DevTree T, DevGeom G, DevCSR m2l, DevCSR p2p, int fold_p2p,
                                                        const int* __restrict__ leaf_of_slot, int slot0, int n,
                                                        double* __restrict__ mpot, const double* __restrict__ pot,
                                                        const double* __restrict__ C, const double* __restrict__ f_ref,
                                                        const double* __restrict__ x, double* __restrict__ out, int mode, int Ls,
                                                        int D, int top, int cta0, const __grid_constant__ FoldList fl) {
    extern __shared__ double sm[];
    __shared__ double tn[NP_MAX * NP_MAX];
    const int np = NP ? NP : T.np, r = np * np, m = blockIdx.x + cta0;  // cta0: first sub-tree of this launch (chunked tail)
    const int root = lvoff(Ls) + m;
    const int n0 = T.uniform_npts;  // > 0: slot ranges follow from the leaf index, no table look-ups
    const int rb = n0 ? (m << (2 * D)) * n0 : T.pt_off[root], rn = n0 ? n0 << (2 * D) : T.npts[root];
    const int p_begin = max(rb, slot0), p_end = min(rb + rn, slot0 + n);
    if (p_begin >= p_end) return;  // no owned target below this root (multi-GPU partition)
    double* __restrict__ RTs = sm + (size_t)lvoff(D + 1) * r;
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) tn[i] = T.tnode[i];
    if (SMEM_R)
        for (int i = threadIdx.x; i < 4 * r * r; i += blockDim.x) RTs[i] = __ldg(T.RT + i);
    const int leaf0 = lvoff(Ls + D) + (m << (2 * D));
    // Everything of a point that does not depend on the local expansions (correction stencil, near-field fold,
    // interpolation weights) is loaded before the L2L levels, so those latencies overlap the L2L chain.
    struct Pre {
        double corr, near, wx[NP ? NP : NP_MAX], wy[NP ? NP : NP_MAX];
        int lo;
    };
    auto pre = [&](int p, Pre& P) {
        const int k = p - slot0;
        P.corr = 0;
        if (mode) {
            const int c2 = G.c2, sz = G.sz;
            const int tc = G.perm[p] / c2;
            const int ti = tc / sz, tj = tc - ti * sz;
            const int nterm = MULTI ? fl.n : 1;
            for (int jt = 0; jt < nterm; ++jt) {
                const double* __restrict__ Cj = MULTI ? fl.C[jt] : C;
                const double* __restrict__ fj = MULTI ? fl.f[jt] : f_ref;
                double cj = 0;
                for (int qq = 0; qq < c2; ++qq) {  // the 18 loads of one Legendre mode are issued together
                    double cv[9], fv[9];
#pragma unroll
                    for (int cs = 0; cs < 9; ++cs) {
                        const int si = ti + cs / 3 - 1, sj = tj + cs % 3 - 1;
                        const bool ok = si >= 0 && si < sz && sj >= 0 && sj < sz;
                        cv[cs] = ok ? Cj[(size_t)(cs * c2 + qq) * n + k] : 0.0;
                        fv[cs] = ok ? fj[(size_t)(si * sz + sj) * c2 + qq] : 0.0;
                    }
#pragma unroll
                    for (int cs = 0; cs < 9; ++cs) cj = fma(cv[cs], fv[cs], cj);
                }
                P.corr = MULTI ? fma(fl.coef[jt], cj, P.corr) : cj;
            }
        }
        const int id = n0 ? leaf0 + (p - rb) / n0 : leaf_of_slot[p];
        P.lo = id - leaf0;
        const int tl = n0 ? (p - rb) % n0 : p - T.pt_off[id];
        // fold_p2p == 2: the near field is added later by tail_kernel (split tail: this pass runs under the P2P stream)
        P.near = MULTI ? fold_node_multi(p2p, fl, fl.p2p_scr, id, tl) : fold_p2p == 2 ? 0.0 : fold_p2p ? fold_node(p2p, id, tl) : pot[p];
        cheb_weights<NP>(tn, np, G.sxn[p], P.wx);
        cheb_weights<NP>(tn, np, G.syn[p], P.wy);
    };
    const double* __restrict__ locals = sm + (size_t)lvoff(D) * r;
    auto post = [&](int p, const Pre& P) {
        const int k = p - slot0;
        const double* __restrict__ mp = locals + (size_t)P.lo * r;
        double acc = 0;
#pragma unroll
        for (int j = 0; j < (NP ? NP : NP_MAX); ++j) {
            if (j >= np) break;
            double row = 0;
#pragma unroll
            for (int i = 0; i < (NP ? NP : NP_MAX); ++i)
                if (i < np) row = fma(P.wx[i], mp[j * np + i], row);
            acc = fma(P.wy[j], row, acc);
        }
        double v = P.near + acc;
        if (mode) {
            v = (v + P.corr) * (M_1_PI / 2.0);
            if (mode == 2) v = x[k] - v;
        }
        out[k] = v;
    };
    __syncthreads();  // tn
    const int p_first = p_begin + threadIdx.x;
    Pre P0;
    subtree_own<FOLD, MULTI>(T, m2l, fl, Ls, m, D, top != 0, sm, mpot);
    if (p_first < p_end) pre(p_first, P0);
    subtree_l2l<SMEM_R>(T, Ls, m, D, false, sm, RTs, mpot);
    if (p_first < p_end) post(p_first, P0);
    for (int p = p_first + blockDim.x; p < p_end; p += blockDim.x) {
        Pre P;
        pre(p, P);
        post(p, P);
    }
}

// Split tail: what is left of the matvec per point once the leaf pass has written the far field (L2P of the finished local
// expansions) without the near field:  v = far + near(p) + C.f;  mode 1: out = v / 2pi;  mode 2: out = x - v / 2pi.
// near(p) = the leaf's P2P scratch rows folded in their fixed order (symmetric storage) or pot[p]; the stencil as in
// down_leaf_kernel.  A pure streaming kernel: one point per thread, every load independent of the others.
__global__ void __launch_bounds__(256) tail_kernel(DevTree T, DevGeom G, DevCSR p2p, int fold_p2p, const int* __restrict__ leaf_of_slot,
                                                   int slot0, int n, const double* __restrict__ far, const double* __restrict__ pot,
                                                   const double* __restrict__ C, const double* __restrict__ f_ref,
                                                   const double* __restrict__ x, double* __restrict__ out, int mode) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int p = slot0 + k;
    const int id = leaf_of_slot[p];
    const int tl = p - T.pt_off[id];
    const double near = fold_p2p ? fold_node(p2p, id, tl) : pot[p];
    double corr = 0;
    if (mode) {
        const int c2 = G.c2, sz = G.sz;
        const int tc = G.perm[p] / c2;
        const int ti = tc / sz, tj = tc - ti * sz;
        for (int qq = 0; qq < c2; ++qq) {
            double cv[9], fv[9];
#pragma unroll
            for (int cs = 0; cs < 9; ++cs) {
                const int si = ti + cs / 3 - 1, sj = tj + cs % 3 - 1;
                const bool ok = si >= 0 && si < sz && sj >= 0 && sj < sz;
                cv[cs] = ok ? C[(size_t)(cs * c2 + qq) * n + k] : 0.0;
                fv[cs] = ok ? f_ref[(size_t)(si * sz + sj) * c2 + qq] : 0.0;
            }
#pragma unroll
            for (int cs = 0; cs < 9; ++cs) corr = fma(cv[cs], fv[cs], corr);
        }
    }
    double v = near + far[k];
    if (mode) {
        v = (v + corr) * (M_1_PI / 2.0);
        if (mode == 2) v = x[k] - v;
    }
    out[k] = v;
}

void launch_tail(const DevTree& T, const DevGeom& G, const DevCSR& p2p, bool fold_p2p, const int* leaf_of_slot, int slot0, int n,
                 const double* far, const double* pot, const double* C, const double* f_ref, const double* x, double* out, int mode,
                 cudaStream_t st) {
    if (n > 0) tail_kernel<<<(n + 255) / 256, 256, 0, st>>>(T, G, p2p, fold_p2p ? 1 : 0, leaf_of_slot, slot0, n, far, pot, C, f_ref, x, out, mode);
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static bool smem_R(int r) { return (size_t)4 * r * r * sizeof(double) <= 36 * 1024; }  // r <= 33 (np <= 5)

// opt-in above 48 KB of dynamic shared memory: once per kernel instantiation, device and size (never while a stream is
// being captured into a graph: the first, uncaptured application of an operator has made the calls)
template <class K>
static void want_smem(K kernel, size_t bytes) {
    if (bytes <= 48 * 1024) return;
    static std::atomic<size_t> largest[64];  // per device: the attribute is per device
    int d = 0;
    cudaGetDevice(&d);
    d &= 63;
    if (bytes > largest[d].load()) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        largest[d].store(bytes);
    }
}

SubtreePlan make_subtree_plan(int L0, int deepest, int r) {
    SubtreePlan s;
    // levels per sub-tree: 2 keeps every launch wide (4^(deepest-2) CTAs at the leaves; the scratch-row folds of the
    // levels above are spread over 16x more CTAs than with 3) at the price of one more launch each way
    static const int denv = getenv("ANISO_SUBTREE_D") ? atoi(getenv("ANISO_SUBTREE_D")) : 2;
    const int dmax = std::max(1, std::min(denv, r <= 36 ? 3 : 2));
    s.leaf_D = std::min(dmax, deepest - L0);
    s.leaf_Ls = deepest - s.leaf_D;
    s.nmid = 0;
    for (int cur = s.leaf_Ls; cur > L0 && s.nmid < 8;) {  // bottom-up order
        const int D = std::min(dmax, cur - L0);
        s.mid_Ls[s.nmid] = cur - D;
        s.mid_D[s.nmid] = D;
        ++s.nmid;
        cur -= D;
    }
    return s;
}

template <int NP>
static void up_leaf_launch(const DevTree& T, const DevGeom& G, const SubtreePlan& sp, const double* q, double* mcharge, cudaStream_t st,
                           int part, int nparts) {
    const int r = T.r, D = sp.leaf_D, all = 1 << (2 * sp.leaf_Ls), ctas = all / nparts, cta0 = part * ctas;
    const bool sr = smem_R(r);
    const size_t bytes = ((size_t)lvoff(D + 1) * r + (size_t)256 * (2 * T.np + 1) + (sr ? (size_t)4 * r * r : 0)) * sizeof(double);
    const bool l16 = T.r <= 16 && T.max_leaf <= 16;
#define UP_LEAF(LANES, SR)                                                             \
    do {                                                                               \
        want_smem(up_leaf_kernel<NP, LANES, SR>, bytes);                               \
        up_leaf_kernel<NP, LANES, SR><<<ctas, 256, bytes, st>>>(T, G, q, mcharge, sp.leaf_Ls, D, cta0); \
    } while (0)
    if (l16 && sr) UP_LEAF(16, true);
    else if (l16) UP_LEAF(16, false);
    else if (sr) UP_LEAF(32, true);
    else UP_LEAF(32, false);
#undef UP_LEAF
}

int launch_up_tree(const DevTree& T, const DevGeom& G, const SubtreePlan& sp, const double* q, double* mcharge, cudaStream_t st,
                   cudaEvent_t leaf_done, int part, int nparts, const std::function<void()>* after_leaf) {
    switch (T.np) {
        case 4: up_leaf_launch<4>(T, G, sp, q, mcharge, st, part, nparts); break;
        case 6: up_leaf_launch<6>(T, G, sp, q, mcharge, st, part, nparts); break;
        case 8: up_leaf_launch<8>(T, G, sp, q, mcharge, st, part, nparts); break;
        default: up_leaf_launch<0>(T, G, sp, q, mcharge, st, part, nparts);
    }
    if (leaf_done) cudaEventRecord(leaf_done, st);
    if (after_leaf) (*after_leaf)();
    const int r = T.r;
    const bool sr = smem_R(r);
    for (int c = 0; c < sp.nmid; ++c) {
        const int D = sp.mid_D[c], ctas = (1 << (2 * sp.mid_Ls[c])) / nparts, cta0 = part * ctas;
        const size_t bytes = ((size_t)lvoff(D + 1) * r + (sr ? (size_t)4 * r * r : 0)) * sizeof(double);
        if (sr) {
            want_smem(up_mid_kernel<true>, bytes);
            up_mid_kernel<true><<<ctas, 256, bytes, st>>>(T, mcharge, sp.mid_Ls[c], D, cta0);
        } else {
            want_smem(up_mid_kernel<false>, bytes);
            up_mid_kernel<false><<<ctas, 256, bytes, st>>>(T, mcharge, sp.mid_Ls[c], D, cta0);
        }
    }
    return 1 + sp.nmid;
}

int launch_down_mid_tree(const DevTree& T, const DevCSR& m2l, bool fold, const SubtreePlan& sp, int L0, double* mpot, cudaStream_t st,
                         const FoldList* flp) {
    const int r = T.r;
    const bool sr = smem_R(r);
    static const FoldList none = FoldList();
    const FoldList& fl = flp ? *flp : none;
    for (int c = sp.nmid - 1; c >= 0; --c) {  // top-down
        const int D = sp.mid_D[c], Ls = sp.mid_Ls[c], ctas = 1 << (2 * Ls), top = Ls == L0;
        const size_t bytes = ((size_t)lvoff(D + 1) * r + (sr ? (size_t)4 * r * r : 0)) * sizeof(double);
#define DOWN_MID(SR, FO)                                                                       \
    do {                                                                                       \
        want_smem(down_mid_kernel<SR, FO, false>, bytes);                                      \
        down_mid_kernel<SR, FO, false><<<ctas, 256, bytes, st>>>(T, m2l, mpot, Ls, D, top, fl); \
    } while (0)
        if (flp && sr) {
            want_smem(down_mid_kernel<true, true, true>, bytes);
            down_mid_kernel<true, true, true><<<ctas, 256, bytes, st>>>(T, m2l, mpot, Ls, D, top, fl);
        } else if (flp) {
            want_smem(down_mid_kernel<false, true, true>, bytes);
            down_mid_kernel<false, true, true><<<ctas, 256, bytes, st>>>(T, m2l, mpot, Ls, D, top, fl);
        } else if (sr && fold) DOWN_MID(true, true);
        else if (sr) DOWN_MID(true, false);
        else if (fold) DOWN_MID(false, true);
        else DOWN_MID(false, false);
#undef DOWN_MID
    }
    return sp.nmid;
}

template <int NP>
static void down_leaf_launch(const DevTree& T, const DevGeom& G, const DevCSR& m2l, bool fold, const DevCSR& p2p, bool fold_p2p,
                             const SubtreePlan& sp, int L0, const int* leaf_of_slot, int slot0, int n, double* mpot, const double* pot,
                             const double* C, const double* f_ref, const double* x, double* out, int mode, cudaStream_t st,
                             const FoldList* flp, int cta0, int nctas, bool no_near) {
    static const FoldList none = FoldList();
    const FoldList& fl = flp ? *flp : none;
    const int r = T.r, D = sp.leaf_D, Ls = sp.leaf_Ls, ctas = nctas >= 0 ? nctas : 1 << (2 * Ls), top = Ls == L0;
    if (ctas == 0) return;
    const bool sr = smem_R(r);
    const size_t bytes = ((size_t)lvoff(D + 1) * r + (sr ? (size_t)4 * r * r : 0)) * sizeof(double);
#define DOWN_LEAF(SR, FO, MU)                                                                                                   \
    do {                                                                                                                        \
        want_smem(down_leaf_kernel<NP, SR, FO, MU>, bytes);                                                                     \
        down_leaf_kernel<NP, SR, FO, MU><<<ctas, 256, bytes, st>>>(T, G, m2l, p2p, no_near ? 2 : fold_p2p ? 1 : 0, leaf_of_slot, slot0, n, mpot, pot, \
                                                                   C, f_ref, x, out, mode, Ls, D, top, cta0, fl);               \
    } while (0)
    if (flp && sr) DOWN_LEAF(true, true, true);
    else if (flp) DOWN_LEAF(false, true, true);
    else if (sr && fold) DOWN_LEAF(true, true, false);
    else if (sr) DOWN_LEAF(true, false, false);
    else if (fold) DOWN_LEAF(false, true, false);
    else DOWN_LEAF(false, false, false);
#undef DOWN_LEAF
}

int launch_down_leaf_tree(const DevTree& T, const DevGeom& G, const DevCSR& m2l, bool fold, const DevCSR& p2p, bool fold_p2p,
                          const SubtreePlan& sp, int L0, const int* leaf_of_slot, int slot0, int n, double* mpot, const double* pot,
                          const double* C, const double* f_ref, const double* x, double* out, int mode, cudaStream_t st,
                          const FoldList* fl, int cta0, int nctas, bool no_near) {
    if (n == 0) return 0;
    switch (T.np) {
        case 4: down_leaf_launch<4>(T, G, m2l, fold, p2p, fold_p2p, sp, L0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode, st, fl, cta0, nctas, no_near); break;
        case 6: down_leaf_launch<6>(T, G, m2l, fold, p2p, fold_p2p, sp, L0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode, st, fl, cta0, nctas, no_near); break;
        case 8: down_leaf_launch<8>(T, G, m2l, fold, p2p, fold_p2p, sp, L0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode, st, fl, cta0, nctas, no_near); break;
        default: down_leaf_launch<0>(T, G, m2l, fold, p2p, fold_p2p, sp, L0, leaf_of_slot, slot0, n, mpot, pot, C, f_ref, x, out, mode, st, fl, cta0, nctas, no_near);
    }
    return 1;
}

}  // namespace an
