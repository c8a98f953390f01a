// Device-side view of one aniso handle and the launch wrappers of kernels.cu.  sm_100a only.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <functional>

namespace an {

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: a kernel instantiation opts in once per device a
// handle runs on (handles of one process may sit on different GPUs, from different host threads)
struct PerDeviceOnce {
    std::atomic<unsigned long long> seen{0};
    bool first() {
        int d = 0;
        cudaGetDevice(&d);
        const unsigned long long bit = 1ull << (d & 63);
        return (seen.fetch_or(bit) & bit) == 0;
    }
};

enum PartMask { PART_REAL = 1, PART_IMAG = 2, PART_BOTH = 3 };  // realParts / imagParts of KernelFactory.h:22-23

// sigma_t interpolant (KernelFactory.cpp:174-207): per-cell Legendre coefficients already divided by the
// legendreNorms, evaluated at GLOBAL coordinates (quirk Q1).
struct SigmaGrid {
    int sz, deg, c2;
    double h;              // cell width 1/sz
    const double* chat;    // [ncell][c2]
    const double* glx;     // [deg] table order
    const double* glw;     // [deg]
    // deg == 1 only (sigma piecewise constant): running integrals of sigma along each grid line, paired with sigma,
    //   ty[cx*sz + j] = { h * sum_{j' < j} sigma[cx][j'], sigma[cx][j] }     (along y inside column cx)
    //   tx[cy*sz + i] = { h * sum_{i' < i} sigma[i'][cy], sigma[i][cy] }     (along x inside row cy)
    const double2* ty = nullptr;
    const double2* tx = nullptr;
    // differences of neighbouring lines, dty[cx*sz + j] = ty[cx*sz + j] - ty[(cx+1)*sz + j] (dtx alike): what a ray adds when
    // it crosses from line cx to cx + 1
    const double2* dty = nullptr;
    const double2* dtx = nullptr;
};

// block-CSR operator (see plan.hpp BlockCSR); all pointers are device memory
struct DevCSR {
    int groups = 0, entries = 0, nrows_total = 0, uniform_nrow = 0, uniform_block = 0;
    int64_t nnz = 0;
    int *row_off = nullptr, *nrow = nullptr, *ent_ptr = nullptr, *col_off = nullptr, *ncol = nullptr;
    int64_t* blk_off = nullptr;
    int *tgt_node = nullptr, *src_node = nullptr, *grow_ptr = nullptr, *row_group = nullptr;
    int symmetric = 0, scratch_rows = 0;                       // symmetric storage, see plan.hpp
    int swz = 0;  // 16 x 16 blocks stored bank-swizzled: element (s, t) at s*16 + (t ^ 4*(s%4)), see stream.cu (DMMA kernel)
    int *mirror_row = nullptr, *srow_ptr = nullptr;
    int* group_of_node = nullptr;                              // tree node -> group (or -1)
    int2* node_rows = nullptr;                                 // tree node -> [first, last) scratch row (empty without a group)
    double* scratch = nullptr;                                 // [scratch_rows][uniform_block]
};

struct DevTree {
    int nnodes = 0, np = 0, r = 0, nlevels = 0;
    double *cx = nullptr, *cy = nullptr, *rx = nullptr, *ry = nullptr;
    int *parent = nullptr, *index = nullptr, *child = nullptr;  // child: [nnodes][4]
    int *pt_off = nullptr, *npts = nullptr;
    unsigned char* flags = nullptr;                             // bit0 leaf, bit1 empty
    double *cheb = nullptr, *tnode = nullptr, *R = nullptr;     // R: [4][r][r]
    double* RT = nullptr;                                       // R transposed per child: [4][b][a]
    int* leaves = nullptr;                                      // all non-empty leaves (DFS order)
    int nleaves = 0, max_leaf = 0;
    int uniform_npts = 0;                                       // > 0: every leaf holds this many points (slot = leaf * npts + k)
    int* level_ptr = nullptr;                                   // nodes of level L: [level_ptr[L], level_ptr[L+1])
    int4* leaf_info = nullptr;                                  // per leaf: node id, pt_off, npts, 0
    int* child0 = nullptr;                                      // first child (children are consecutive) or -1
    // Rank-major multipole storage of a sharded handle on a complete tree (pack_logP >= 0): the nodes rank q owns on the
    // levels pack_L0 .. deepest (the q-th of 2^pack_logP equal Morton ranges of every level) are stored back to back,
    // pack_tot nodes per rank, so that the multipoles of all levels travel in ONE in-place ncclAllGather.  Only the
    // multipole array (mcharge) uses this numbering; see mc_node().
    int pack_logP = -1, pack_L0 = 0, pack_tot = 0;
};
// position (in nodes) of node i of level L in the multipole array; complete trees only when packed
__host__ __device__ __forceinline__ int mc_node(const DevTree& T, int L, int i) {
    const int off = ((1 << (2 * L)) - 1) / 3;  // nodes above level L
    if (T.pack_logP < 0) return off + i;
    const int sh = 2 * L - T.pack_logP;        // 4^L / P nodes of this level per rank
    const int base = (off - ((1 << (2 * T.pack_L0)) - 1) / 3) >> T.pack_logP;
    return (i >> sh) * T.pack_tot + base + (i & ((1 << sh) - 1));
}

// cell-based correction tables (Geometry / singPrecompute data, uploaded once)
struct DevGeom {
    int sz, deg, c2, rq, nsing, N;
    double dx;
    const double *px, *py, *pw;          // slot order
    const double *sxn, *syn;             // box-local coordinates (x - cx)/rx, (y - cy)/ry of each slot in its leaf
    const int *perm, *inv_perm;          // slot <-> reference node index
    const double *sqrtw, *interp, *lnorm, *nearMap, *rqx, *rqy, *rqw, *singX, *singY, *singW;
};

// ---- launch wrappers (kernels.cu) -------------------------------------------------------------
// K1/K2 evaluator: fills A for every entry of `m`.  cheb != 0: Chebyshev nodes of the two boxes (M2L),
// else the points of the two nodes (P2P).  A[blk_off[e] + s*nrow + t] = eval(source s, target t).
void launch_fill(const DevCSR& m, const DevTree& T, const DevGeom& G, const SigmaGrid& S, int id, int part, int cheb,
                 double* A, cudaStream_t st);
// same for kernel id > 0, part BOTH, from the id-0 operators A0 (one multiply by cos(id * theta) per entry, no ray march)
void launch_fill_from_base(const DevCSR& m, const DevTree& T, const DevGeom& G, int id, int cheb, const double* A0, double* A,
                           cudaStream_t st);
// y[row] (=) sum over entries: the HBM-streaming cached apply (K3 / K4)
// sign = (-1)^id: parity of the kernel family under source<->target exchange (symmetric storage only)
// reduce = false (symmetric storage only): leave the per-target sums in the scratch rows for the consumer to fold
void launch_bsr_apply(const DevCSR& m, const double* A, const double* x, double* y, double sign, bool reduce, cudaStream_t st);
void launch_bsr_reduce(const DevCSR& m, double* y, cudaStream_t st);
// cache-free P2P (kernel::run / downPass, bbfmm.h:507-532, 863-946): evaluates the U/W tiles of kernel (id, part) while
// applying them to q; leaves its results where launch_bsr_apply(..., reduce = false) would (scratch rows under symmetric
// storage, y otherwise).  Leaves of at most 64 points.
bool p2p_onfly_ok(const DevCSR& m, int max_leaf);
void launch_p2p_onfly(const DevCSR& m, const DevGeom& G, const SigmaGrid& S, int id, int part, const double* q, double* y, double sign,
                      cudaStream_t st);
// the same operator applied to nrhs vectors x + r*xstride in one pass (symmetric storage, 16x16 / 64x64 blocks); results
// stay in the scratch sets scr + r*sstride (laid out like m.scratch).  Returns the number of launches.
int launch_bsr_apply_multi(const DevCSR& m, const double* A, const double* x, size_t xstride, int nrhs, double* scr, size_t sstride,
                           double sign, cudaStream_t st);
// q = f.*w, f = x or sig_s.*x; f_ref (optional) = f scattered into reference node order
void launch_scale(int n, const double* x, const double* sig_s, const double* w, const int* perm, double* f_ref, double* q,
                  cudaStream_t st);
void launch_p2m(const DevTree& T, const DevGeom& G, const double* q, double* mcharge, cudaStream_t st);
void launch_m2m(const DevTree& T, int first, int count, double* mcharge, cudaStream_t st);
// fold: the M2L result of each node is still spread over its scratch rows (symmetric storage) and is reduced inside
void launch_l2l(const DevTree& T, const DevCSR& m2l, bool fold, int first, int count, double* mpot, cudaStream_t st);
// levels [Llo, Lhi] (at most 1024 nodes each) in one single-CTA launch
void launch_m2m_top(const DevTree& T, int Lhi, int Llo, double* mcharge, cudaStream_t st);
void launch_l2l_top(const DevTree& T, const DevCSR& m2l, bool fold, int Llo, int Lhi, double* mpot, cudaStream_t st);
// The terms of one output mode of the multi-mode block operator (aniso.m:121-156): out = sum_j coef[j] * mapping_{k_j}(u_{a_j}).
// The streams of (k_j, a_j) left their per-target sums in the scratch sets m2l_scr[j] / p2p_scr[j]; the downward pass folds
// and combines them, so one downward pass serves all 2K-1 terms.
struct FoldList {
    static constexpr int MAXF = 16;
    int n = 0;
    double coef[MAXF];
    const double* m2l_scr[MAXF];
    const double* p2p_scr[MAXF];
    const double* C[MAXF];  // fused correction stencil of kernel id k_j
    const double* f[MAXF];  // unscaled source of input mode a_j, reference node order
    int nscr = -1;          // terms that carry scratch sets (the first nscr; -1: all); the others only apply their stencil
};

// operators of several kernel ids applied in one pass (block operator on the tensor cores, stream.cu): id k multiplies the
// nrhs vectors x[k] + r * xstride with A[k] (all ids share one block-CSR structure) and everything accumulates into the
// nrhs result sets
struct IdList {
    static constexpr int MAXI = 15;
    int n = 0;
    const double* A[MAXI];
    const double* x[MAXI];
    double sign[MAXI];
};
// dst = ca * a + cb * b (b may be null), n entries of one length
struct LinList {
    static constexpr int MAXL = 96;
    int n = 0;
    double* dst[MAXL];
    const double* a[MAXL];
    const double* b[MAXL];
    double ca[MAXL], cb[MAXL];
};
void launch_lincomb(const LinList& L, size_t len, cudaStream_t st);
// requires m.swz (16 x 16 blocks, swizzled storage), nrhs <= 8; one launch
void launch_bsr_apply_ids(const DevCSR& m, const IdList& ids, size_t xstride, int nrhs, double* scr, size_t sstride, cudaStream_t st);

// ---- sub-tree passes for complete uniform trees (subtree.cu) ----
struct SubtreePlan {
    int leaf_Ls = 0, leaf_D = 0;          // sub-trees holding the leaves: roots at level leaf_Ls, leaf_D levels below
    int nmid = 0, mid_Ls[8], mid_D[8];    // the chunks above, bottom-up
};
SubtreePlan make_subtree_plan(int first_m2l_level, int deepest, int r);
// P2M + M2M of every level down to first_m2l_level; leaf_done (optional) is recorded once the leaf multipoles exist.
// part / nparts: only the sub-trees of share `part` out of `nparts` equal Morton ranges (sharded upward pass; nparts must
// divide the number of sub-tree roots of every chunk, i.e. 4^first_m2l_level).  Return value: kernels launched.
// after_leaf (optional): called once the leaf-level kernel has been launched (e.g. to push the leaf multipoles to the peers)
int launch_up_tree(const DevTree& T, const DevGeom& G, const SubtreePlan& sp, const double* q, double* mcharge, cudaStream_t st,
                   cudaEvent_t leaf_done, int part = 0, int nparts = 1, const std::function<void()>* after_leaf = nullptr);
// fl != nullptr: multi-term folds (block operator; needs fold == fold_p2p == true); C / f_ref / pot are then unused
int launch_down_mid_tree(const DevTree& T, const DevCSR& m2l, bool fold, const SubtreePlan& sp, int L0, double* mpot, cudaStream_t st,
                         const FoldList* fl = nullptr);
int launch_down_leaf_tree(const DevTree& T, const DevGeom& G, const DevCSR& m2l, bool fold, const DevCSR& p2p, bool fold_p2p,
                          const SubtreePlan& sp, int L0, const int* leaf_of_slot, int slot0, int n, double* mpot, const double* pot,
                          const double* C, const double* f_ref, const double* x, double* out, int mode, cudaStream_t st,
                          const FoldList* fl = nullptr, int cta0 = 0, int nctas = -1,  // [cta0, cta0 + nctas): sub-trees of one chunk
                          bool no_near = false);  // no_near: far field only (mode 0); launch_tail adds the rest
// split tail: out = finalise(far + near-field fold + correction stencil), see tail_kernel (subtree.cu)
void launch_tail(const DevTree& T, const DevGeom& G, const DevCSR& p2p, bool fold_p2p, const int* leaf_of_slot, int slot0, int n,
                 const double* far, const double* pot, const double* C, const double* f_ref, const double* x, double* out, int mode,
                 cudaStream_t st);
// L2P + tail of the matvec for slots [slot0, slot0+n):  fmm = pot[p] + L2P(p)
//   mode 0: out = fmm;  mode 1: out = (fmm + C.f)/2pi;  mode 2: out = x - (fmm + C.f)/2pi   (C: fused stencil, f_ref: full unscaled source in reference order)
void launch_l2p(const DevTree& T, const DevGeom& G, const DevCSR& p2p, bool fold_p2p, const int* leaf_of_slot, int slot0, int n,
                const double* mpot,
                const double* pot, const double* C, const double* f_ref, const double* x, double* out, int mode,
                cudaStream_t st);
// correction stencils: C[(slotcell*c2+q)*nown + (p-slot0)], slotcell in 0..8 ((dr+1)*3+(dc+1))
void launch_build_stencils(const DevGeom& G, int id, int slot0, int nown, double* c7, double* c8, double* c9, cudaStream_t st);
// c7/c8 act on the scaled source, c9 ([c2][nown]) on the Legendre coefficients of the unscaled source;
// call = fused stencil on the unscaled source
void launch_combine_stencils(const DevGeom& G, int slot0, int nown, const double* c7, const double* c8, const double* c9,
                             double* call, cudaStream_t st);
void launch_sing_apply(const DevGeom& G, int slot0, int nown, const double* c9, const double* coeff, double* acc, cudaStream_t st);
// acc[p-slot0] (+)= sum stencil * f (f: full vector, slot order); accumulate 0: overwrite, 1: add
void launch_stencil_apply(const DevGeom& G, int slot0, int nown, const double* C, const double* f, double* acc, int accumulate,
                          cudaStream_t st);
// out = (fmm + corr) / 2pi, optionally y = x - that (GMRES operator of main.cpp:123-136)
void launch_finalize(int n, const double* fmm, const double* corr, const double* x, double* out, int subtract_from_x,
                     cudaStream_t st);
void launch_gather(int n, const double* src, const int* idx, double* dst, cudaStream_t st);  // dst[i] = src[idx[i]]
void launch_scatter(int n, const double* src, const int* idx, double* dst, cudaStream_t st);  // dst[idx[i]] = src[i]
void launch_add3(long n, const double* a, const double* b, const double* c, double* out, cudaStream_t st);
// sigma interpolation on device: chat[c][r] = (sum_j interp[r][j] sqrtw[j] h[c*c2+j]) / (lnorm ? lnorm[r] : 1), h in REFERENCE order
void launch_interpolate(const DevGeom& G, const double* h_ref, const double* lnorm_or_null, double* coeff, cudaStream_t st);
// (re)builds S.ty / S.tx from S.chat (deg == 1)
void launch_sigma_prefix(const SigmaGrid& S, double2* ty, double2* tx, cudaStream_t st);
// probes for parity tests
void launch_probe_line_integral(const SigmaGrid& S, int n, const double* pts4, double* out, cudaStream_t st);
void launch_probe_kernel(const SigmaGrid& S, int id, int part, int n, const double* pts4, double* out, cudaStream_t st);

// ---- GMRES vector ops (gmres.cu) ----------------------------------------------------------------
struct DotWork {
    double* partial = nullptr;  // [blocks]
    unsigned int* counter = nullptr;
    int blocks = 0;
};
void launch_dot(int n, const double* a, const double* b, double* result, DotWork& wk, cudaStream_t st);
// w -= h[0] * v, then result = w . u (u == nullptr: w . w): one MGS step in one pass
void launch_axpy_dot(int n, const double* h, const double* v, double* w, const double* u, double* result, DotWork& wk,
                     cudaStream_t st);
// CGS2 pieces: out[k] = w . V_k for k < nv (V_k = V + k*ldv; partial: nv * wk.blocks doubles of scratch), and
// w -= sum_k h[k] V_k with an optional ||w||^2 of the result
void launch_multi_dot(int n, int nv, const double* w, const double* V, size_t ldv, double* out, double* partial, DotWork& wk,
                      cudaStream_t st);
bool cgs_pass_ok(int n, int nv, const double* V, size_t ldv, const double* w);
// Small all-reduce over peer memory (peer.cu) fused into the epilogue of a reduction kernel: the block that holds the final
// sums pushes them to a slot of its own on every rank, waits for the slots of all ranks and adds them in rank order.
// P == 0: no exchange.  Filled by peer_red_args().
struct PeerRed {
    int P = 0, me = 0, cap = 0;
    double* red[16];
    unsigned long long* flags[16];
    const double* my_red = nullptr;
    const unsigned long long* my_flags = nullptr;
    unsigned long long epoch = 0;
    int* err = nullptr;
};
// pr (optional, modes 0 - 3): the outputs are all-reduced over the ranks inside the kernel.  Mode 4 with vnext: the updated
// vector is not written back to w but, scaled to unit length (its norm follows from h_in), to vnext (and vnext2).
void launch_cgs_pass(int mode, int n, int nv, const double* h_in, const double* V, size_t ldv, double* w, double* out, double* partial,
                     DotWork& wk, cudaStream_t st, const PeerRed* pr = nullptr, double* vnext = nullptr, double* vnext2 = nullptr);
void launch_multi_axpy(int n, int nv, const double* h, const double* V, size_t ldv, double* w, double* norm2_out, double* partial,
                       DotWork& wk, cudaStream_t st);
// w -= h[0] * v  (h read from device memory)
void launch_axpy_neg_dev(int n, const double* h, const double* v, double* w, cudaStream_t st);
void launch_axpy(int n, double alpha, const double* x, double* y, cudaStream_t st);
// out = in * (1/ *norm), or 1/sqrt(*norm) when `squared` (norm holds w . w)
// out2 (optional): a second copy of the result (the fixed input buffer of the captured operator graph)
void launch_scale_inv_dev(int n, const double* norm, const double* in, double* out, int squared, cudaStream_t st,
                          double* out2 = nullptr);
void launch_sqrt1(double* v, cudaStream_t st);
void launch_sub(int n, const double* a, const double* b, double* out, cudaStream_t st);  // out = a - b
// scalar part of Arnoldi step i on the device (gmres.cpp:123-144): column i of H from the projections h1 (+ h2 for CGS2) and
// norm2 = w . w, previous rotations applied, new rotation generated, s updated; report[0] = |s(i+1)| / normb,
// report[1] = 1 when that is below tol.  beta is read at i == 0.
// norm2 == nullptr: w . w = h2[i + 1] - sum_k h2[k]^2 (the two-reduction form of CGS2)
void launch_givens_step(int i, const double* h1, const double* h2, const double* norm2, const double* beta, double* H, int ldh,
                        double* cs, double* sn, double* s, double normb, double tol, double* report, cudaStream_t st);
// yneg = -(H(0:k,0:k)^-1 s(0:k)) (Update, gmres.cpp:12-24)
void launch_gm_solve(int k, const double* H, int ldh, const double* s, double* yneg, cudaStream_t st);

}  // namespace an
