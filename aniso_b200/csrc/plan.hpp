// Host-side set-up of the aniso hot path: geometry tables, quadtree, interaction lists, and their
// flattening into the CSR / SoA arrays the CUDA kernels consume.  Everything here runs once per
// handle (the reference rebuilds the tree twice per matvec, KernelFactory.cpp:356,380 -> bbfmm.h:481).
//
// Reference behaviour restated (file:line into the reference tree):
//   Geometry ctor            Geometry.cpp:10-114      nodes, weights, interpolate, nearMapping
//   makeLegendreMatrix       Geometry.cpp:129-154
//   get_legendre_data/affine Quadrature.cpp:5418-22199
//   singPrecompute / duffy   KernelFactory.cpp:929-986, 863-927
//   tree::populate ...       bbfmm/bbfmm.h:176-447
//   kernel::initialize       bbfmm/bbfmm.h:476-505, 597-693
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace an {

struct TreeNode {
    double cx, cy, rx, ry;
    int parent = -1;
    int child[4] = {-1, -1, -1, -1};
    int level = 0, index = 0;   // index = which child of the parent (2*ybit + xbit)
    bool leaf = false, empty = false;
    int npts = 0;               // nSource == nTarget (sources and targets are the same node set)
    int pt_off = 0;             // first slot of this node's points in `perm` (DFS leaf order)
    std::vector<int> pts;       // point indices in the reference's insertion order (build-time only)
    std::vector<int> U, V, W, X;
};

// One block-sparse operator in the layout the streaming kernel reads:
//   y[row_off[g] + t] = sum_{e in [ent_ptr[g], ent_ptr[g+1])} sum_{s < ncol[e]}
//                         A[blk_off[e] + s*nrow[g] + t] * x[col_off[e] + s]
struct BlockCSR {
    std::vector<int> row_off, nrow;       // per row group (target node / target leaf)
    std::vector<int> group_node;          // tree node of each group
    std::vector<int> ent_ptr;             // size groups+1
    std::vector<int> col_off, ncol;       // per entry
    std::vector<int64_t> blk_off;         // per entry, in doubles
    std::vector<int> tgt_node, src_node;  // per entry (tree node ids, new numbering)
    std::vector<int> grow_ptr;            // groups+1: launch rows of group g are [grow_ptr[g], grow_ptr[g+1])
    std::vector<int> row_group;           // per launch row -> group
    int uniform_nrow = 0;                 // >0 when every group has this many rows (M2L: r)
    int uniform_block = 0;                // >0 when every block is uniform_block x uniform_block
    // symmetric storage (uniform blocks only): K(tgt A <- src B) = (-1)^id K(tgt B <- src A)^T, so only one block of
    // each pair is stored (the one with A < B); while streaming it the kernel also forms the transposed product and
    // drops it into the partner's scratch row `mirror_row`.  Scratch rows of group g: [srow_ptr[g], srow_ptr[g+1]),
    // the first one holding the group's own partial result.
    bool symmetric = false;
    std::vector<int> mirror_row;          // per entry: scratch row of the partner group, or -1
    std::vector<int> srow_ptr;            // groups+1
    int scratch_rows = 0;
    int64_t nnz = 0;                      // total doubles of operator storage
    int nrows_total = 0;                  // launch rows
};

constexpr int kDeepChunks = 4;

struct HostPlan {
    // parameters (Aniso ctor, Aniso.h:14-15)
    int sz = 0, deg = 0, K = 0, kernelSize = 0, sr = 0, np = 0, maxLevelArg = 0;
    double g = 0, dx = 0;
    int N = 0, ncell = 0, c2 = 0, rq = 0, nsing = 0, rank = 0;
    // geometry
    std::vector<double> nx, ny, w;            // nodes + weights (reference order)
    std::vector<double> glx, glw;             // volume GL rule, table order
    std::vector<double> sgx, sgw;             // singular rule mapped to [0,1]
    std::vector<double> sqrtw;                // sqrt of the in-cell tensor weights
    std::vector<double> interp;               // c2 x c2, row (n,k), col I
    std::vector<double> lnorm;                // legendreNorms (refined rule, Q3)
    std::vector<double> rqx, rqy, rqw;        // refined rule (16*c2 points)
    std::vector<double> nearMap;              // rq x c2
    std::vector<double> singX, singY, singW;  // c2 x nsing Duffy points in the reference cell
    // tree (level-major numbering, Morton order inside a level)
    std::vector<TreeNode> nodes;
    std::vector<int> level_ptr;               // nodes of level L are [level_ptr[L], level_ptr[L+1])
    int treeMaxLevel = 0;
    std::vector<int> perm;                    // perm[slot] = original point index
    std::vector<int> inv_perm;                // inv_perm[orig] = slot
    std::vector<int> leaf_of_slot;            // tree node owning each slot
    std::vector<double> cheb;                 // np Chebyshev nodes
    std::vector<double> tnode;                // T_k(t_m), m*np+k
    std::vector<double> R;                    // 4 x r x r transfer matrices, R[c][(a)*r + b]
    // generic point-set plan (bbfmm::kernel with a caller-supplied eval, bbfmm.h:451-505): arbitrary points instead of the
    // quadrature grid, leaf capacity independent of the Chebyshev order, no symmetric storage (nothing is known about eval)
    bool generic = false;
    int leaf_cap = 0;                         // > 0: leaf rule nSource <= leaf_cap (bbfmm.h:264); 0: rank (= np^2, KernelFactory.cpp:286)
    BlockCSR m2l, p2p;
    int deep_chunks = 1;                      // > 1: the owned deepest-level M2L rows form this many equal Morton ranges whose
                                              //      symmetric pairs are kept by the earlier range (see symmetrize)
    std::vector<int> leaves;                  // non-empty leaves (new numbering)
    // statistics in the order of the golden fixtures
    long stats[9] = {0};
};

// Gauss-Legendre rule of degree n in the reference's table order.  Returns false for n outside [1,128]
// (the reference prints "The degree is not implemented." and leaves the table empty).
bool gl_table(int n, double* x, double* w);

// Builds everything above.  Throws std::runtime_error on invalid arguments.  flatten = false leaves m2l / p2p empty for a
// caller that flattens a leaf range itself (sharded handles).
void build_plan(HostPlan& p, int sz, int deg, int K, double g, int sr, int np, int maxLevel, bool flatten = true);

// kernel::initialize for an arbitrary point set with sources == targets (bbfmm.h:476-505 -> tree::populate 176-209): tree,
// lists, Chebyshev operators and the flattened (unsymmetric) block-CSR operators; no quadrature geometry.
void build_plan_points(HostPlan& p, int np, int n, const double* x, const double* y, int leaf_cap, int maxLevel);

// Tabulates eval(source, target) into operator storage laid out like `b` (kernelEvalIndex / kernelEvalChebyshev,
// bbfmm.h:770-804): A[blk_off[e] + s * nrow + t] = eval(source s of entry e, target t).  cheb: the np x np Chebyshev nodes of
// the two boxes (flat index j*np + i, y slow), else the points of the two leaves.  OpenMP over the entries; eval must be
// callable from several threads at once, as under the reference's OpenMP tasks.
typedef double (*plan_eval_fn)(void* ctx, double sx, double sy, double tx, double ty);
void tabulate_operator(const HostPlan& p, const BlockCSR& b, bool cheb, plan_eval_fn eval, void* ctx, int threads, double* A);

// sigma interpolation (KernelFactory.cpp:212-227 / 988-1005): coeff[c*c2 + r] = sum_j interp[r][j] sqrtw[j] h[c*c2+j]
void interpolate_field(const HostPlan& p, const double* h, double* coeff);

// Contiguous ownership of the target leaves for `nranks` devices: leaves in Morton/DFS order are
// split into ranges of near-equal operator bytes.  out_begin/out_end are indices into p.leaves.
void partition_leaves(const HostPlan& p, int nranks, std::vector<int>& begin, std::vector<int>& end);

// (Re)builds p.m2l / p.p2p for the targets owned by leaves [leaf_begin, leaf_end) of p.leaves: P2P rows for those
// leaves, M2L rows for them and all their ancestors.  build_plan() calls it for the full range.
void flatten_lists(HostPlan& p, int leaf_begin, int leaf_end);

}  // namespace an
