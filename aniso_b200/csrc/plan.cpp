// Host set-up for the aniso hot path (see plan.hpp).  Plain C++17, no CUDA, usable on a CPU-only box.
#include "plan.hpp"

#include <omp.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <functional>
#include <numeric>
#include <stdexcept>

namespace an {

static constexpr double kEps = 1e-12;  // bbfmm/utils.h:46 (EPS)

// ---------------------------------------------------------------------------------------------
// Gauss-Legendre tables.  The reference ships 58-digit decimal tables (Quadrature.cpp:5418-22191);
// correctly rounded doubles are reproduced here by Newton iteration carried to binary128, and the
// rows are emitted in the table's (non-monotone) order: centre outwards in (-x,+x) pairs, with the
// three irregular degrees 6, 7 and 9 listed explicitly (quirk Q9).
// ---------------------------------------------------------------------------------------------
typedef __float128 quad_t;

static void legendre_pair(int n, quad_t z, quad_t& pn, quad_t& pnm1) {
    quad_t a = 1, b = 0;
    for (int j = 1; j <= n; ++j) {
        quad_t c = b;
        b = a;
        a = ((2 * j - 1) * z * b - (j - 1) * c) / j;
    }
    pn = a;
    pnm1 = b;
}

static void gl_sorted(int n, std::vector<quad_t>& x, std::vector<quad_t>& w) {
    x.resize(n);
    w.resize(n);
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 0; i < n; ++i) {
        long double zl = -cosl(pi * (i + 0.75L) / (n + 0.5L));
        for (int it = 0; it < 100; ++it) {
            long double a = 1, b = 0;
            for (int j = 1; j <= n; ++j) {
                long double c = b;
                b = a;
                a = ((2 * j - 1) * zl * b - (j - 1) * c) / j;
            }
            long double step = a / (n * (zl * a - b) / (zl * zl - 1));
            zl -= step;
            if (fabsl(step) < 1e-19L) break;
        }
        quad_t z = zl, pn, pm, dp = 0;
        for (int it = 0; it < 3; ++it) {
            legendre_pair(n, z, pn, pm);
            dp = n * (z * pn - pm) / (z * z - 1);
            if (it < 2) z -= pn / dp;
        }
        x[i] = z;
        w[i] = 2 / ((1 - z * z) * dp * dp);
    }
    if (n & 1) x[n / 2] = 0;
}

static void table_order(int n, std::vector<int>& ord) {
    ord.resize(n);
    static const int o6[6] = {4, 1, 2, 3, 0, 5};
    static const int o7[7] = {3, 4, 2, 1, 5, 0, 6};
    static const int o9[9] = {4, 1, 7, 0, 8, 3, 5, 2, 6};
    const int* irregular = n == 6 ? o6 : n == 7 ? o7 : n == 9 ? o9 : nullptr;
    if (irregular) {
        std::copy(irregular, irregular + n, ord.begin());
        return;
    }
    int k = 0;
    int lo = (n & 1) ? n / 2 - 1 : n / 2 - 1, hi = (n & 1) ? n / 2 + 1 : n / 2;
    if (n & 1) ord[k++] = n / 2;
    while (lo >= 0) {
        ord[k++] = lo--;
        ord[k++] = hi++;
    }
}

bool gl_table(int n, double* x, double* w) {
    if (n < 1 || n > 128) return false;
    std::vector<quad_t> xs, ws;
    std::vector<int> ord;
    gl_sorted(n, xs, ws);
    table_order(n, ord);
    for (int s = 0; s < n; ++s) {
        x[s] = (double)xs[ord[s]];
        w[s] = (double)ws[ord[s]];
    }
    return true;
}

// libstdc++ std::tr1::legendre (tr1/legendre_function.tcc:82-114), used via bbfmm/utils.h:81
static double legendre_tr1(int l, double x) {
    if (x == 1.0) return 1.0;
    if (x == -1.0) return (l & 1) ? -1.0 : 1.0;
    if (l == 0) return 1.0;
    if (l == 1) return x;
    double pm2 = 1.0, pm1 = x, p = 0;
    for (int ll = 2; ll <= l; ++ll) {
        p = 2.0 * x * pm1 - pm2 - (x * pm1 - pm2) / (double)ll;
        pm2 = pm1;
        pm1 = p;
    }
    return p;
}

// Geometry::makeLegendreMatrix (Geometry.cpp:129-154): rows (n,k), normalised; norms go to p.lnorm
static void legendre_matrix(HostPlan& p, std::vector<double>& M, int npts, const double* x, const double* y,
                            const double* w) {
    const int deg = p.deg, c2 = p.c2;
    M.assign((size_t)c2 * npts, 0.0);
    for (int n = 0; n < deg; ++n)
        for (int k = 0; k < deg; ++k) {
            double* row = &M[(size_t)(n * deg + k) * npts];
            double ss = 0;
            for (int I = 0; I < npts; ++I) {
                row[I] = legendre_tr1(n, x[I]) * legendre_tr1(k, y[I]) * std::sqrt(w[I]);
                ss += row[I] * row[I];
            }
            const double nrm = std::sqrt(ss);
            p.lnorm[n * deg + k] = nrm;  // the second call (refined rule) overwrites the first: Q3
            for (int I = 0; I < npts; ++I) row[I] /= nrm;
        }
}

static void build_geometry(HostPlan& p) {
    const int sz = p.sz, deg = p.deg, c2 = p.c2, rq = p.rq, sr = p.sr;
    p.glx.resize(deg);
    p.glw.resize(deg);
    gl_table(deg, p.glx.data(), p.glw.data());
    p.sgx.resize(sr);
    p.sgw.resize(sr);
    gl_table(sr, p.sgx.data(), p.sgw.data());
    for (int k = 0; k < sr; ++k) {  // affine(), Quadrature.cpp:22194-22199
        p.sgw[k] /= 2.0;
        p.sgx[k] += 1.0;
        p.sgx[k] /= 2.0;
    }
    std::vector<double> qx(c2), qy(c2), qw(c2);
    p.sqrtw.resize(c2);
    for (int a = 0; a < deg; ++a)
        for (int b = 0; b < deg; ++b) {
            qx[a * deg + b] = p.glx[a];
            qy[a * deg + b] = p.glx[b];
            qw[a * deg + b] = p.glw[a] * p.glw[b];
            p.sqrtw[a * deg + b] = std::sqrt(qw[a * deg + b]);
        }
    p.nx.resize(p.N);
    p.ny.resize(p.N);
    p.w.resize(p.N);
    const double dx = p.dx;
    for (int i = 0; i < sz; ++i)
        for (int j = 0; j < sz; ++j)
            for (int k = 0; k < c2; ++k) {
                const int id = (i * sz + j) * c2 + k;
                p.nx[id] = (0.5 + i) * dx + 0.5 * (qx[k]) * dx;
                p.ny[id] = (0.5 + j) * dx + 0.5 * qy[k] * dx;
                p.w[id] = qw[k] * 0.25 * (dx * dx);
            }
    p.lnorm.assign(c2, 0.0);
    legendre_matrix(p, p.interp, c2, qx.data(), qy.data(), qw.data());
    // two rounds of 4-way refinement in the reference's push order (Geometry.cpp:79-107)
    std::vector<double> rx = qx, ry = qy, rw = qw;
    for (int round = 0; round < 2; ++round) {
        std::vector<double> tx, ty, tw;
        tx.reserve(rx.size() * 4);
        ty.reserve(rx.size() * 4);
        tw.reserve(rx.size() * 4);
        static const int sx[4] = {+1, +1, -1, -1}, sy[4] = {+1, -1, +1, -1};
        for (size_t id = 0; id < rx.size(); ++id)
            for (int q = 0; q < 4; ++q) {
                tx.push_back((rx[id] + sx[q]) / 2.0);
                ty.push_back((ry[id] + sy[q]) / 2.0);
                tw.push_back(rw[id] / 4.0);
            }
        rx.swap(tx);
        ry.swap(ty);
        rw.swap(tw);
    }
    p.rqx = rx;
    p.rqy = ry;
    p.rqw = rw;
    std::vector<double> refinements;
    legendre_matrix(p, refinements, rq, rx.data(), ry.data(), rw.data());
    // nearMapping = refinements^T * interpolate (Geometry.cpp:109-113), stored [a][q]
    p.nearMap.assign((size_t)rq * c2, 0.0);
    for (int a = 0; a < rq; ++a)
        for (int q = 0; q < c2; ++q) {
            double s = 0;
            for (int r = 0; r < c2; ++r) s += refinements[(size_t)r * rq + a] * p.interp[(size_t)r * c2 + q];
            p.nearMap[(size_t)a * c2 + q] = s;
        }
    // singPrecompute (KernelFactory.cpp:929-986): 8 Duffy triangles around each in-cell node
    const int ns = p.nsing;
    p.singX.assign((size_t)c2 * ns, 0.0);
    p.singY.assign((size_t)c2 * ns, 0.0);
    p.singW.assign((size_t)c2 * ns, 0.0);
    for (int a = 0; a < deg; ++a)
        for (int b = 0; b < deg; ++b) {
            const int t = a * deg + b;
            const double x = p.glx[a], y = p.glx[b];
            const double tri[8][6] = {{x, y, 1., y, 1., 1.},   {x, y, 1., 1, x, 1.},    {x, y, x, 1., -1., 1.},
                                      {x, y, -1., 1., -1., y}, {x, y, -1., y, -1., -1.}, {x, y, -1., -1, x, -1.},
                                      {x, y, x, -1., 1., -1.}, {x, y, 1., -1, 1., y}};
            for (int k = 0; k < 8; ++k) {  // duffy_transform, KernelFactory.cpp:863-927
                const double* v = tri[k];
                const double a11 = v[2] - v[0], a12 = v[4] - v[2], a21 = v[3] - v[1], a22 = v[5] - v[3];
                const double det = a11 * a22 - a12 * a21;
                size_t o = (size_t)t * ns + (size_t)k * sr * sr;
                for (int ri = 0; ri < sr; ++ri)
                    for (int ci = 0; ci < sr; ++ci, ++o) {
                        const double u = p.sgx[ri], vv = p.sgx[ci], ww = p.sgw[ri] * p.sgw[ci];
                        const double zx = u, zy = u * vv, zw = ww * u;
                        p.singX[o] = a11 * zx + a12 * zy + v[0];
                        p.singY[o] = a21 * zx + a22 * zy + v[1];
                        p.singW[o] = det * zw;
                    }
            }
        }
}

void interpolate_field(const HostPlan& p, const double* h, double* coeff) {
    const int c2 = p.c2;
    for (int c = 0; c < p.ncell; ++c)
        for (int r = 0; r < c2; ++r) {
            double s = 0;
            for (int j = 0; j < c2; ++j) s += p.interp[(size_t)r * c2 + j] * (p.sqrtw[j] * h[(size_t)c * c2 + j]);
            coeff[(size_t)c * c2 + r] = s;
        }
}

// ---------------------------------------------------------------------------------------------
// Quadtree and interaction lists (bbfmm.h:176-447).  Built in creation order, then renumbered
// level-major (Morton order inside a level) so that every level is one contiguous index range.
// ---------------------------------------------------------------------------------------------
namespace {

struct TreeBuilder {
    HostPlan& p;
    std::vector<TreeNode> t;
    int rank, maxLevel, deepest = 0;
    explicit TreeBuilder(HostPlan& plan) : p(plan), rank(plan.leaf_cap > 0 ? plan.leaf_cap : plan.rank), maxLevel(plan.maxLevelArg) {}

    void split(int id) {  // assignChildren, bbfmm.h:250-317
        if (t[id].npts == 0) {
            t[id].leaf = t[id].empty = true;
            return;
        }
        if (t[id].npts <= rank || t[id].level == maxLevel) {
            t[id].leaf = true;
            deepest = std::max(deepest, t[id].level);
            return;
        }
        int kids[4];
        for (int c = 0; c < 4; ++c) {
            TreeNode n;
            const TreeNode& par = t[id];
            n.parent = id;
            n.level = par.level + 1;
            n.index = c;
            n.cx = par.cx + ((c & 1) - 0.5) * par.rx;
            n.cy = par.cy + (((c >> 1) & 1) - 0.5) * par.ry;
            n.rx = par.rx * 0.5;
            n.ry = par.ry * 0.5;
            kids[c] = (int)t.size();
            t.push_back(n);
            t[id].child[c] = kids[c];
        }
        for (int idx : t[id].pts) {
            const int xbit = p.nx[idx] < t[id].cx ? 0 : 1, ybit = p.ny[idx] < t[id].cy ? 0 : 1;
            TreeNode& q = t[kids[2 * ybit + xbit]];
            q.pts.push_back(idx);
            q.npts++;
        }
        for (int c = 0; c < 4; ++c) split(kids[c]);
    }

    int locate(double px, double py) const {  // findNode, bbfmm.h:416-429
        int id = 0;
        for (;;) {
            const TreeNode& n = t[id];
            if (std::fabs(n.cx - px) < kEps && std::fabs(n.cy - py) < kEps) return id;
            if (n.leaf) return id;
            const int xbit = n.cx > px ? 0 : 1, ybit = n.cy > py ? 0 : 1;
            id = n.child[2 * ybit + xbit];
        }
    }

    bool adjacent(int a, int b) const {  // isAdjacent, bbfmm.h:431-447
        const TreeNode &A = t[a], &B = t[b];
        const double dx = std::fabs(A.cx - B.cx), dy = std::fabs(A.cy - B.cy);
        const double rx = std::fabs(A.rx + B.rx), ry = std::fabs(A.ry + B.ry);
        const bool rdx = rx >= dx - kEps, rdy = ry >= dy - kEps;
        return ((std::fabs(dx - rx) < kEps) && rdy) || ((std::fabs(dy - ry) < kEps) && rdx);
    }

    static void add_unique(std::vector<int>& v, int x) {
        if (std::find(v.begin(), v.end(), x) == v.end()) v.push_back(x);
    }

    void lists(int id, double minx, double miny, double maxx, double maxy) {  // buildNode, bbfmm.h:334-413
        TreeNode& n = t[id];
        if (n.parent >= 0) {
            const TreeNode& par = t[n.parent];
            const double hx = n.rx, hy = n.ry, x0 = par.cx - hx, y0 = par.cy - hy;
            for (int xi = -2; xi < 4; ++xi)
                for (int yi = -2; yi < 4; ++yi) {
                    const double cx = x0 + 2 * xi * hx, cy = y0 + 2 * yi * hy;
                    const bool inside = cx <= maxx + kEps && cy <= maxy + kEps && cx >= minx - kEps && cy >= miny - kEps;
                    const bool self = std::fabs(cx - n.cx) < kEps && std::fabs(cy - n.cy) < kEps;
                    if (!inside || self) continue;
                    const int cur = locate(cx, cy);
                    const bool adj = adjacent(id, cur);
                    if (t[cur].level < n.level) {
                        if (adj) {
                            if (t[cur].leaf) add_unique(n.U, cur);
                        } else
                            add_unique(n.X, cur);
                    }
                    if (t[cur].level == n.level) {
                        if (!adj)
                            add_unique(n.V, cur);
                        else if (n.leaf) {
                            std::vector<int> queue{cur};
                            for (size_t h = 0; h < queue.size(); ++h) {
                                const int f = queue[h];
                                if (!adjacent(f, id))
                                    add_unique(n.W, f);
                                else if (t[f].leaf)
                                    add_unique(n.U, f);
                                else
                                    for (int c = 0; c < 4; ++c) queue.push_back(t[f].child[c]);
                            }
                        }
                    }
                }
        }
        if (n.leaf) add_unique(n.U, id);  // bbfmm.h:405-407
    }
};

}  // namespace

static void build_tree(HostPlan& p) {
    TreeBuilder B(p);
    TreeNode root;
    double xmax = p.nx[0], xmin = p.nx[0], ymax = p.ny[0], ymin = p.ny[0];
    for (int i = 0; i < p.N; ++i) {  // getCenterRadius, bbfmm.h:231-248 (bounding box of the nodes, Q4)
        xmax = std::max(xmax, p.nx[i]);
        xmin = std::min(xmin, p.nx[i]);
        ymax = std::max(ymax, p.ny[i]);
        ymin = std::min(ymin, p.ny[i]);
    }
    root.cx = (xmax + xmin) / 2.0;
    root.cy = (ymax + ymin) / 2.0;
    root.rx = (xmax - xmin) / 2.0;
    root.ry = (ymax - ymin) / 2.0;
    root.npts = p.N;
    root.pts.resize(p.N);
    for (int i = 0; i < p.N; ++i) root.pts[i] = i;
    B.t.reserve((size_t)p.N / 2 + 16);
    B.t.push_back(root);
    B.split(0);
    const double minx = B.t[0].cx - B.t[0].rx, miny = B.t[0].cy - B.t[0].ry;
    const double maxx = B.t[0].cx + B.t[0].rx, maxy = B.t[0].cy + B.t[0].ry;
    const int nn = (int)B.t.size();
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < nn; ++i) B.lists(i, minx, miny, maxx, maxy);
    p.treeMaxLevel = B.deepest;

    // statistics (same order as tests/golden tree_stats)
    long leaves = 0, su = 0, sv = 0, sw = 0, sx = 0, pairs = 0, mx = 0;
    for (const TreeNode& n : B.t) {
        if (n.leaf) {
            ++leaves;
            mx = std::max<long>(mx, n.npts);
            for (int e : n.U) pairs += (long)B.t[e].npts * n.npts;
            for (int e : n.W) pairs += (long)B.t[e].npts * n.npts;
        }
        su += n.U.size();
        sv += n.V.size();
        sw += n.W.size();
        sx += n.X.size();
    }
    long st[9] = {nn, leaves, B.deepest, su, sv, sw, sx, pairs, mx};
    std::copy(st, st + 9, p.stats);

    // renumber: breadth-first with children visited 0..3 => Morton order per level
    std::vector<int> order, newid(nn, -1);
    order.reserve(nn);
    p.level_ptr.assign(1, 0);
    std::vector<int> cur{0};
    while (!cur.empty()) {
        std::vector<int> next;
        for (int id : cur) {
            newid[id] = (int)order.size();
            order.push_back(id);
            if (!B.t[id].leaf)
                for (int c = 0; c < 4; ++c) next.push_back(B.t[id].child[c]);
        }
        p.level_ptr.push_back((int)order.size());
        cur.swap(next);
    }
    p.nodes.resize(nn);
    for (int k = 0; k < nn; ++k) {
        TreeNode n = std::move(B.t[order[k]]);
        if (n.parent >= 0) n.parent = newid[n.parent];
        for (int c = 0; c < 4; ++c)
            if (n.child[c] >= 0) n.child[c] = newid[n.child[c]];
        for (auto* L : {&n.U, &n.V, &n.W, &n.X})
            for (int& e : *L) e = newid[e];
        p.nodes[k] = std::move(n);
    }
    // point slots: depth-first leaf order, so that every subtree owns one contiguous slot range
    p.perm.clear();
    p.perm.reserve(p.N);
    p.leaves.clear();
    std::function<void(int)> dfs = [&](int id) {
        TreeNode& n = p.nodes[id];
        n.pt_off = (int)p.perm.size();
        if (n.leaf) {
            if (!n.empty) p.leaves.push_back(id);
            for (int idx : n.pts) p.perm.push_back(idx);
        } else
            for (int c = 0; c < 4; ++c) dfs(n.child[c]);
    };
    dfs(0);
    p.inv_perm.assign(p.N, -1);
    p.leaf_of_slot.assign(p.N, -1);
    for (int s = 0; s < p.N; ++s) p.inv_perm[p.perm[s]] = s;
    for (int id : p.leaves)
        for (int k = 0; k < p.nodes[id].npts; ++k) p.leaf_of_slot[p.nodes[id].pt_off + k] = id;
}

// ---------------------------------------------------------------------------------------------
// Chebyshev machinery of kernel::initialize (bbfmm.h:476-505, 597-693)
// ---------------------------------------------------------------------------------------------
static double cheb_weight(const HostPlan& p, double x, int m) {  // S(x,m) = (2 sum_k T_k(x) T_k(t_m) - 1)/np
    const int np = p.np;
    double s = -1.0, t0 = 1.0, t1 = x;
    for (int k = 0; k < np; ++k) {
        double tk = k == 0 ? 1.0 : x;
        if (k >= 2) {
            tk = 2.0 * x * t1 - t0;
            t0 = t1;
            t1 = tk;
        }
        s += 2.0 * tk * p.tnode[(size_t)m * np + k];
    }
    return s / np;
}

static void build_chebyshev(HostPlan& p) {
    const int np = p.np, r = p.rank;
    p.cheb.resize(np);
    p.tnode.resize((size_t)np * np);
    for (int i = 0; i < np; ++i) p.cheb[i] = -std::cos((i + 0.5) * M_PI / np);
    for (int m = 0; m < np; ++m)
        for (int k = 0; k < np; ++k)
            p.tnode[(size_t)m * np + k] =
                k == 0 ? 1.0 : k == 1 ? p.cheb[m] : 2.0 * p.cheb[m] * p.tnode[(size_t)m * np + k - 1] - p.tnode[(size_t)m * np + k - 2];
    // R[c][(i*np+j), (k*np+l)] = S(child node i along y, k) * S(child node j along x, l)   (bbfmm.h:661-693)
    p.R.assign((size_t)4 * r * r, 0.0);
    for (int c = 0; c < 4; ++c) {
        const int bx = c & 1, by = (c >> 1) & 1;
        for (int i = 0; i < np; ++i)
            for (int j = 0; j < np; ++j) {
                const double yi = (by ? 0.5 : -0.5) + 0.5 * p.cheb[i];
                const double xj = (bx ? 0.5 : -0.5) + 0.5 * p.cheb[j];
                for (int k = 0; k < np; ++k)
                    for (int l = 0; l < np; ++l)
                        p.R[((size_t)c * r + (i * np + j)) * r + (k * np + l)] = cheb_weight(p, yi, k) * cheb_weight(p, xj, l);
            }
    }
}

// ---------------------------------------------------------------------------------------------
// Flattening of the V/X (M2L) and U/W (P2P) lists into block-CSR form
// ---------------------------------------------------------------------------------------------
static void finish_csr(BlockCSR& b) {
    const int groups = (int)b.nrow.size();
    b.grow_ptr.assign(groups + 1, 0);
    for (int g = 0; g < groups; ++g) b.grow_ptr[g + 1] = b.grow_ptr[g] + b.nrow[g];
    b.nrows_total = b.grow_ptr[groups];
    b.uniform_block = groups ? b.nrow[0] : 0;
    for (int g = 0; g < groups; ++g)
        if (b.nrow[g] != b.uniform_block) b.uniform_block = 0;
    for (int nc : b.ncol)
        if (nc != b.uniform_block) b.uniform_block = 0;
    b.row_group.resize(b.nrows_total);
    for (int g = 0; g < groups; ++g)
        for (int k = b.grow_ptr[g]; k < b.grow_ptr[g + 1]; ++k) b.row_group[k] = g;
}

// Rewrites a uniform-block CSR into symmetric storage (see BlockCSR::symmetric).
// chunk (optional): node -> chunk number.  A pair whose two nodes sit in different chunks is kept by the node of the LOWER
// chunk: when the rows of the chunks are streamed one launch after the other, everything a target of chunk c receives -
// its own products and the transposed products other targets drop into its scratch rows - has been written once the
// launches of chunks <= c are done, so the consumer of chunk c (L2L / L2P of its sub-trees) can start while the later
// chunks are still streaming.
static void symmetrize(BlockCSR& b, const std::vector<int>* chunk = nullptr) {
    const int groups = (int)b.nrow.size(), R = b.uniform_block;
    if (!groups || R < 4 || R > 64 || getenv("ANISO_NO_SYM")) return;  // R = 16, 64: stream.cu bsr_apply_sym_kernel; else symg
    // partner[e]: the entry (tgt s <- src t) of the same pair as entry e = (tgt t <- src s), or -1.  Found by a binary search
    // in the entries of s's group sorted by source node (a hash map of all entries took 1 s of the 1.9 s plan build at C5).
    const int ne = (int)b.col_off.size();
    int max_node = 0;
    for (int e = 0; e < ne; ++e) max_node = std::max(max_node, std::max(b.tgt_node[e], b.src_node[e]));
    std::vector<int> group_of(max_node + 1, -1), by_src(ne), partner(ne, -1);
    for (int g = 0; g < groups; ++g) group_of[b.group_node[g]] = g;
#pragma omp parallel for schedule(static)
    for (int g = 0; g < groups; ++g) {
        int* lo = by_src.data() + b.ent_ptr[g];
        int* hi = by_src.data() + b.ent_ptr[g + 1];
        std::iota(lo, hi, b.ent_ptr[g]);
        std::sort(lo, hi, [&b](int x, int y) { return b.src_node[x] < b.src_node[y]; });
    }
#pragma omp parallel for schedule(static)
    for (int e = 0; e < ne; ++e) {
        const int t = b.tgt_node[e], s = b.src_node[e], g = group_of[s];
        if (t == s || g < 0) continue;
        const int* lo = by_src.data() + b.ent_ptr[g];
        const int* hi = by_src.data() + b.ent_ptr[g + 1];
        const int* it = std::lower_bound(lo, hi, t, [&b](int x, int node) { return b.src_node[x] < node; });
        if (it != hi && b.src_node[*it] == t) partner[e] = *it;
    }
    // Of a pair (A,B) the block is kept by A when A+B is even and A < B, or A+B is odd and A > B: every target keeps
    // about half of its pairs, so the warps of the streaming kernel stay balanced.  An entry is a mirror (dropped)
    // when its partner exists and the partner owns the pair.
    auto owns = [chunk](int t, int s) {
        if (chunk && (*chunk)[t] != (*chunk)[s]) return (*chunk)[t] < (*chunk)[s];
        return ((t + s) & 1) ? t > s : t < s;
    };
    std::vector<char> is_mirror(b.col_off.size(), 0);
    std::vector<int> row_of_entry(b.col_off.size(), -1);
    b.srow_ptr.assign(groups + 1, 0);
    for (int g = 0; g < groups; ++g) {
        int rows = 1;
        for (int e = b.ent_ptr[g]; e < b.ent_ptr[g + 1]; ++e) {
            const int t = b.tgt_node[e], s = b.src_node[e];
            if (partner[e] >= 0 && !owns(t, s)) {
                is_mirror[e] = 1;
                row_of_entry[e] = b.srow_ptr[g] + rows++;
            }
        }
        b.srow_ptr[g + 1] = b.srow_ptr[g] + rows;
    }
    b.scratch_rows = b.srow_ptr[groups];
    BlockCSR n;
    const size_t kept = (size_t)ne - (size_t)(b.scratch_rows - groups);
    for (auto* v : {&n.col_off, &n.ncol, &n.tgt_node, &n.src_node, &n.mirror_row}) v->reserve(kept);
    n.blk_off.reserve(kept);
    n.ent_ptr.reserve(groups + 1);
    n.ent_ptr.push_back(0);
    for (int g = 0; g < groups; ++g) {
        for (int e = b.ent_ptr[g]; e < b.ent_ptr[g + 1]; ++e) {
            if (is_mirror[e]) continue;
            const int t = b.tgt_node[e], s = b.src_node[e];
            const int mrow = partner[e] >= 0 && owns(t, s) ? row_of_entry[partner[e]] : -1;
            n.col_off.push_back(b.col_off[e]);
            n.ncol.push_back(b.ncol[e]);
            n.blk_off.push_back(n.nnz);
            n.tgt_node.push_back(t);
            n.src_node.push_back(s);
            n.mirror_row.push_back(mrow);
            n.nnz += (int64_t)R * R;
        }
        n.ent_ptr.push_back((int)n.col_off.size());
    }
    b.ent_ptr.swap(n.ent_ptr);
    b.col_off.swap(n.col_off);
    b.ncol.swap(n.ncol);
    b.blk_off.swap(n.blk_off);
    b.tgt_node.swap(n.tgt_node);
    b.src_node.swap(n.src_node);
    b.mirror_row.swap(n.mirror_row);
    b.nnz = n.nnz;
    b.symmetric = true;
}

void flatten_lists(HostPlan& p, int leaf_begin, int leaf_end) {
    const int r = p.rank, nn = (int)p.nodes.size();
    std::vector<char> need(nn, 0);
    for (int k = leaf_begin; k < leaf_end; ++k)
        for (int id = p.leaves[k]; id >= 0 && !need[id]; id = p.nodes[id].parent) need[id] = 1;
    BlockCSR m, q;
    m.uniform_nrow = r;
    {
        size_t nm = 0, nq = 0;
        for (int id = 0; id < nn; ++id)
            if (need[id]) nm += p.nodes[id].V.size() + p.nodes[id].X.size();
        for (int k = leaf_begin; k < leaf_end; ++k) nq += p.nodes[p.leaves[k]].U.size() + p.nodes[p.leaves[k]].W.size();
        for (auto* v : {&m.col_off, &m.ncol, &m.tgt_node, &m.src_node}) v->reserve(nm);
        m.blk_off.reserve(nm);
        for (auto* v : {&q.col_off, &q.ncol, &q.tgt_node, &q.src_node}) v->reserve(nq);
        q.blk_off.reserve(nq);
    }
    m.ent_ptr.push_back(0);
    q.ent_ptr.push_back(0);
    for (int id = 0; id < nn; ++id) {  // level-major order: M2L blocks of one level are contiguous
        const TreeNode& n = p.nodes[id];
        if (!need[id] || n.empty || n.parent < 0) continue;
        int cnt = 0;
        for (const std::vector<int>* L : {&n.V, &n.X})  // X handled as M2L: Q6
            for (int s : *L) {
                if (p.nodes[s].empty) continue;
                m.col_off.push_back(s * r);
                m.ncol.push_back(r);
                m.blk_off.push_back(m.nnz);
                m.tgt_node.push_back(id);
                m.src_node.push_back(s);
                m.nnz += (int64_t)r * r;
                ++cnt;
            }
        if (!cnt) continue;
        m.row_off.push_back(id * r);
        m.group_node.push_back(id);
        m.nrow.push_back(r);
        m.ent_ptr.push_back((int)m.col_off.size());
    }
    for (int k = leaf_begin; k < leaf_end; ++k) {
        const int id = p.leaves[k];
        const TreeNode& n = p.nodes[id];
        for (const std::vector<int>* L : {&n.U, &n.W})  // W handled as P2P: Q6
            for (int s : *L) {
                if (p.nodes[s].empty) continue;
                q.col_off.push_back(p.nodes[s].pt_off);
                q.ncol.push_back(p.nodes[s].npts);
                q.blk_off.push_back(q.nnz);
                q.tgt_node.push_back(id);
                q.src_node.push_back(s);
                q.nnz += (int64_t)n.npts * p.nodes[s].npts;
            }
        q.row_off.push_back(n.pt_off);
        q.group_node.push_back(id);
        q.nrow.push_back(n.npts);
        q.ent_ptr.push_back((int)q.col_off.size());
    }
    finish_csr(m);
    finish_csr(q);
    // Chunks of the deepest level (complete trees: all owned leaves on the deepest level, in Morton order): kDeepChunks
    // equal ranges of the owned leaves, aligned to whole sub-trees of the leaf pass (64 leaves).  Other nodes: chunk 0.
    p.deep_chunks = 1;
    std::vector<int> chunk(nn, 0);
    {
        const int nown = leaf_end - leaf_begin, deepest = p.treeMaxLevel;
        bool ok = nown > 0 && nown % (kDeepChunks * 64) == 0 && leaf_begin % 64 == 0 && getenv("ANISO_DEEP_CHUNKS");  // opt-in: measured slower
        for (int k = leaf_begin; ok && k < leaf_end; ++k) ok = p.nodes[p.leaves[k]].level == deepest && p.leaves[k] == p.leaves[leaf_begin] + (k - leaf_begin);
        if (ok) {
            p.deep_chunks = kDeepChunks;
            for (int k = leaf_begin; k < leaf_end; ++k) chunk[p.leaves[k]] = (int)((int64_t)(k - leaf_begin) * kDeepChunks / nown);
        }
    }
    if (!p.generic) {  // the symmetry K_AB = (-1)^k K_BA^T belongs to the aniso kernel family, not to an arbitrary eval
        symmetrize(m, p.deep_chunks > 1 ? &chunk : nullptr);
        symmetrize(q);
    }
    p.m2l = std::move(m);
    p.p2p = std::move(q);
}

void partition_leaves(const HostPlan& p, int nranks, std::vector<int>& begin, std::vector<int>& end) {
    // cost of a leaf = operator doubles streamed for it per apply: its P2P tiles plus its own M2L blocks
    // (ancestors' M2L blocks are shared by the ranks below them and are small in total).
    const int nl = (int)p.leaves.size(), r = p.rank;
    std::vector<double> cost(nl);
    double total = 0;
    for (int k = 0; k < nl; ++k) {
        const TreeNode& n = p.nodes[p.leaves[k]];
        double c = 0;
        for (int s : n.U) c += (double)n.npts * p.nodes[s].npts;
        for (int s : n.W) c += (double)n.npts * p.nodes[s].npts;
        c += (double)(n.V.size() + n.X.size()) * r * r;
        cost[k] = c;
        total += c;
    }
    begin.assign(nranks, 0);
    end.assign(nranks, 0);
    {
        // uniform leaves that divide evenly: equal shares (the per-leaf cost differs only along the domain boundary), so
        // that every rank owns the same number of slots and the per-matvec exchange is one ncclAllGather
        bool same = nl > 0 && nl % nranks == 0;
        for (int q = 1; same && q < nl; ++q) same = p.nodes[p.leaves[q]].npts == p.nodes[p.leaves[0]].npts;
        if (same) {
            for (int rk = 0; rk < nranks; ++rk) {
                begin[rk] = rk * (nl / nranks);
                end[rk] = (rk + 1) * (nl / nranks);
            }
            return;
        }
    }
    int k = 0;
    double acc = 0;
    for (int rk = 0; rk < nranks; ++rk) {
        begin[rk] = k;
        const double target = total * (rk + 1) / nranks;
        while (k < nl && (rk == nranks - 1 || acc + 0.5 * cost[k] <= target)) acc += cost[k++];
        end[rk] = k;
    }
}

void build_plan_points(HostPlan& p, int np, int n, const double* x, const double* y, int leaf_cap, int maxLevel) {
    if (n < 1) throw std::runtime_error("at least one point is required");
    if (np < 1) throw std::runtime_error("np must be >= 1");
    if (np > 16) throw std::runtime_error("np > 16 (rank > 256) is not supported by the device kernels");
    if (leaf_cap < 1) throw std::runtime_error("leaf capacity (rank) must be >= 1");
    for (int i = 0; i < n; ++i)
        if (!std::isfinite(x[i]) || !std::isfinite(y[i])) throw std::runtime_error("point coordinates must be finite");
    p = HostPlan();
    p.generic = true;
    p.K = 1; p.kernelSize = 1; p.np = np; p.maxLevelArg = maxLevel; p.leaf_cap = leaf_cap;
    p.N = n;
    p.rank = np * np;  // kernel::rank = nChebyshev^2 (bbfmm.h:489)
    p.nx.assign(x, x + n);
    p.ny.assign(y, y + n);
    p.w.assign(n, 1.0);
    build_tree(p);
    build_chebyshev(p);
    flatten_lists(p, 0, (int)p.leaves.size());
    for (TreeNode& nd : p.nodes) std::vector<int>().swap(nd.pts);
}

void tabulate_operator(const HostPlan& p, const BlockCSR& b, bool cheb, plan_eval_fn eval, void* ctx, int threads, double* A) {
    const int np = p.np, ne = (int)b.col_off.size();
    std::vector<int> group_of_entry(ne, 0);
    for (size_t g = 0; g + 1 < b.ent_ptr.size(); ++g)
        for (int e = b.ent_ptr[g]; e < b.ent_ptr[g + 1]; ++e) group_of_entry[e] = (int)g;
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads > 0 ? threads : omp_get_max_threads())
    for (int e = 0; e < ne; ++e) {
        const TreeNode &T = p.nodes[b.tgt_node[e]], &S = p.nodes[b.src_node[e]];
        const int nrow = b.nrow[group_of_entry[e]], ncol = b.ncol[e];
        double* out = A + b.blk_off[e];
        for (int s = 0; s < ncol; ++s) {
            double sx, sy;
            if (cheb) {  // box nodes at (c.x + r.x t_i, c.y + r.y t_j), flat index j*np + i (bbfmm.h:699-705, 787-800)
                sx = S.cx + S.rx * p.cheb[s % np];
                sy = S.cy + S.ry * p.cheb[s / np];
            } else {
                const int idx = p.perm[S.pt_off + s];
                sx = p.nx[idx];
                sy = p.ny[idx];
            }
            for (int t = 0; t < nrow; ++t) {
                double tx, ty;
                if (cheb) {
                    tx = T.cx + T.rx * p.cheb[t % np];
                    ty = T.cy + T.ry * p.cheb[t / np];
                } else {
                    const int idx = p.perm[T.pt_off + t];
                    tx = p.nx[idx];
                    ty = p.ny[idx];
                }
                out[(size_t)s * nrow + t] = eval(ctx, sx, sy, tx, ty);
            }
        }
    }
}

void build_plan(HostPlan& p, int sz, int deg, int K, double g, int sr, int np, int maxLevel, bool flatten) {
    if (sz < 1) throw std::runtime_error("geometry_size must be >= 1");
    if (deg < 1 || deg > 128) throw std::runtime_error("The degree is not implemented.");  // Quadrature.cpp:22189-22191
    if (sr < 1 || sr > 128) throw std::runtime_error("The degree is not implemented.");
    if (K < 1) throw std::runtime_error("kernel_size must be >= 1");
    if (np < 1) throw std::runtime_error("np must be >= 1");
    if (np > 16) throw std::runtime_error("np > 16 (rank > 256) is not supported by the device kernels");
    if ((double)sz * sz * deg * deg > 2.0e9) throw std::runtime_error("number of nodes exceeds index_t (int)");
    p = HostPlan();
    p.sz = sz; p.deg = deg; p.K = K; p.kernelSize = 2 * K - 1; p.g = g; p.sr = sr; p.np = np; p.maxLevelArg = maxLevel;
    p.dx = 1.0 / sz;
    p.ncell = sz * sz;
    p.c2 = deg * deg;
    p.N = p.ncell * p.c2;
    p.rq = 16 * p.c2;      // DEFAULT_REFINE_LEVEL 2, Geometry.h:27
    p.nsing = 8 * sr * sr;
    p.rank = np * np;      // KernelFactory.cpp:284-286 passes np*np as the leaf capacity
    build_geometry(p);
    build_tree(p);
    build_chebyshev(p);
    if (flatten) flatten_lists(p, 0, (int)p.leaves.size());
    for (TreeNode& n : p.nodes) std::vector<int>().swap(n.pts);  // build-time only
}

}  // namespace an
