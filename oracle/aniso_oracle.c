/* ============================================================================================
 * TEST INFRASTRUCTURE ONLY - CPU ORACLE.  Plain-C restatement of the lowrank/aniso hot path
 * (integral-operator matvec `mapping` + restarted GMRES), written from the algorithm, not from
 * the reference's code.  Every function cites the reference file:line it follows.
 *
 * PINNED: this restatement is checked in tests/test_oracle.py against
 *   (1) the compiled, unmodified reference (oracle/_ref/libaniso_ref.so, built by oracle/Makefile
 *       from /root/reference) whenever that library is present, and
 *   (2) golden vectors generated from that library and committed under tests/golden/
 *       (generator: tests/golden/make_golden.py), including the known-answer anchors of SURVEY 8(c).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.  The
 * product (aniso_b200/) never does.
 *
 * Conventions restated from the reference (SURVEY "quirks" Q1-Q16):
 *   cell c = i*sz + j with i <-> x, j <-> y;  node index = c*deg^2 + a*deg + b (a <-> x, b <-> y)
 *   Gauss-Legendre tables are in the reference's TABLE order (not sorted)            [Q9]
 *   sigma interpolant is evaluated at GLOBAL coordinates                              [Q1]
 *   legendreNorms are those of the refined rule                                        [Q3]
 *   tree root box = bounding box of the nodes                                          [Q4]
 *   X list handled as M2L, W list handled as P2P                                       [Q6]
 *   kernel arguments are (source a, target b)                                          [Q8]
 *   imag_0(a,a) = +sigma_t(a)                                                          [Q2]
 * ============================================================================================ */
#define _GNU_SOURCE
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_EPS 1e-12 /* bbfmm/utils.h:46 */

/* ------------------------------------------------------------------------------------------ */
/* Gauss-Legendre rules in the reference's table order (Quadrature.cpp:5418-22191).             */
/* The reference tabulates 58-digit decimals; here the nodes are recomputed by Newton iteration  */
/* in long double, polished in binary128 and rounded to double, then permuted into the table's order: centre-outwards,  */
/* (-x,+x) pairs by increasing |x| (odd degrees start with 0), except degrees 6, 7 and 9 whose   */
/* table rows are irregular.                                                                     */
/* ------------------------------------------------------------------------------------------ */
typedef __float128 qfloat; /* compiler soft-float: +,-,*,/ only, no libquadmath needed */
static void gl_ascending(int n, qfloat *x, qfloat *w) {
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 0; i < n; ++i) {
        long double zl = -cosl(pi * (i + 0.75L) / (n + 0.5L));
        for (int it = 0; it < 100; ++it) { /* Newton in long double */
            long double p1 = 1, p2 = 0;
            for (int j = 1; j <= n; ++j) {
                long double p3 = p2;
                p2 = p1;
                p1 = ((2 * j - 1) * zl * p2 - (j - 1) * p3) / j;
            }
            long double dz = p1 / (n * (zl * p1 - p2) / (zl * zl - 1));
            zl -= dz;
            if (fabsl(dz) < 1e-19L) break;
        }
        qfloat z = zl, pp = 0;
        for (int it = 0; it < 3; ++it) { /* polish in binary128 so that rounding to double is correct */
            qfloat p1 = 1, p2 = 0;
            for (int j = 1; j <= n; ++j) {
                qfloat p3 = p2;
                p2 = p1;
                p1 = ((2 * j - 1) * z * p2 - (j - 1) * p3) / j;
            }
            pp = n * (z * p1 - p2) / (z * z - 1);
            if (it < 2) z -= p1 / pp;
        }
        x[i] = z;
        w[i] = 2 / ((1 - z * z) * pp * pp);
    }
    if (n % 2 == 1) x[n / 2] = 0;
}

static void gl_rank(int n, int *rank) { /* rank[slot] = index in ascending order */
    static const int r6[6] = {4, 1, 2, 3, 0, 5};
    static const int r7[7] = {3, 4, 2, 1, 5, 0, 6};
    static const int r9[9] = {4, 1, 7, 0, 8, 3, 5, 2, 6};
    if (n == 6) { memcpy(rank, r6, sizeof r6); return; }
    if (n == 7) { memcpy(rank, r7, sizeof r7); return; }
    if (n == 9) { memcpy(rank, r9, sizeof r9); return; }
    int c = 0;
    if (n % 2 == 1) {
        int mid = n / 2;
        rank[c++] = mid;
        for (int j = 1; j <= mid; ++j) { rank[c++] = mid - j; rank[c++] = mid + j; }
    } else {
        int lo = n / 2 - 1, hi = n / 2;
        for (int j = 0; j < n / 2; ++j) { rank[c++] = lo - j; rank[c++] = hi + j; }
    }
}

/* get_legendre_data, Quadrature.cpp:5418 */
void oracle_gl_table(int n, double *x, double *w) {
    qfloat *xa = malloc(sizeof(qfloat) * n), *wa = malloc(sizeof(qfloat) * n);
    int *rank = malloc(sizeof(int) * n);
    gl_ascending(n, xa, wa);
    gl_rank(n, rank);
    for (int s = 0; s < n; ++s) {
        x[s] = (double)xa[rank[s]];
        w[s] = (double)wa[rank[s]];
    }
    free(xa); free(wa); free(rank);
}

/* std::tr1::legendre as used through bbfmm/utils.h:81 (libstdc++ tr1/legendre_function.tcc:82-114) */
static double legendre_tr1(int l, double x) {
    if (x == 1.0) return 1.0;
    if (x == -1.0) return (l % 2 == 1) ? -1.0 : 1.0;
    double pm2 = 1.0;
    if (l == 0) return pm2;
    double pm1 = x;
    if (l == 1) return pm1;
    double p = 0;
    for (int ll = 2; ll <= l; ++ll) {
        p = 2.0 * x * pm1 - pm2 - (x * pm1 - pm2) / (double)ll;
        pm2 = pm1;
        pm1 = p;
    }
    return p;
}

/* ------------------------------------------------------------------------------------------ */
typedef struct {
    double cx, cy, rx, ry;
    int parent, child[4], level, index, leaf, empty, n; /* n = nSource = nTarget */
    int *pts;                                            /* point indices, reference order */
    int nU, nV, nW, nX, *U, *V, *W, *X;
    double *mcharge, *mpot; /* nodeCharge, nodePotential (rank) */
} onode;

typedef struct {
    int nblk;
    double **blk; /* blocks in traversal order: per node V, X, U, W */
} ocache;

typedef struct {
    int sz, deg, Karg, kernelSize, sr, np, maxLevelArg, N, ncell, c2, rq, nsing;
    double g, dx;
    double *nx, *ny, *w, *glx, *glw, *sgx, *sgw, *sqrtw, *interp, *lnorm, *rqx, *rqy, *rqw, *nearMap;
    double *sig_s, *sig_t, *cs, *ct, *singX, *singY, *singW;
    /* tree */
    onode *nodes;
    int nnodes, capnodes, treeMaxLevel, rank;
    double *cheb, *tnode, *R[4];
    ocache *cache_real, *cache_imag; /* [kernelSize] */
    int use_cache;
    long evals;
} oracle_t;

/* Geometry::makeLegendreMatrix, Geometry.cpp:129-154. K is rows x npts, row-major here. */
static void make_legendre_matrix(oracle_t *o, double *K, int deg, int npts, const double *x, const double *y,
                                 const double *w) {
    int row = 0;
    for (int n = 0; n < deg; ++n)
        for (int k = 0; k < deg; ++k) {
            for (int I = 0; I < npts; ++I) K[row * npts + I] = legendre_tr1(n, x[I]) * legendre_tr1(k, y[I]) * sqrt(w[I]);
            ++row;
        }
    for (row = 0; row < deg * deg; ++row) {
        double nrm = 0;
        for (int c = 0; c < npts; ++c) nrm += K[row * npts + c] * K[row * npts + c];
        nrm = sqrt(nrm);
        o->lnorm[row] = nrm; /* overwritten by the second call: Q3 */
        for (int c = 0; c < npts; ++c) K[row * npts + c] /= nrm;
    }
}

static void duffy(oracle_t *o, const double p[6], double *X, double *Y, double *W); /* fwd */

/* Geometry ctor (Geometry.cpp:10-114) + KernelFactory ctor (KernelFactory.cpp:7-54) + Aniso ctor (Aniso.cpp:7-13) */
oracle_t *oracle_create(int sz, int deg, int K, double g, int sr, int np, int maxLevel) {
    oracle_t *o = calloc(1, sizeof *o);
    o->sz = sz; o->deg = deg; o->Karg = K; o->kernelSize = 2 * K - 1; o->sr = sr; o->np = np;
    o->maxLevelArg = maxLevel; o->g = g;
    o->dx = 1.0 / sz;
    o->ncell = sz * sz;
    o->c2 = deg * deg;
    o->N = o->ncell * o->c2;
    o->rq = 16 * o->c2; /* DEFAULT_REFINE_LEVEL 2, Geometry.h:27 */
    o->nsing = 8 * sr * sr;
    int c2 = o->c2, rq = o->rq;
    o->glx = malloc(sizeof(double) * deg); o->glw = malloc(sizeof(double) * deg);
    oracle_gl_table(deg, o->glx, o->glw);
    o->sgx = malloc(sizeof(double) * sr); o->sgw = malloc(sizeof(double) * sr);
    oracle_gl_table(sr, o->sgx, o->sgw);
    for (int k = 0; k < sr; ++k) { /* affine(), Quadrature.cpp:22194-22199 */
        o->sgw[k] /= 2.0;
        o->sgx[k] += 1.0;
        o->sgx[k] /= 2.0;
    }
    double *qx = malloc(sizeof(double) * c2), *qy = malloc(sizeof(double) * c2), *qw = malloc(sizeof(double) * c2);
    o->sqrtw = malloc(sizeof(double) * c2);
    for (int r = 0; r < deg; ++r)
        for (int c = 0; c < deg; ++c) {
            qx[r * deg + c] = o->glx[r];
            qy[r * deg + c] = o->glx[c];
            qw[r * deg + c] = o->glw[r] * o->glw[c];
            o->sqrtw[r * deg + c] = sqrt(qw[r * deg + c]);
        }
    o->nx = malloc(sizeof(double) * o->N); o->ny = malloc(sizeof(double) * o->N); o->w = malloc(sizeof(double) * o->N);
    double dx = o->dx;
    for (int i = 0; i < sz; ++i)
        for (int j = 0; j < sz; ++j)
            for (int k = 0; k < c2; ++k) {
                int id = (i * sz + j) * c2 + k;
                o->nx[id] = (0.5 + i) * dx + 0.5 * (qx[k]) * dx;
                o->ny[id] = (0.5 + j) * dx + 0.5 * qy[k] * dx;
                o->w[id] = qw[k] * 0.25 * (dx * dx);
            }
    o->lnorm = malloc(sizeof(double) * c2);
    o->interp = malloc(sizeof(double) * c2 * c2);
    make_legendre_matrix(o, o->interp, deg, c2, qx, qy, qw);
    /* two levels of 4-way refinement, push order of Geometry.cpp:79-107 */
    int cur = c2;
    double *rx = malloc(sizeof(double) * rq), *ry = malloc(sizeof(double) * rq), *rw = malloc(sizeof(double) * rq);
    double *tx = malloc(sizeof(double) * rq), *ty = malloc(sizeof(double) * rq), *tw = malloc(sizeof(double) * rq);
    memcpy(rx, qx, sizeof(double) * c2); memcpy(ry, qy, sizeof(double) * c2); memcpy(rw, qw, sizeof(double) * c2);
    for (int level = 0; level < 2; ++level) {
        int m = 0;
        for (int id = 0; id < cur; ++id) {
            static const int sx[4] = {+1, +1, -1, -1}, sy[4] = {+1, -1, +1, -1};
            for (int q = 0; q < 4; ++q) {
                tx[m] = (rx[id] + sx[q]) / 2.0;
                ty[m] = (ry[id] + sy[q]) / 2.0;
                tw[m] = rw[id] / 4.0;
                ++m;
            }
        }
        cur = m;
        memcpy(rx, tx, sizeof(double) * cur); memcpy(ry, ty, sizeof(double) * cur); memcpy(rw, tw, sizeof(double) * cur);
    }
    o->rqx = rx; o->rqy = ry; o->rqw = rw;
    free(tx); free(ty); free(tw);
    double *refinements = malloc(sizeof(double) * c2 * rq);
    make_legendre_matrix(o, refinements, deg, rq, rx, ry, rw);
    /* nearMapping = refinements^T * interpolate (Geometry.cpp:112-113), stored [q'][q] */
    o->nearMap = malloc(sizeof(double) * rq * c2);
    for (int a = 0; a < rq; ++a)
        for (int b = 0; b < c2; ++b) {
            double s = 0;
            for (int r = 0; r < c2; ++r) s += refinements[r * rq + a] * o->interp[r * c2 + b];
            o->nearMap[a * c2 + b] = s;
        }
    free(refinements); free(qx); free(qy); free(qw);
    o->sig_s = calloc(o->N, sizeof(double)); o->sig_t = calloc(o->N, sizeof(double));
    o->cs = calloc(o->N, sizeof(double)); o->ct = calloc(o->N, sizeof(double));
    o->singX = malloc(sizeof(double) * c2 * o->nsing);
    o->singY = malloc(sizeof(double) * c2 * o->nsing);
    o->singW = malloc(sizeof(double) * c2 * o->nsing);
    /* singPrecompute, KernelFactory.cpp:929-986 (8 triangles around the in-cell target) */
    for (int a = 0; a < deg; ++a)
        for (int b = 0; b < deg; ++b) {
            int t = a * deg + b;
            double x = o->glx[a], y = o->glx[b];
            double tri[8][6] = {{x, y, 1., y, 1., 1.},   {x, y, 1., 1, x, 1.},    {x, y, x, 1., -1., 1.},
                                {x, y, -1., 1., -1., y}, {x, y, -1., y, -1., -1.}, {x, y, -1., -1, x, -1.},
                                {x, y, x, -1., 1., -1.}, {x, y, 1., -1, 1., y}};
            for (int k = 0; k < 8; ++k)
                duffy(o, tri[k], o->singX + t * o->nsing + k * sr * sr, o->singY + t * o->nsing + k * sr * sr,
                      o->singW + t * o->nsing + k * sr * sr);
        }
    o->cache_real = calloc(o->kernelSize, sizeof(ocache));
    o->cache_imag = calloc(o->kernelSize, sizeof(ocache));
    return o;
}

/* duffy_transform, KernelFactory.cpp:863-927 */
static void duffy(oracle_t *o, const double p[6], double *X, double *Y, double *W) {
    double a = p[0], b = p[1];
    double a11 = p[2] - p[0], a12 = p[4] - p[2], a21 = p[3] - p[1], a22 = p[5] - p[3];
    double detA = a11 * a22 - a12 * a21;
    int n = o->sr, id = 0;
    for (int ri = 0; ri < n; ++ri)
        for (int ci = 0; ci < n; ++ci) {
            double u = o->sgx[ri], v = o->sgx[ci], w = o->sgw[ri] * o->sgw[ci];
            double z1x = u, z1y = u * v, z1w = w * u;
            X[id] = a11 * z1x + a12 * z1y + a;
            Y[id] = a21 * z1x + a22 * z1y + b;
            W[id] = detA * z1w;
            ++id;
        }
}

int oracle_num_nodes(oracle_t *o) { return o->N; }
int oracle_kernel_size(oracle_t *o) { return o->kernelSize; }
void oracle_get_nodes(oracle_t *o, double *x, double *y, double *w) {
    for (int i = 0; i < o->N; ++i) {
        if (x) x[i] = o->nx[i];
        if (y) y[i] = o->ny[i];
        if (w) w[i] = o->w[i];
    }
}

/* KernelFactory::interpolation(h, coeff), KernelFactory.cpp:988-1005 (and 212-227 for sigma) */
static void interpolate_field(oracle_t *o, const double *h, double *coeff) {
    int c2 = o->c2;
    for (int c = 0; c < o->ncell; ++c)
        for (int r = 0; r < c2; ++r) {
            double s = 0;
            for (int j = 0; j < c2; ++j) s += o->interp[r * c2 + j] * (o->sqrtw[j] * h[c * c2 + j]);
            coeff[c * c2 + r] = s;
        }
}

static void tree_build(oracle_t *o);
static void cache_free(oracle_t *o);

/* 'setCoeff', AnisoWrapper.cpp:46-69 */
void oracle_set_coeff(oracle_t *o, const double *sigma_s, const double *sigma_t) {
    memcpy(o->sig_s, sigma_s, sizeof(double) * o->N);
    memcpy(o->sig_t, sigma_t, sizeof(double) * o->N);
    interpolate_field(o, o->sig_t, o->ct);
    interpolate_field(o, o->sig_s, o->cs);
    cache_free(o);
    if (!o->nodes) tree_build(o);
}

static int cell_of(oracle_t *o, double v) { return (int)floor(v * o->sz); } /* getRow/getCol, KernelFactory.cpp:392-401 */

/* sigma_t interpolant at GLOBAL coordinates with the coefficients of cell (col,row): the inner
 * loops of integral_helper / evaluate, KernelFactory.cpp:180-187, 198-205 */
static double sigma_poly(oracle_t *o, int col, int row, double x, double y) {
    int deg = o->deg;
    const double *c = o->ct + (size_t)(col * o->sz + row) * o->c2;
    double s = 0;
    for (int n = 0; n < deg; ++n)
        for (int k = 0; k < deg; ++k)
            s += (legendre_tr1(n, x) * legendre_tr1(k, y) / o->lnorm[n * deg + k]) * c[n * deg + k];
    return s;
}

/* integral_helper, KernelFactory.cpp:174-190 */
static double seg_integral(oracle_t *o, double x0, double y0, double x1, double y1) {
    int col = cell_of(o, (x0 + x1) / 2), row = cell_of(o, (y0 + y1) / 2);
    double ret = 0;
    for (int i = 0; i < o->deg; ++i) {
        double x = (x0 + x1) / 2 + (x0 - x1) / 2 * o->glx[i];
        double y = (y0 + y1) / 2 + (y0 - y1) / 2 * o->glx[i];
        ret += sigma_poly(o, col, row, x, y) * o->glw[i];
    }
    return ret * sqrt((x0 - x1) * (x0 - x1) + (y0 - y1) * (y0 - y1)) / 2.0;
}

/* lineIntegral, KernelFactory.cpp:67-166: bisect until the end cells coincide or touch, then split
 * at the cell edges. */
double oracle_line_integral(oracle_t *o, double x0, double y0, double x1, double y1) {
    int col0 = cell_of(o, x0), col1 = cell_of(o, x1), row0 = cell_of(o, y0), row1 = cell_of(o, y1);
    int dc = col0 - col1, dr = row0 - row1;
    double side = 1.0 / o->dx;
    if (dc == 0 && dr == 0) return seg_integral(o, x0, y0, x1, y1);
    if (dc == 0 && (dr == 1 || dr == -1)) {
        double ybar = (double)(dr == 1 ? row0 : row1) / side;
        double xbar = ((y1 - ybar) * x0 + (ybar - y0) * x1) / (y1 - y0);
        return seg_integral(o, x0, y0, xbar, ybar) + seg_integral(o, xbar, ybar, x1, y1);
    }
    if (dr == 0 && (dc == 1 || dc == -1)) {
        double xbar = (double)(dc == 1 ? col0 : col1) / side;
        double ybar = ((x1 - xbar) * y0 + (xbar - x0) * y1) / (x1 - x0);
        return seg_integral(o, x0, y0, xbar, ybar) + seg_integral(o, xbar, ybar, x1, y1);
    }
    if ((dc == 1 || dc == -1) && (dr == 1 || dr == -1)) {
        double xbar = (double)(dc == 1 ? col0 : col1) / side;
        double ybar2 = (double)(dr == 1 ? row0 : row1) / side;
        double ybar = ((x1 - xbar) * y0 + (xbar - x0) * y1) / (x1 - x0);
        double xbar2 = ((y1 - ybar2) * x0 + (ybar2 - y0) * x1) / (y1 - y0);
        int vertical_first = (dc == 1) ? (xbar < xbar2) : (xbar > xbar2);
        if (vertical_first)
            return seg_integral(o, x1, y1, xbar, ybar) + seg_integral(o, xbar, ybar, xbar2, ybar2) +
                   seg_integral(o, xbar2, ybar2, x0, y0);
        return seg_integral(o, x1, y1, xbar2, ybar2) + seg_integral(o, xbar, ybar, xbar2, ybar2) +
               seg_integral(o, xbar, ybar, x0, y0);
    }
    double xm = (x0 + x1) / 2, ym = (y0 + y1) / 2;
    return oracle_line_integral(o, x0, y0, xm, ym) + oracle_line_integral(o, xm, ym, x1, y1);
}

/* evaluate, KernelFactory.cpp:193-207 */
double oracle_evaluate(oracle_t *o, double x, double y) { return sigma_poly(o, cell_of(o, x), cell_of(o, y), x, y); }

/* makeKernels, KernelFactory.cpp:240-267. part 0 = realParts, 1 = imagParts; a = source, b = target */
double oracle_kernel_eval(oracle_t *o, int id, int part, double ax, double ay, double bx, double by) {
    double dist = sqrt((ax - bx) * (ax - bx) + (ay - by) * (ay - by));
    double ang = atan2(ay - by, ax - bx);
    if (!part) {
        if (id == 0) return dist == 0. ? 0. : 1 / dist;
        return dist == 0. ? 0. : cos(id * ang) / dist;
    }
    if (id == 0) return dist == 0. ? oracle_evaluate(o, ax, ay) : (exp(-oracle_line_integral(o, ax, ay, bx, by)) - 1) / dist;
    return dist == 0. ? 0. : (exp(-oracle_line_integral(o, ax, ay, bx, by)) - 1) * cos(id * ang) / dist;
}

/* ------------------------------------------------------------------------------------------ */
/* quadtree: tree::populate / assignChildren / buildNode / findNode / isAdjacent, bbfmm.h:176-447 */
/* ------------------------------------------------------------------------------------------ */
static int node_new(oracle_t *o, int level, int index) {
    if (o->nnodes == o->capnodes) {
        o->capnodes = o->capnodes ? 2 * o->capnodes : 1024;
        o->nodes = realloc(o->nodes, sizeof(onode) * o->capnodes);
    }
    onode *n = &o->nodes[o->nnodes];
    memset(n, 0, sizeof *n);
    n->parent = -1;
    for (int i = 0; i < 4; ++i) n->child[i] = -1;
    n->level = level;
    n->index = index;
    return o->nnodes++;
}

static void assign_children(oracle_t *o, int id, int maxLevel) { /* bbfmm.h:250-317 */
    if (o->nodes[id].n == 0) {
        o->nodes[id].leaf = 1;
        o->nodes[id].empty = 1;
        return;
    }
    if (o->nodes[id].n <= o->rank || o->nodes[id].level == maxLevel) {
        o->nodes[id].leaf = 1;
        if (o->treeMaxLevel < o->nodes[id].level) o->treeMaxLevel = o->nodes[id].level;
        return;
    }
    int ch[4];
    for (int i = 0; i < 4; ++i) {
        int c = node_new(o, o->nodes[id].level + 1, i); /* may realloc */
        onode *p = &o->nodes[id], *q = &o->nodes[c];
        ch[i] = c;
        p->child[i] = c;
        q->parent = id;
        q->cx = p->cx + ((i & 1) - 0.5) * p->rx;
        q->cy = p->cy + (((i >> 1) & 1) - 0.5) * p->ry;
        q->rx = p->rx * 0.5;
        q->ry = p->ry * 0.5;
        q->pts = malloc(sizeof(int) * (p->n ? p->n : 1));
    }
    onode *p = &o->nodes[id];
    for (int i = 0; i < p->n; ++i) {
        int idx = p->pts[i];
        int ybit = o->ny[idx] < p->cy ? 0 : 1, xbit = o->nx[idx] < p->cx ? 0 : 1;
        onode *q = &o->nodes[ch[2 * ybit + xbit]];
        q->pts[q->n++] = idx;
    }
    for (int i = 0; i < 4; ++i) assign_children(o, ch[i], maxLevel);
}

static int find_node(oracle_t *o, int id, double px, double py) { /* bbfmm.h:416-429 */
    for (;;) {
        onode *n = &o->nodes[id];
        if (fabs(n->cx - px) < ORACLE_EPS && fabs(n->cy - py) < ORACLE_EPS) return id;
        if (n->leaf) return id;
        int xbit = n->cx > px ? 0 : 1, ybit = n->cy > py ? 0 : 1;
        id = n->child[2 * ybit + xbit];
    }
}

static int is_adjacent(oracle_t *o, int a, int b) { /* bbfmm.h:431-447 */
    onode *A = &o->nodes[a], *B = &o->nodes[b];
    double dx = fabs(A->cx - B->cx), dy = fabs(A->cy - B->cy);
    double rx = fabs(A->rx + B->rx), ry = fabs(A->ry + B->ry);
    int rdx = rx >= dx - ORACLE_EPS, rdy = ry >= dy - ORACLE_EPS;
    int xadj = (fabs(dx - rx) < ORACLE_EPS) && rdy, yadj = (fabs(dy - ry) < ORACLE_EPS) && rdx;
    return xadj || yadj;
}

static void set_insert(int **list, int *n, int v) { /* unordered_set semantics: no duplicates */
    for (int i = 0; i < *n; ++i)
        if ((*list)[i] == v) return;
    *list = realloc(*list, sizeof(int) * (*n + 1));
    (*list)[(*n)++] = v;
}

static void build_node(oracle_t *o, int id, double minx, double miny, double maxx, double maxy) { /* bbfmm.h:334-413 */
    onode *n = &o->nodes[id];
    if (n->parent != -1) {
        onode *pn = &o->nodes[n->parent];
        double dx = n->rx, dy = n->ry, xs = pn->cx - dx, ys = pn->cy - dy;
        for (int xi = -2; xi < 4; ++xi)
            for (int yi = -2; yi < 4; ++yi) {
                double cx = xs + 2 * xi * dx, cy = ys + 2 * yi * dy;
                int inside = (cx <= maxx + ORACLE_EPS) && (cy <= maxy + ORACLE_EPS) && (cx >= minx - ORACLE_EPS) &&
                             (cy >= miny - ORACLE_EPS);
                int self = fabs(cx - n->cx) < ORACLE_EPS && fabs(cy - n->cy) < ORACLE_EPS;
                if (!inside || self) continue;
                int cur = find_node(o, 0, cx, cy);
                int adj = is_adjacent(o, id, cur);
                onode *cn = &o->nodes[cur];
                if (cn->level < n->level) {
                    if (adj) {
                        if (cn->leaf) set_insert(&n->U, &n->nU, cur);
                    } else
                        set_insert(&n->X, &n->nX, cur);
                }
                if (cn->level == n->level) {
                    if (!adj)
                        set_insert(&n->V, &n->nV, cur);
                    else if (n->leaf) {
                        int *queue = malloc(sizeof(int) * (o->nnodes + 4)), qh = 0, qt = 0;
                        queue[qt++] = cur;
                        while (qh < qt) {
                            int f = queue[qh++];
                            if (!is_adjacent(o, f, id))
                                set_insert(&n->W, &n->nW, f);
                            else if (o->nodes[f].leaf)
                                set_insert(&n->U, &n->nU, f);
                            else
                                for (int i = 0; i < 4; ++i) queue[qt++] = o->nodes[f].child[i];
                        }
                        free(queue);
                    }
                }
            }
    }
    if (n->leaf) set_insert(&n->U, &n->nU, id);
}

/* Chebyshev machinery of kernel::initialize, bbfmm.h:476-505, 597-693 */
static double cheb_interp(oracle_t *o, double x, int m) { /* S(x,m) = (2 sum_k T_k(x) T_k(t_m) - 1)/np, bbfmm.h:635-656,742-746 */
    int np = o->np;
    double s = -1.0, t0 = 1.0, t1 = x;
    for (int k = 0; k < np; ++k) {
        double tk = (k == 0) ? 1.0 : (k == 1 ? x : 0.0);
        if (k >= 2) {
            tk = 2.0 * x * t1 - t0;
            t0 = t1;
            t1 = tk;
        }
        s += 2.0 * tk * o->tnode[m * np + k];
    }
    return s / np;
}

static void tree_build(oracle_t *o) {
    int np = o->np;
    o->rank = np * np; /* KernelFactory.cpp:284-286 passes np*np as _rank */
    double xmax = o->nx[0], xmin = o->nx[0], ymax = o->ny[0], ymin = o->ny[0];
    for (int i = 0; i < o->N; ++i) { /* getCenterRadius, bbfmm.h:231-248 */
        if (o->nx[i] > xmax) xmax = o->nx[i];
        if (o->nx[i] < xmin) xmin = o->nx[i];
        if (o->ny[i] > ymax) ymax = o->ny[i];
        if (o->ny[i] < ymin) ymin = o->ny[i];
    }
    int root = node_new(o, 0, 0);
    onode *r = &o->nodes[root];
    r->cx = (xmax + xmin) / 2.0; r->cy = (ymax + ymin) / 2.0;
    r->rx = (xmax - xmin) / 2.0; r->ry = (ymax - ymin) / 2.0;
    r->n = o->N;
    r->pts = malloc(sizeof(int) * o->N);
    for (int i = 0; i < o->N; ++i) r->pts[i] = i;
    o->treeMaxLevel = 0;
    assign_children(o, root, o->maxLevelArg);
    r = &o->nodes[root];
    double minx = r->cx - r->rx, miny = r->cy - r->ry, maxx = r->cx + r->rx, maxy = r->cy + r->ry;
    for (int i = 0; i < o->nnodes; ++i) build_node(o, i, minx, miny, maxx, maxy);
    /* Chebyshev nodes and T_k(t_m) (bbfmm.h:597-630) */
    o->cheb = malloc(sizeof(double) * np);
    o->tnode = malloc(sizeof(double) * np * np);
    for (int i = 0; i < np; ++i) o->cheb[i] = -cos((i + 0.5) * M_PI / np);
    for (int m = 0; m < np; ++m) {
        double x = o->cheb[m];
        for (int k = 0; k < np; ++k) o->tnode[m * np + k] = (k == 0) ? 1.0 : (k == 1 ? x : 2.0 * x * o->tnode[m * np + k - 1] - o->tnode[m * np + k - 2]);
    }
    /* M2M/L2L operators, bbfmm.h:661-693: R[c][(i,j),(k,l)] = S_{ybit}(i,k) * S_{xbit}(j,l) */
    int rk = o->rank;
    for (int c = 0; c < 4; ++c) {
        o->R[c] = malloc(sizeof(double) * rk * rk);
        int bx = c & 1, by = (c >> 1) & 1;
        for (int i = 0; i < np; ++i)
            for (int j = 0; j < np; ++j)
                for (int k = 0; k < np; ++k)
                    for (int l = 0; l < np; ++l) {
                        double ci = (by ? 0.5 : -0.5) + 0.5 * o->cheb[i]; /* child node i in parent coordinates */
                        double cj = (bx ? 0.5 : -0.5) + 0.5 * o->cheb[j];
                        o->R[c][(i * np + j) * rk + (k * np + l)] = cheb_interp(o, ci, k) * cheb_interp(o, cj, l);
                    }
    }
    for (int i = 0; i < o->nnodes; ++i) {
        o->nodes[i].mcharge = calloc(rk, sizeof(double));
        o->nodes[i].mpot = calloc(rk, sizeof(double));
    }
}

/* ------------------------------------------------------------------------------------------ */
/* FMM apply: kernel::run / runCache / runFast, bbfmm.h:507-592, 825-1129                        */
/* ------------------------------------------------------------------------------------------ */
static void cheb_point(oracle_t *o, onode *n, int flat, double *x, double *y) { /* flat = j*np + i, y slow: bbfmm.h:787-800 */
    int i = flat % o->np, j = flat / o->np;
    *x = n->cx + n->rx * o->cheb[i];
    *y = n->cy + n->ry * o->cheb[j];
}

static void up_pass(oracle_t *o, int id, const double *charge) { /* bbfmm.h:825-861 */
    onode *n = &o->nodes[id];
    int rk = o->rank, np = o->np;
    memset(n->mcharge, 0, sizeof(double) * rk);
    memset(n->mpot, 0, sizeof(double) * rk);
    if (n->leaf) {
        for (int k = 0; k < n->n; ++k) { /* R(k, j*np+i) = Sx(k,i) Sy(k,j), bbfmm.h:723-759 */
            double sx = (o->nx[n->pts[k]] - n->cx) / n->rx, sy = (o->ny[n->pts[k]] - n->cy) / n->ry;
            for (int j = 0; j < np; ++j)
                for (int i = 0; i < np; ++i) n->mcharge[j * np + i] += cheb_interp(o, sx, i) * cheb_interp(o, sy, j) * charge[n->pts[k]];
        }
    } else {
        for (int c = 0; c < 4; ++c) up_pass(o, n->child[c], charge);
        for (int c = 0; c < 4; ++c) {
            onode *ch = &o->nodes[n->child[c]];
            if (ch->empty) continue;
            for (int a = 0; a < rk; ++a) /* parent += R[c]^T child */
                for (int b = 0; b < rk; ++b) n->mcharge[b] += o->R[c][a * rk + b] * ch->mcharge[a];
        }
    }
}

static double *eval_block(oracle_t *o, int id, int part, onode *tgt, onode *src, int cheb) {
    /* K[t][s] = eval(source_s, target_t): kernelEvalChebyshev (bbfmm.h:782-804) / kernelEvalIndex (770-780) */
    int nt = cheb ? o->rank : tgt->n, ns = cheb ? o->rank : src->n;
    double *K = malloc(sizeof(double) * (nt * ns ? nt * ns : 1));
    for (int s = 0; s < ns; ++s)
        for (int t = 0; t < nt; ++t) {
            double ax, ay, bx, by;
            if (cheb) {
                cheb_point(o, src, s, &ax, &ay);
                cheb_point(o, tgt, t, &bx, &by);
            } else {
                ax = o->nx[src->pts[s]]; ay = o->ny[src->pts[s]];
                bx = o->nx[tgt->pts[t]]; by = o->ny[tgt->pts[t]];
            }
            K[t * ns + s] = oracle_kernel_eval(o, id, part, ax, ay, bx, by);
        }
    o->evals += (long)nt * ns;
    return K;
}

static void down_pass(oracle_t *o, int nid, int id, int part, const double *charge, double *pot, ocache *cache,
                      int *slot, int filling) { /* bbfmm.h:863-947 (run), 949-1039 (cache), 1041-1129 (fast) */
    onode *n = &o->nodes[nid];
    int rk = o->rank, np = o->np;
    if (n->parent != -1) {
        for (int pass = 0; pass < 2; ++pass) { /* V then X, both as M2L: Q6 */
            int cnt = pass ? n->nX : n->nV, *lst = pass ? n->X : n->V;
            for (int e = 0; e < cnt; ++e) {
                onode *s = &o->nodes[lst[e]];
                if (s->empty) continue;
                double *K;
                if (cache && !filling) K = cache->blk[(*slot)++];
                else {
                    K = eval_block(o, id, part, n, s, 1);
                    if (cache) cache->blk[(*slot)++] = K;
                }
                for (int t = 0; t < rk; ++t) {
                    double acc = 0;
                    for (int q = 0; q < rk; ++q) acc += K[t * rk + q] * s->mcharge[q];
                    n->mpot[t] += acc;
                }
                if (!cache) free(K);
            }
        }
        onode *p = &o->nodes[n->parent]; /* L2L: child += R[index] * parent */
        for (int a = 0; a < rk; ++a) {
            double acc = 0;
            for (int b = 0; b < rk; ++b) acc += o->R[n->index][a * rk + b] * p->mpot[b];
            n->mpot[a] += acc;
        }
    }
    if (n->leaf && n->n != 0) {
        double *lp = calloc(n->n, sizeof(double));
        for (int pass = 0; pass < 2; ++pass) { /* U then W, both as P2P: Q6 */
            int cnt = pass ? n->nW : n->nU, *lst = pass ? n->W : n->U;
            for (int e = 0; e < cnt; ++e) {
                onode *s = &o->nodes[lst[e]];
                if (s->empty) continue;
                double *K;
                if (cache && !filling) K = cache->blk[(*slot)++];
                else {
                    K = eval_block(o, id, part, n, s, 0);
                    if (cache) cache->blk[(*slot)++] = K;
                }
                for (int t = 0; t < n->n; ++t) {
                    double acc = 0;
                    for (int q = 0; q < s->n; ++q) acc += K[t * s->n + q] * charge[s->pts[q]];
                    lp[t] += acc;
                }
                if (!cache) free(K);
            }
        }
        for (int k = 0; k < n->n; ++k) { /* L2T: pot += L * nodePotential */
            double sx = (o->nx[n->pts[k]] - n->cx) / n->rx, sy = (o->ny[n->pts[k]] - n->cy) / n->ry;
            double acc = 0;
            for (int j = 0; j < np; ++j)
                for (int i = 0; i < np; ++i) acc += cheb_interp(o, sx, i) * cheb_interp(o, sy, j) * n->mpot[j * np + i];
            lp[k] += acc;
        }
        for (int k = 0; k < n->n; ++k) pot[n->pts[k]] += lp[k];
        free(lp);
    }
    if (!n->leaf)
        for (int c = 0; c < 4; ++c) down_pass(o, n->child[c], id, part, charge, pot, cache, slot, filling);
}

static long count_blocks(oracle_t *o) {
    long c = 0;
    for (int i = 0; i < o->nnodes; ++i) c += o->nodes[i].nU + o->nodes[i].nV + o->nodes[i].nW + o->nodes[i].nX;
    return c;
}

static void cache_free(oracle_t *o) {
    for (int part = 0; part < 2; ++part) {
        ocache *arr = part ? o->cache_imag : o->cache_real;
        for (int id = 0; id < o->kernelSize; ++id) {
            for (int b = 0; b < arr[id].nblk; ++b) free(arr[id].blk[b]);
            free(arr[id].blk);
            arr[id].blk = NULL;
            arr[id].nblk = 0;
        }
    }
}

/* runKernelsFast / runKernelsFastSing (KernelFactory.cpp:354-387) = initialize + runFast; without a
 * prior oracle_cache(id) this is kernel::run (evaluate on the fly) */
void oracle_fmm(oracle_t *o, int id, int part, const double *charge, double *pot) {
    ocache *c = part ? &o->cache_imag[id] : &o->cache_real[id];
    memset(pot, 0, sizeof(double) * o->N);
    up_pass(o, 0, charge);
    int slot = 0;
    down_pass(o, 0, id, part, charge, pot, c->blk ? c : NULL, &slot, 0);
}

/* 'cache', AnisoWrapper.cpp:72-90: runKernelsCache + runKernelsCacheSing (the two correction caches are
 * pure memoisation and are recomputed on the fly below) */
void oracle_cache(oracle_t *o, int id) {
    double *zero = calloc(o->N, sizeof(double)), *pot = calloc(o->N, sizeof(double));
    long nb = count_blocks(o);
    for (int part = 0; part < 2; ++part) {
        ocache *c = part ? &o->cache_imag[id] : &o->cache_real[id];
        if (c->blk) continue;
        c->blk = calloc(nb, sizeof(double *));
        up_pass(o, 0, zero);
        int slot = 0;
        down_pass(o, 0, id, part, zero, pot, c, &slot, 1);
        c->nblk = slot;
    }
    free(zero); free(pot);
}

/* nearRemoval, KernelFactory.cpp:445-478 */
void oracle_near_removal(oracle_t *o, int id, const double *f, double *ret) {
    int sz = o->sz, c2 = o->c2;
    for (int tc = 0; tc < o->ncell; ++tc)
        for (int tq = 0; tq < c2; ++tq) {
            int t = tc * c2 + tq, trow = tc / sz, tcol = tc - sz * trow;
            for (int dr = -1; dr < 2; ++dr)
                for (int dc = -1; dc < 2; ++dc) {
                    if (dr + trow < 0 || dr + trow >= sz || dc + tcol < 0 || dc + tcol >= sz) continue;
                    int sc = tc + dr * sz + dc;
                    for (int q = 0; q < c2; ++q) {
                        int s = sc * c2 + q;
                        ret[t] -= oracle_kernel_eval(o, id, 0, o->nx[s], o->ny[s], o->nx[t], o->ny[t]) * f[s];
                    }
                }
        }
}

/* refineAddOnCache + refineAddOnFast, KernelFactory.cpp:550-609, 662-709 */
void oracle_refine_add(oracle_t *o, int id, const double *f, double *ret) {
    int sz = o->sz, c2 = o->c2, rq = o->rq;
    double dx = o->dx;
    double *oldv = malloc(sizeof(double) * c2), *newv = malloc(sizeof(double) * rq);
    for (int tc = 0; tc < o->ncell; ++tc)
        for (int tq = 0; tq < c2; ++tq) {
            int t = tc * c2 + tq, trow = tc / sz, tcol = tc - sz * trow;
            for (int dr = -1; dr < 2; ++dr)
                for (int dc = -1; dc < 2; ++dc) {
                    int sc = tc + dr * sz + dc;
                    if (sc == tc) continue;
                    if (dr + trow < 0 || dr + trow >= sz || dc + tcol < 0 || dc + tcol >= sz) continue;
                    for (int q = 0; q < c2; ++q) oldv[q] = f[sc * c2 + q] / o->sqrtw[q];
                    for (int a = 0; a < rq; ++a) {
                        double s = 0;
                        for (int q = 0; q < c2; ++q) s += o->nearMap[a * c2 + q] * oldv[q];
                        newv[a] = s;
                    }
                    for (int a = 0; a < rq; ++a) {
                        double px = (0.5 + (trow + dr)) * dx + 0.5 * (o->rqx[a]) * dx;
                        double py = (0.5 + (tcol + dc)) * dx + 0.5 * (o->rqy[a]) * dx;
                        double inter = oracle_kernel_eval(o, id, 0, px, py, o->nx[t], o->ny[t]) * sqrt(o->rqw[a]);
                        ret[t] += inter * newv[a];
                    }
                }
        }
    free(oldv); free(newv);
}

/* singularAddCache + singularAddFast, KernelFactory.cpp:752-788, 828-860 */
void oracle_singular_add(oracle_t *o, int id, const double *fcoeff, double *ret) {
    int sz = o->sz, c2 = o->c2, deg = o->deg;
    double dx = o->dx;
    for (int tc = 0; tc < o->ncell; ++tc)
        for (int tq = 0; tq < c2; ++tq) {
            int col = tc / sz, row = tc - col * sz, t = tc * c2 + tq;
            for (int q = 0; q < o->nsing; ++q) {
                double x = (0.5 + col) * dx + 0.5 * (o->singX[tq * o->nsing + q]) * dx;
                double y = (0.5 + row) * dx + 0.5 * (o->singY[tq * o->nsing + q]) * dx;
                double w = o->singW[tq * o->nsing + q] * (dx * dx) / 4.0;
                double inter = oracle_kernel_eval(o, id, 0, x, y, o->nx[t], o->ny[t]) * w;
                double fv = 0;
                for (int n = 0; n < deg; ++n)
                    for (int k = 0; k < deg; ++k)
                        fv += (legendre_tr1(n, x) * legendre_tr1(k, y) / o->lnorm[n * deg + k]) * fcoeff[(col * sz + row) * c2 + n * deg + k];
                ret[t] += fv * inter;
            }
        }
}

/* 'mapping', AnisoWrapper.cpp:92-136 == apply_mapping, main.cpp:78-119; optional stage outputs */
void oracle_mapping_stages(oracle_t *o, const double *charge, int id, double *out_s, double *real_fmm, double *k7,
                           double *k8, double *k9, double *out) {
    int N = o->N;
    double *scaled = malloc(sizeof(double) * N), *coeff = malloc(sizeof(double) * N);
    double *os = malloc(sizeof(double) * N), *ot = malloc(sizeof(double) * N);
    for (int i = 0; i < N; ++i) scaled[i] = charge[i] * o->w[i];
    interpolate_field(o, charge, coeff);
    oracle_fmm(o, id, 1, scaled, os);
    if (out_s) memcpy(out_s, os, sizeof(double) * N);
    oracle_fmm(o, id, 0, scaled, ot);
    if (real_fmm) memcpy(real_fmm, ot, sizeof(double) * N);
    oracle_near_removal(o, id, scaled, ot);
    if (k7) memcpy(k7, ot, sizeof(double) * N);
    oracle_refine_add(o, id, scaled, ot);
    if (k8) memcpy(k8, ot, sizeof(double) * N);
    oracle_singular_add(o, id, coeff, ot);
    if (k9) memcpy(k9, ot, sizeof(double) * N);
    for (int i = 0; i < N; ++i) ot[i] = (ot[i] + os[i]) * (M_1_PI / 2.0);
    if (out) memcpy(out, ot, sizeof(double) * N);
    free(scaled); free(coeff); free(os); free(ot);
}

void oracle_mapping(oracle_t *o, const double *charge, int id, double *out) {
    oracle_mapping_stages(o, charge, id, NULL, NULL, NULL, NULL, NULL, out);
}

/* ------------------------------------------------------------------------------------------ */
/* GMRES, gmres.cpp:12-169, operator of main.cpp:123-136: A x = x - mapping_0(sigma_s .* x)      */
/* ------------------------------------------------------------------------------------------ */
static double vdot(int n, const double *a, const double *b) {
    double s = 0;
    for (int i = 0; i < n; ++i) s += a[i] * b[i];
    return s;
}
static double vnrm2(int n, const double *a) { return sqrt(vdot(n, a, a)); }

static void apply_A(oracle_t *o, const double *x, double *y, double *tmp) {
    for (int i = 0; i < o->N; ++i) tmp[i] = x[i] * o->sig_s[i];
    oracle_mapping(o, tmp, 0, y);
    for (int i = 0; i < o->N; ++i) y[i] = x[i] - y[i];
}

static void plane_gen(double dx, double dy, double *cs, double *sn) { /* gmres.cpp:26-39 */
    if (dy == 0.0) { *cs = 1.0; *sn = 0.0; }
    else if (fabs(dy) > fabs(dx)) { double t = dx / dy; *sn = 1.0 / sqrt(1.0 + t * t); *cs = t * *sn; }
    else { double t = dy / dx; *cs = 1.0 / sqrt(1.0 + t * t); *sn = t * *cs; }
}
static void plane_apply(double *dx, double *dy, double cs, double sn) { /* gmres.cpp:42-46 */
    double t = cs * *dx + sn * *dy;
    *dy = -sn * *dx + cs * *dy;
    *dx = t;
}
static void gm_update(int n, double *x, int k, const double *H, int ldh, const double *s, double **v) { /* gmres.cpp:12-24 */
    double *y = malloc(sizeof(double) * (k + 2));
    memcpy(y, s, sizeof(double) * (k + 1));
    for (int i = k; i >= 0; i--) {
        y[i] /= H[i + i * ldh];
        for (int j = i - 1; j >= 0; j--) y[j] -= H[j + i * ldh] * y[i];
    }
    for (int j = 0; j <= k; j++)
        for (int i = 0; i < n; ++i) x[i] += y[j] * v[j][i];
    free(y);
}

/* returns 0 converged / 1 not (gmres.cpp:53-169). hist[j-1] = residual printed at iteration j
 * (i.e. the residual BEFORE that iteration); hist[*iters] = final residual. */
int oracle_gmres(oracle_t *o, const double *b, double *x, int m, int max_iter, double tol, int *iters, double *hist,
                 int hist_cap) {
    int n = o->N, j = 1, i, k, ldh = m + 1;
    double *s = calloc(m + 1, sizeof(double)), *cs = calloc(m + 1, sizeof(double)), *sn = calloc(m + 1, sizeof(double));
    double *w = malloc(sizeof(double) * n), *r = malloc(sizeof(double) * n), *tmp = malloc(sizeof(double) * n);
    double *H = calloc((size_t)(m + 1) * m, sizeof(double));
    double **v = calloc(m + 1, sizeof(double *));
    for (i = 0; i <= m; ++i) v[i] = malloc(sizeof(double) * n);
    double normb = vnrm2(n, b), resid, beta;
    int rc = 1, done = 0;
    apply_A(o, x, w, tmp);
    for (i = 0; i < n; ++i) r[i] = b[i] - w[i];
    beta = vnrm2(n, r);
    if (normb == 0.0) normb = 1;
    *iters = 0;
    if ((resid = vnrm2(n, r) / normb) <= tol) { rc = 0; done = 1; }
    while (!done && j <= max_iter) {
        for (i = 0; i < n; ++i) v[0][i] = r[i] * (1.0 / beta);
        memset(s, 0, sizeof(double) * (m + 1));
        s[0] = beta;
        for (i = 0; i < m && j <= max_iter; i++, j++) {
            if (hist && j - 1 < hist_cap) hist[j - 1] = resid;
            *iters = j;
            apply_A(o, v[i], w, tmp);
            for (k = 0; k <= i; k++) {
                H[k + i * ldh] = vdot(n, w, v[k]);
                for (int q = 0; q < n; ++q) w[q] -= H[k + i * ldh] * v[k][q];
            }
            H[(i + 1) + i * ldh] = vnrm2(n, w);
            for (int q = 0; q < n; ++q) v[i + 1][q] = w[q] * (1.0 / H[(i + 1) + i * ldh]);
            for (k = 0; k < i; k++) plane_apply(&H[k + i * ldh], &H[k + 1 + i * ldh], cs[k], sn[k]);
            plane_gen(H[i + i * ldh], H[i + 1 + i * ldh], &cs[i], &sn[i]);
            plane_apply(&H[i + i * ldh], &H[i + 1 + i * ldh], cs[i], sn[i]);
            plane_apply(&s[i], &s[i + 1], cs[i], sn[i]);
            if ((resid = fabs(s[i + 1]) / normb) < tol) {
                gm_update(n, x, i, H, ldh, s, v);
                rc = 0; done = 1;
                break;
            }
        }
        if (done) break;
        gm_update(n, x, i - 1, H, ldh, s, v);
        apply_A(o, x, w, tmp);
        for (int q = 0; q < n; ++q) r[q] = b[q] - w[q];
        beta = vnrm2(n, r);
        if ((resid = beta / normb) < tol) { rc = 0; done = 1; }
    }
    if (hist && *iters < hist_cap) hist[*iters] = resid;
    for (i = 0; i <= m; ++i) free(v[i]);
    free(v); free(H); free(s); free(cs); free(sn); free(w); free(r); free(tmp);
    return rc;
}

/* ------------------------------------------------------------------------------------------ */
/* introspection for host-side parity tests                                                     */
/* ------------------------------------------------------------------------------------------ */
void oracle_tree_stats(oracle_t *o, long *out) {
    long leaves = 0, su = 0, sv = 0, sw = 0, sx = 0, pairs = 0, mx = 0;
    for (int i = 0; i < o->nnodes; ++i) {
        onode *n = &o->nodes[i];
        if (n->leaf) {
            leaves++;
            if (n->n > mx) mx = n->n;
            for (int e = 0; e < n->nU; ++e) pairs += (long)o->nodes[n->U[e]].n * n->n;
            for (int e = 0; e < n->nW; ++e) pairs += (long)o->nodes[n->W[e]].n * n->n;
        }
        su += n->nU; sv += n->nV; sw += n->nW; sx += n->nX;
    }
    out[0] = o->nnodes; out[1] = leaves; out[2] = o->treeMaxLevel; out[3] = su; out[4] = sv; out[5] = sw;
    out[6] = sx; out[7] = pairs; out[8] = mx;
}

int oracle_tree_nodes(oracle_t *o, int cap, double *geom, int *info) {
    for (int i = 0; i < o->nnodes && i < cap; ++i) {
        onode *n = &o->nodes[i];
        geom[4 * i] = n->cx; geom[4 * i + 1] = n->cy; geom[4 * i + 2] = n->rx; geom[4 * i + 3] = n->ry;
        info[4 * i] = n->level; info[4 * i + 1] = n->leaf; info[4 * i + 2] = n->n; info[4 * i + 3] = n->parent;
    }
    return o->nnodes;
}

int oracle_node_lists(oracle_t *o, int node, int cap, int *ids, int *kinds) {
    onode *n = &o->nodes[node];
    int c = 0;
    for (int e = 0; e < n->nV; ++e, ++c) if (c < cap) { ids[c] = n->V[e]; kinds[c] = 0; }
    for (int e = 0; e < n->nX; ++e, ++c) if (c < cap) { ids[c] = n->X[e]; kinds[c] = 1; }
    for (int e = 0; e < n->nU; ++e, ++c) if (c < cap) { ids[c] = n->U[e]; kinds[c] = 2; }
    for (int e = 0; e < n->nW; ++e, ++c) if (c < cap) { ids[c] = n->W[e]; kinds[c] = 3; }
    return c;
}

void oracle_geometry_tables(oracle_t *o, double *interp, double *nearMap, double *lnorm, double *rqx, double *rqy,
                            double *rw) {
    if (interp) memcpy(interp, o->interp, sizeof(double) * o->c2 * o->c2);
    if (nearMap) memcpy(nearMap, o->nearMap, sizeof(double) * o->rq * o->c2);
    if (lnorm) memcpy(lnorm, o->lnorm, sizeof(double) * o->c2);
    if (rqx) memcpy(rqx, o->rqx, sizeof(double) * o->rq);
    if (rqy) memcpy(rqy, o->rqy, sizeof(double) * o->rq);
    if (rw) memcpy(rw, o->rqw, sizeof(double) * o->rq);
}

void oracle_sing_tables(oracle_t *o, double *sx, double *sy, double *sw) {
    memcpy(sx, o->singX, sizeof(double) * o->c2 * o->nsing);
    memcpy(sy, o->singY, sizeof(double) * o->c2 * o->nsing);
    memcpy(sw, o->singW, sizeof(double) * o->c2 * o->nsing);
}

void oracle_sigma_coeff(oracle_t *o, double *ct, double *cs) {
    if (ct) memcpy(ct, o->ct, sizeof(double) * o->N);
    if (cs) memcpy(cs, o->cs, sizeof(double) * o->N);
}

long oracle_evals(oracle_t *o) { return o->evals; }

void oracle_destroy(oracle_t *o) {
    if (!o) return;
    cache_free(o);
    for (int i = 0; i < o->nnodes; ++i) {
        onode *n = &o->nodes[i];
        free(n->pts); free(n->U); free(n->V); free(n->W); free(n->X); free(n->mcharge); free(n->mpot);
    }
    free(o->nodes);
    for (int c = 0; c < 4; ++c) free(o->R[c]);
    free(o->cheb); free(o->tnode); free(o->cache_real); free(o->cache_imag);
    free(o->nx); free(o->ny); free(o->w); free(o->glx); free(o->glw); free(o->sgx); free(o->sgw); free(o->sqrtw);
    free(o->interp); free(o->lnorm); free(o->rqx); free(o->rqy); free(o->rqw); free(o->nearMap);
    free(o->sig_s); free(o->sig_t); free(o->cs); free(o->ct); free(o->singX); free(o->singY); free(o->singW);
    free(o);
}
