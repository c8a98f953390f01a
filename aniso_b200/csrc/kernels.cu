// Hand-written sm_100a kernels of the aniso matvec.  FP64 throughout (parity gate 1e-10).
//   fill_blocks      K1/K2 evaluator: optical-depth ray march + kernel family, writes the cached operators
//   bsr_apply        K3/K4 cached apply: one HBM pass over the operator blocks
//   p2m/m2m/l2l/l2p  K5/K6 Chebyshev passes
//   build_stencils   K7+K8+K9 corrections folded into one 9-cell stencil per target (built once per id)
//   stencil_apply    the per-matvec correction
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "device.cuh"

namespace an {

// ------------------------------------------------------------------------------------------------
// sigma_t interpolant and its line integral
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int cell_index(double v, int sz) {  // getRow/getCol, KernelFactory.cpp:392-401
    int c = (int)floor(v * sz);
    return min(max(c, 0), sz - 1);  // the reference does not clamp: see magic_floor
}

// sum_{n,k} P_n(x) P_k(y) chat[n*deg+k] with the libstdc++ tr1 Legendre recurrence at GLOBAL coordinates
// (KernelFactory.cpp:180-187, 198-205; quirk Q1).  chat already holds coeff / legendreNorms.
__device__ __forceinline__ double sigma_poly(const double* __restrict__ chat, int deg, double x, double y) {
    if (deg == 1) return __ldg(chat);
    double acc = 0, pxm2 = 1.0, pxm1 = x;
    for (int n = 0; n < deg; ++n) {
        double pxn;
        if (n == 0) pxn = 1.0;
        else if (n == 1) pxn = x;
        else {
            pxn = 2.0 * x * pxm1 - pxm2 - (x * pxm1 - pxm2) / (double)n;
            pxm2 = pxm1;
            pxm1 = pxn;
        }
        double pym2 = 1.0, pym1 = y, inner = 0;
        for (int k = 0; k < deg; ++k) {
            double pyk;
            if (k == 0) pyk = 1.0;
            else if (k == 1) pyk = y;
            else {
                pyk = 2.0 * y * pym1 - pym2 - (y * pym1 - pym2) / (double)k;
                pym2 = pym1;
                pym1 = pyk;
            }
            inner += pyk * __ldg(chat + n * deg + k);
        }
        acc += pxn * inner;
    }
    return acc;
}

__device__ __forceinline__ double sigma_eval(const SigmaGrid& S, double x, double y) {  // evaluate, KernelFactory.cpp:193-207
    const int col = cell_index(x, S.sz), row = cell_index(y, S.sz);
    return sigma_poly(S.chat + (size_t)(col * S.sz + row) * S.c2, S.deg, x, y);
}

// Optical depth tau(p0,p1) = exact integral of the piecewise-polynomial interpolant along the segment
// (lineIntegral + integral_helper, KernelFactory.cpp:67-190).  The reference bisects until the end cells touch and
// then splits at the cell edges; every piece is integrated with the deg-point Gauss rule, which is exact for the
// degree-2(deg-1) polynomial along the ray, so the value equals the sum over the cells the ray crosses.  That sum
// is what this cell walk computes: crossing parameters are evaluated directly (not accumulated) per step.
// DEG > 0: the degree is a compile-time constant (2 ... 5 are instantiated).  Both one-dimensional Legendre recurrences are
// run once per point into registers and followed by the double sum, instead of re-running the y recurrence (one FP64
// division per step) for every x order; the operations and their order are those of sigma_poly, so the values are the same.
template <int DEG>
__device__ __forceinline__ double sigma_poly_t(const double* __restrict__ chat, int deg, double x, double y) {
    if (DEG == 0) return sigma_poly(chat, deg, x, y);
    constexpr int D = DEG > 0 ? DEG : 1;
    double px[D], py[D];
    px[0] = 1.0;
    py[0] = 1.0;
    if (D > 1) {
        px[1] = x;
        py[1] = y;
    }
#pragma unroll
    for (int n = 2; n < D; ++n) {
        px[n] = 2.0 * x * px[n - 1] - px[n - 2] - (x * px[n - 1] - px[n - 2]) / (double)n;
        py[n] = 2.0 * y * py[n - 1] - py[n - 2] - (y * py[n - 1] - py[n - 2]) / (double)n;
    }
    double acc = 0;
#pragma unroll
    for (int n = 0; n < D; ++n) {
        double inner = 0;
#pragma unroll
        for (int k = 0; k < D; ++k) inner += py[k] * __ldg(chat + n * D + k);
        acc += px[n] * inner;
    }
    return acc;
}

template <int DEG>
__device__ double line_integral_t(const SigmaGrid& S, double x0, double y0, double x1, double y1) {
    const int sz = S.sz;
    const double ddx = x1 - x0, ddy = y1 - y0;
    const double len = sqrt(ddx * ddx + ddy * ddy);
    int cx = cell_index(x0, sz), cy = cell_index(y0, sz);
    const int ex = cell_index(x1, sz), ey = cell_index(y1, sz);
    const int sx = ex >= cx ? 1 : -1, sy = ey >= cy ? 1 : -1;
    const double ivx = ddx != 0.0 ? 1.0 / ddx : 0.0, ivy = ddy != 0.0 ? 1.0 / ddy : 0.0;
    const double inf = 2.0;
    int bx = cx + (sx > 0), by = cy + (sy > 0);
    double tx = cx != ex ? ((double)bx * S.h - x0) * ivx : inf;
    double ty = cy != ey ? ((double)by * S.h - y0) * ivy : inf;
    const int nsteps = abs(ex - cx) + abs(ey - cy);
    double tcur = 0.0, acc = 0.0;
    const int deg = DEG ? DEG : S.deg;
    constexpr int GD = DEG > 0 ? DEG : 1;
    double gx[GD], gw[GD];  // the Gauss rule in registers when the degree is known
    if (DEG > 0) {
#pragma unroll
        for (int i = 0; i < GD; ++i) {
            gx[i] = __ldg(S.glx + i);
            gw[i] = __ldg(S.glw + i);
        }
    }
    for (int step = 0; step <= nsteps; ++step) {
        const double tn = step == nsteps ? 1.0 : fmin(tx, ty);
        const double* __restrict__ chat = S.chat + (size_t)(cx * sz + cy) * S.c2;
        if (deg == 1) {
            acc += __ldg(chat) * (tn - tcur);
        } else {
            // sub-segment end points and its Gauss points (integral_helper, KernelFactory.cpp:174-190)
            const double xa = x0 + tcur * ddx, ya = y0 + tcur * ddy, xb = x0 + tn * ddx, yb = y0 + tn * ddy;
            const double mx = (xa + xb) / 2, my = (ya + yb) / 2, hx = (xa - xb) / 2, hy = (ya - yb) / 2;
            double s = 0;
            if (DEG > 0) {
#pragma unroll
                for (int i = 0; i < GD; ++i) s += sigma_poly_t<DEG>(chat, deg, mx + hx * gx[i], my + hy * gx[i]) * gw[i];
            } else {
                for (int i = 0; i < deg; ++i) {
                    const double g = __ldg(S.glx + i);
                    s += sigma_poly(chat, deg, mx + hx * g, my + hy * g) * __ldg(S.glw + i);
                }
            }
            acc += s * (tn - tcur) * 0.5;
        }
        if (step < nsteps) {
            if (tx <= ty) {
                cx += sx;
                bx += sx;
                tx = cx != ex ? ((double)bx * S.h - x0) * ivx : inf;
            } else {
                cy += sy;
                by += sy;
                ty = cy != ey ? ((double)by * S.h - y0) * ivy : inf;
            }
        }
        tcur = tn;
    }
    return acc * len;
}

__device__ double line_integral(const SigmaGrid& S, double x0, double y0, double x1, double y1) {
    switch (S.deg) {
        case 2: return line_integral_t<2>(S, x0, y0, x1, y1);
        case 3: return line_integral_t<3>(S, x0, y0, x1, y1);  // data.cfg (BASELINE config 1)
        case 4: return line_integral_t<4>(S, x0, y0, x1, y1);
        case 5: return line_integral_t<5>(S, x0, y0, x1, y1);
        default: return line_integral_t<0>(S, x0, y0, x1, y1);
    }
}

// deg == 1 fast path.  sigma is constant per cell, so along a grid line its integral is piecewise linear:
//   F_c(v) = int_0^v sigma(c, v') dv' = P[c][j] + sigma[c][j] (v - j h),  j = floor(v / h)   (table S.ty or S.tx).
// Take v = the coordinate with the larger extent and u = the other one.  The ray crosses |cell(u1) - cell(u0)| grid
// lines u = b h; between two crossings it stays inside one line of cells c, where it contributes
// (len/dv) (F_c(v_out) - F_c(v_in)).  Summed over the runs,
//   tau = (len/dv) [ F_c1(v1) - F_c0(v0) + sum_crossings (F_c(v_b) - F_c'(v_b)) ],   c -> c' the line change at v_b,
// one independent term per crossed line of the MINOR axis (on average 3.4x fewer terms than cells crossed, none at
// all for axis-aligned rays), each two 16-byte table reads and a handful of FMAs.  Exact for the piecewise-constant
// interpolant like the cell walk (the reference's bisection + one-point Gauss rule, KernelFactory.cpp:67-190).
// CLAMP: the cell index is clamped to [0, sz - 1].  The reference's getRow/getCol are a bare int(floor(v * sz))
// (KernelFactory.cpp:392-401): for a coordinate of exactly 1 they index cell sz, one past the grid (undefined behaviour there;
// the compiled reference crashes on such probe points).  Here a caller-supplied point on the upper edge is counted to the
// last cell instead; the probe entry points use this form.  Every point the matvec path itself evaluates - quadrature
// nodes, Chebyshev nodes of tree boxes, which lie inside the bounding box of the nodes - is at least half a cell inside the
// square, where round(v*sz - 0.5) is in [0, sz - 1] by itself; the evaluator instantiates CLAMP = false (the double
// fmin / fmax pair was 15 % of its instructions, profiles/r02_ncu_raw.csv + source page).
template <bool CLAMP>
__device__ __forceinline__ double magic_floor(double v, double szd, double top, int& j) {
    // t = round(v*sz - 0.5) = floor(v*sz) away from integers; at an integer either neighbour is fine (F is continuous)
    const double MAGIC = 6755399441055744.0;  // 2^52 + 2^51: the integer sits in the low mantissa bits
    double t = fma(v, szd, -0.5) + MAGIC;     // (MAGIC - 0.5 is not representable: the shift needs its own rounding)
    if (CLAMP) t = fmin(fmax(t, MAGIC), top);
    j = __double2loint(t);
    return t - MAGIC;
}

// crossing points lie between the two (clamped) end points: no clamp needed (the tables carry one entry of padding for the
// one-ulp case v == 1)
// CLAMP (probe points only): a ray that ends on the upper edge v == 1 exactly AND on a grid line of its minor axis has its
// last crossing at the end point itself, where the unclamped index would be sz
template <bool CLAMP>
__device__ __forceinline__ double magic_floor_inner(double v, double szd, double top, int& j) {
    const double MAGIC = 6755399441055744.0;
    double t = fma(v, szd, -0.5) + MAGIC;
    if (CLAMP) t = fmin(t, top);
    j = __double2loint(t);
    return t - MAGIC;
}

template <bool CLAMP>
__device__ __forceinline__ double line_integral_d1(const SigmaGrid& S, double x0, double y0, double x1, double y1, double len) {
    const double ddx = x1 - x0, ddy = y1 - y0;
    const bool steep = fabs(ddy) >= fabs(ddx);
    const double u0 = steep ? x0 : y0, v0 = steep ? y0 : x0, u1 = steep ? x1 : y1, v1 = steep ? y1 : x1;
    const double du = u1 - u0, dv = v1 - v0;
    const double2* __restrict__ tab = steep ? S.ty : S.tx;
    const int sz = S.sz;
    const double szd = (double)sz, h = S.h, top = 6755399441055744.0 + (double)(sz - 1);
    int c0, c1, j0, j1;
    magic_floor<CLAMP>(u0, szd, top, c0);
    magic_floor<CLAMP>(u1, szd, top, c1);
    const double jd0 = magic_floor<CLAMP>(v0, szd, top, j0), jd1 = magic_floor<CLAMP>(v1, szd, top, j1);
    const double2 e0 = __ldg(tab + (size_t)c0 * sz + j0), e1 = __ldg(tab + (size_t)c1 * sz + j1);
    double acc = (e1.x - e0.x) + (fma(e1.y, fma(-jd1, h, v1), -e0.y * fma(-jd0, h, v0)));
    if (c0 != c1) {
        // Crossing from line c to c + 1 at v_b contributes F_c(v_b) - F_{c+1}(v_b) = D[c][j].x + D[c][j].y (v_b - j h) with the
        // difference table D[c][j] = tab[c][j] - tab[c+1][j] (built with the tables): ONE 16-byte read per crossing.  Going
        // down (c -> c - 1) the term is -D[c-1][j]; the sign is applied once to the sum.
        const double slope = dv * __drcp_rn(du);  // crossing positions tolerate the extra ulp
        const int sgn = c1 > c0 ? 1 : -1, ncross = abs(c1 - c0);
        double bd = (double)(c1 > c0 ? c0 + 1 : c0);
        const double sd = (double)sgn;
        const double2* __restrict__ row = (steep ? S.dty : S.dtx) + (size_t)(c1 > c0 ? c0 : c0 - 1) * sz;
        const ptrdiff_t rstep = (ptrdiff_t)sgn * sz;
        double cr0 = 0, cr1 = 0;
        int k = 0;
        for (; k + 2 <= ncross; k += 2) {  // two crossings per trip: two independent table reads in flight
            const double vb0 = fma(fma(bd, h, -u0), slope, v0), vb1 = fma(fma(bd + sd, h, -u0), slope, v0);
            int ja, jb;
            const double jda = magic_floor_inner<CLAMP>(vb0, szd, top, ja), jdb = magic_floor_inner<CLAMP>(vb1, szd, top, jb);
            const double2 d0 = __ldg(row + ja), d1 = __ldg(row + rstep + jb);
            cr0 += fma(d0.y, fma(-jda, h, vb0), d0.x);
            cr1 += fma(d1.y, fma(-jdb, h, vb1), d1.x);
            bd += 2.0 * sd;
            row += 2 * rstep;
        }
        if (k < ncross) {
            const double vb0 = fma(fma(bd, h, -u0), slope, v0);
            int ja;
            const double jda = magic_floor_inner<CLAMP>(vb0, szd, top, ja);
            const double2 d0 = __ldg(row + ja);
            cr0 += fma(d0.y, fma(-jda, h, vb0), d0.x);
        }
        acc = fma(sd, cr0 + cr1, acc);
    }
    return acc * (len * __drcp_rn(dv));
}

// Kernel family of makeKernels (KernelFactory.cpp:240-267): a = source, b = target (Q8).
//   REAL: cos(k theta)/d        IMAG: (exp(-tau)-1) cos(k theta)/d, diagonal +sigma_t(a) for k = 0 (Q2)
//   BOTH: their sum exp(-tau) cos(k theta)/d (the physical kernel; what the fused matvec caches)
__device__ __forceinline__ double cos_k_theta(int k, double c, double s) {  // Re (c + i s)^k
    double cr = 1.0, ci = 0.0;
    for (int i = 0; i < k; ++i) {
        const double nr = cr * c - ci * s;
        ci = cr * s + ci * c;
        cr = nr;
    }
    return cr;
}

// exp(x) for -700 <= x <= 700 (the attenuation e^{-tau}: tau is an optical depth of order 1 - 100).  libm's exp spent two
// thirds of its instructions on range and special-case handling this argument never needs.  n = round(x / ln 2) by the
// magic-number shift, r = x - n ln 2 in two FMAs (ln 2 split hi + lo), e^r by the degree-13 Taylor polynomial, even and odd powers as
// two Horner chains in r^2 (|r| <= 0.347: truncation r^14 / 14! < 5e-18), scaled by 2^n through the exponent field.  Within 2 ulp of the correctly
// rounded value, i.e. at rounding level against the reference's libm (parity gate 1e-10); below -700 the result is 0.
__device__ __forceinline__ double exp_reduced(double x) {
    const double MAGIC = 6755399441055744.0;
    const double t = fma(x, 1.4426950408889634074, MAGIC);
    const int n = __double2loint(t);
    const double nd = t - MAGIC;
    double r = fma(nd, -6.93147180559945286227e-01, x);
    r = fma(nd, -2.31904681384629955842e-17, r);
    // even and odd powers as two independent Horner chains in r^2 (dependency depth 8 instead of 13: the evaluator runs 8
    // warps per scheduler and waits on dependent FP64 results)
    const double r2 = r * r;
    double pe = 2.0876756987868098979e-09;           // 1/12!
    double po = 1.6059043836821614599e-10;           // 1/13!
    pe = fma(pe, r2, 2.7557319223985890653e-07);     // 1/10!
    po = fma(po, r2, 2.5052108385441718775e-08);     // 1/11!
    pe = fma(pe, r2, 2.4801587301587301566e-05);     // 1/8!
    po = fma(po, r2, 2.7557319223985892511e-06);     // 1/9!
    pe = fma(pe, r2, 1.3888888888888889419e-03);     // 1/6!
    po = fma(po, r2, 1.9841269841269841253e-04);     // 1/7!
    pe = fma(pe, r2, 4.1666666666666664354e-02);     // 1/4!
    po = fma(po, r2, 8.3333333333333332177e-03);     // 1/5!
    pe = fma(pe, r2, 0.5);
    po = fma(po, r2, 1.6666666666666665741e-01);     // 1/3!
    pe = fma(pe, r2, 1.0);
    po = fma(po, r2, 1.0);
    const double p = fma(po, r, pe);
    const double e = __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
    return x < -700.0 ? 0.0 : e;
}

// D1: the deg-1 table path is known at compile time (keeps the general cell walk and its registers out of the kernel)
// CLAMP: see magic_floor (false on the evaluator's own points)
template <bool D1, bool CLAMP = true>
__device__ __forceinline__ double kernel_eval_t(const SigmaGrid& S, int id, int part, double ax, double ay, double bx, double by) {
    const double ddx = ax - bx, ddy = ay - by;
    const double d2 = ddx * ddx + ddy * ddy;
    if (d2 == 0.0) return ((part & PART_IMAG) && id == 0) ? sigma_eval(S, ax, ay) : 0.0;
    double dist, inv;
    if (D1) {  // rsqrt + one Newton step on the product (<= 1 ulp from the IEEE pair, far inside the 1e-10 gate)
        inv = rsqrt(d2);
        dist = d2 * inv;
        inv = fma(fma(-dist, inv, 1.0), inv, inv);
    } else {
        dist = sqrt(d2);
        inv = 1.0 / dist;
    }
    const double ck = id == 0 ? 1.0 : cos_k_theta(id, ddx * inv, ddy * inv);
    if (part == PART_REAL) return ck * inv;
    const double tau = D1 ? line_integral_d1<CLAMP>(S, ax, ay, bx, by, dist) : line_integral(S, ax, ay, bx, by);
    const double e = D1 ? exp_reduced(-tau) : exp(-tau);
    if (part == PART_IMAG) return (e - 1.0) * ck * inv;
    return e * ck * inv;
}
__device__ double kernel_eval(const SigmaGrid& S, int id, int part, double ax, double ay, double bx, double by) {
    return S.ty ? kernel_eval_t<true>(S, id, part, ax, ay, bx, by) : kernel_eval_t<false>(S, id, part, ax, ay, bx, by);
}

// ------------------------------------------------------------------------------------------------
// K1/K2 evaluator.  One CTA per operator block; threads stride over the block column-major in the target
// index so the rays of one warp join neighbouring points of the same two boxes (similar step counts).
// ------------------------------------------------------------------------------------------------
// NR > 0: every block is NR x NR (index arithmetic by shifts)
// A CTA evaluates EPC consecutive blocks.  The metadata of all of them (entry -> nodes -> box geometry or point offsets,
// a chain of dependent global loads) is fetched once by the first EPC threads and parked in shared memory, so the chain
// is paid once per EPC * 256 interactions instead of once per block.
static constexpr int EPC = 8;
template <bool D1, int NR>
__global__ void __launch_bounds__(256, D1 ? 4 : 2) fill_blocks_kernel(DevCSR m, DevTree T, DevGeom G, SigmaGrid S, int id, int part,
                                                                      int cheb, double* __restrict__ A) {
    __shared__ double s_geo[EPC][8];  // cheb: target cx, cy, rx, ry, source cx, cy, rx, ry
    __shared__ int s_off[EPC][4];     // points: toff, soff, nrow, ncol
    __shared__ int64_t s_blk[EPC];
    __shared__ double s_cheb[16];
    const int e_base = blockIdx.x * EPC, cnt = min(EPC, m.entries - e_base);
    const int np = NR == 16 ? 4 : T.np;
    if (threadIdx.x < cnt) {
        const int e = e_base + threadIdx.x;
        const int tn = m.tgt_node[e], sn = m.src_node[e];
        s_blk[threadIdx.x] = m.blk_off[e];
        if (cheb) {
            const double g0 = T.cx[tn], g1 = T.cy[tn], g2 = T.rx[tn], g3 = T.ry[tn];
            const double g4 = T.cx[sn], g5 = T.cy[sn], g6 = T.rx[sn], g7 = T.ry[sn];
            double* __restrict__ gg = s_geo[threadIdx.x];
            gg[0] = g0; gg[1] = g1; gg[2] = g2; gg[3] = g3; gg[4] = g4; gg[5] = g5; gg[6] = g6; gg[7] = g7;
            s_off[threadIdx.x][2] = s_off[threadIdx.x][3] = NR ? NR : T.r;
        } else {
            const int o0 = T.pt_off[tn], o1 = T.pt_off[sn];
            const int n0 = NR ? NR : T.npts[tn], n1 = NR ? NR : T.npts[sn];
            s_off[threadIdx.x][0] = o0; s_off[threadIdx.x][1] = o1; s_off[threadIdx.x][2] = n0; s_off[threadIdx.x][3] = n1;
        }
    } else if (threadIdx.x >= 32 && threadIdx.x < 32 + np && cheb)
        s_cheb[threadIdx.x - 32] = T.cheb[threadIdx.x - 32];
    __syncthreads();
    for (int j = 0; j < cnt; ++j) {
        const int nrow = NR ? NR : s_off[j][2], ncol = NR ? NR : s_off[j][3];  // NR: the block shape is a compile-time constant
        double* __restrict__ out = A + s_blk[j];
        const int total = nrow * ncol;
        for (int idx0 = threadIdx.x; idx0 < total; idx0 += blockDim.x) {
            int idx = idx0;
            int s = NR ? idx / NR : idx / nrow;
            int t = (NR ? idx % NR : idx % nrow) ^ (m.swz ? (s & 3) << 2 : 0);  // idx is the stored position
            if (NR == 16 && cheb) {
                // 4 x 4 Chebyshev nodes per box: the number of grid lines a ray crosses along its minor axis - the trip count
                // of the line integral - is fixed by the two node indices along that axis.  Give the 16 lanes of a half-warp
                // the same pair of minor-axis indices (and all 16 pairs of major-axis indices), so that the lanes of a warp
                // leave the crossing loop together; the minor axis is taken from the box centres.
                const double* __restrict__ gg = s_geo[j];
                const bool xmajor = fabs(gg[4] - gg[0]) >= fabs(gg[5] - gg[1]);
                const int hi = (idx0 >> 4) & 15, lo = idx0 & 15;  // hi: constant over a half-warp
                const int mn = xmajor ? hi : lo, mj = xmajor ? lo : hi;  // (source, target) index pairs along minor / major
                const int sj = xmajor ? mn >> 2 : mj >> 2, tj = xmajor ? mn & 3 : mj & 3;  // y indices
                const int si = xmajor ? mj >> 2 : mn >> 2, ti = xmajor ? mj & 3 : mn & 3;  // x indices
                s = sj * 4 + si;
                t = tj * 4 + ti;
                idx = s * 16 + (t ^ (m.swz ? (s & 3) << 2 : 0));
            }
            double ax, ay, bx, by;
            if (cheb) {  // flat = j*np + i, y slow (bbfmm.h:787-800)
                const double* __restrict__ gg = s_geo[j];
                ax = gg[4] + gg[6] * s_cheb[s % np];
                ay = gg[5] + gg[7] * s_cheb[s / np];
                bx = gg[0] + gg[2] * s_cheb[t % np];
                by = gg[1] + gg[3] * s_cheb[t / np];
            } else {
                const int toff = s_off[j][0], soff = s_off[j][1];
                ax = G.px[soff + s];
                ay = G.py[soff + s];
                bx = G.px[toff + t];
                by = G.py[toff + t];
            }
            out[idx] = kernel_eval_t<D1, false>(S, id, part, ax, ay, bx, by);  // nodes / Chebyshev nodes: interior points
        }
    }
}

// Operators of kernel id k > 0 from those of id 0 (the optical depth does not depend on the id):
//   K_k = exp(-tau) cos(k theta) / d = K_0 * cos(k theta),  diagonal 0 (makeKernels, KernelFactory.cpp:240-267),
// one streaming pass over K_0 instead of a second ray march per entry.  Same CTA layout as fill_blocks_kernel.
__global__ void __launch_bounds__(256) fill_from_base_kernel(DevCSR m, DevTree T, DevGeom G, int id, int cheb,
                                                             const double* __restrict__ A0, double* __restrict__ A) {
    __shared__ double s_geo[EPC][8];
    __shared__ int s_off[EPC][4];
    __shared__ int64_t s_blk[EPC];
    __shared__ double s_cheb[16];
    const int e_base = blockIdx.x * EPC, cnt = min(EPC, m.entries - e_base);
    const int np = T.np;
    if (threadIdx.x < cnt) {
        const int e = e_base + threadIdx.x;
        const int tn = m.tgt_node[e], sn = m.src_node[e];
        s_blk[threadIdx.x] = m.blk_off[e];
        if (cheb) {
            double* __restrict__ gg = s_geo[threadIdx.x];
            gg[0] = T.cx[tn]; gg[1] = T.cy[tn]; gg[2] = T.rx[tn]; gg[3] = T.ry[tn];
            gg[4] = T.cx[sn]; gg[5] = T.cy[sn]; gg[6] = T.rx[sn]; gg[7] = T.ry[sn];
            s_off[threadIdx.x][2] = s_off[threadIdx.x][3] = T.r;
        } else {
            s_off[threadIdx.x][0] = T.pt_off[tn]; s_off[threadIdx.x][1] = T.pt_off[sn];
            s_off[threadIdx.x][2] = T.npts[tn]; s_off[threadIdx.x][3] = T.npts[sn];
        }
    } else if (threadIdx.x >= 32 && threadIdx.x < 32 + np && cheb)
        s_cheb[threadIdx.x - 32] = T.cheb[threadIdx.x - 32];
    __syncthreads();
    for (int j = 0; j < cnt; ++j) {
        const int nrow = s_off[j][2], ncol = s_off[j][3];
        const double* __restrict__ in = A0 + s_blk[j];
        double* __restrict__ out = A + s_blk[j];
        const int total = nrow * ncol;
        for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
            const double base = in[idx];
            const int s = idx / nrow, t = (idx % nrow) ^ (m.swz ? (s & 3) << 2 : 0);
            double ax, ay, bx, by;
            if (cheb) {
                const double* __restrict__ gg = s_geo[j];
                ax = gg[4] + gg[6] * s_cheb[s % np];
                ay = gg[5] + gg[7] * s_cheb[s / np];
                bx = gg[0] + gg[2] * s_cheb[t % np];
                by = gg[1] + gg[3] * s_cheb[t / np];
            } else {
                const int toff = s_off[j][0], soff = s_off[j][1];
                ax = G.px[soff + s]; ay = G.py[soff + s];
                bx = G.px[toff + t]; by = G.py[toff + t];
            }
            const double ddx = ax - bx, ddy = ay - by, d2 = ddx * ddx + ddy * ddy;
            double v = 0.0;
            if (d2 != 0.0) {
                double inv = rsqrt(d2);
                inv = fma(fma(-d2 * inv, inv, 1.0), 0.5 * inv, inv);
                v = base * cos_k_theta(id, ddx * inv, ddy * inv);
            }
            out[idx] = v;
        }
    }
}
// The same for uniform 16 x 16 blocks (every BASELINE configuration at np 4: M2L blocks between 4 x 4 Chebyshev grids, P2P
// tiles between 16-point leaves).  256 threads = one block entry each, so a thread's (source, target) position - and for
// the Chebyshev case its four node offsets - are the same in every block and leave the loop; the EPC id-0 values a thread
// needs are requested before any arithmetic (EPC loads in flight per thread instead of one).  The generic kernel ran at
// 2.3 TB/s, bound by its index arithmetic (idx / nrow, % np per entry).  Same expressions, same values.
template <bool CHEB>
__global__ void __launch_bounds__(256) fill_from_base16_kernel(DevCSR m, DevTree T, DevGeom G, int id, const double* __restrict__ A0,
                                                               double* __restrict__ A) {
    __shared__ double s_geo[EPC][8];
    __shared__ int s_off[EPC][2];
    __shared__ int64_t s_blk[EPC];
    __shared__ double s_cheb[4];
    const int e_base = blockIdx.x * EPC, cnt = min(EPC, m.entries - e_base);
    if (threadIdx.x < cnt) {
        const int e = e_base + threadIdx.x;
        const int tn = m.tgt_node[e], sn = m.src_node[e];
        s_blk[threadIdx.x] = m.blk_off[e];
        if (CHEB) {
            double* __restrict__ gg = s_geo[threadIdx.x];
            gg[0] = T.cx[tn]; gg[1] = T.cy[tn]; gg[2] = T.rx[tn]; gg[3] = T.ry[tn];
            gg[4] = T.cx[sn]; gg[5] = T.cy[sn]; gg[6] = T.rx[sn]; gg[7] = T.ry[sn];
        } else {
            s_off[threadIdx.x][0] = T.pt_off[tn];
            s_off[threadIdx.x][1] = T.pt_off[sn];
        }
    } else if (CHEB && threadIdx.x >= 32 && threadIdx.x < 36)
        s_cheb[threadIdx.x - 32] = T.cheb[threadIdx.x - 32];
    __syncthreads();
    const int idx = threadIdx.x, sI = idx >> 4, tI = (idx & 15) ^ (m.swz ? (sI & 3) << 2 : 0);
    double base[EPC];
#pragma unroll
    for (int j = 0; j < EPC; ++j) base[j] = j < cnt ? A0[s_blk[j] + idx] : 0.0;
    const double csx = CHEB ? s_cheb[sI & 3] : 0.0, csy = CHEB ? s_cheb[sI >> 2] : 0.0;
    const double ctx = CHEB ? s_cheb[tI & 3] : 0.0, cty = CHEB ? s_cheb[tI >> 2] : 0.0;
#pragma unroll
    for (int j = 0; j < EPC; ++j) {
        if (j >= cnt) break;
        double ax, ay, bx, by;
        if (CHEB) {
            const double* __restrict__ gg = s_geo[j];
            ax = gg[4] + gg[6] * csx;
            ay = gg[5] + gg[7] * csy;
            bx = gg[0] + gg[2] * ctx;
            by = gg[1] + gg[3] * cty;
        } else {
            const int toff = s_off[j][0], soff = s_off[j][1];
            ax = G.px[soff + sI]; ay = G.py[soff + sI];
            bx = G.px[toff + tI]; by = G.py[toff + tI];
        }
        const double ddx = ax - bx, ddy = ay - by, d2 = ddx * ddx + ddy * ddy;
        double v = 0.0;
        if (d2 != 0.0) {
            double inv = rsqrt(d2);
            inv = fma(fma(-d2 * inv, inv, 1.0), 0.5 * inv, inv);
            v = base[j] * cos_k_theta(id, ddx * inv, ddy * inv);
        }
        A[s_blk[j] + idx] = v;
    }
}

void launch_fill_from_base(const DevCSR& m, const DevTree& T, const DevGeom& G, int id, int cheb, const double* A0, double* A,
                           cudaStream_t st) {
    if (m.entries == 0) return;
    const unsigned ctas = (m.entries + EPC - 1) / EPC;
    static const bool generic = getenv("ANISO_FROM_BASE_GENERIC") != nullptr;  // A/B switch
    if (m.uniform_block == 16 && !generic && (!cheb || T.np == 4)) {
        if (cheb) fill_from_base16_kernel<true><<<ctas, 256, 0, st>>>(m, T, G, id, A0, A);
        else fill_from_base16_kernel<false><<<ctas, 256, 0, st>>>(m, T, G, id, A0, A);
        return;
    }
    fill_from_base_kernel<<<ctas, 256, 0, st>>>(m, T, G, id, cheb, A0, A);
}

void launch_fill(const DevCSR& m, const DevTree& T, const DevGeom& G, const SigmaGrid& S, int id, int part, int cheb,
                 double* A, cudaStream_t st) {
    if (m.entries == 0) return;
    int threads = 256;
    if (cheb && T.r <= 8) threads = 64;
    else if (cheb && T.r <= 11) threads = 128;
    const unsigned ctas = (m.entries + EPC - 1) / EPC;
    if (S.ty && m.uniform_block == 16) fill_blocks_kernel<true, 16><<<ctas, threads, 0, st>>>(m, T, G, S, id, part, cheb, A);
    else if (S.ty) fill_blocks_kernel<true, 0><<<ctas, threads, 0, st>>>(m, T, G, S, id, part, cheb, A);
    else fill_blocks_kernel<false, 0><<<ctas, threads, 0, st>>>(m, T, G, S, id, part, cheb, A);
}

// ------------------------------------------------------------------------------------------------
// Cache-free P2P (kernel::run -> downPass, bbfmm.h:507-532, 863-946: the reference's uncached path evaluates every U/W
// entry during the apply).  For memory-constrained runs: the near-field tiles are not stored; the warp of a
// target leaf evaluates the tiles of its stored pairs with the same kernel_eval_t as the evaluator and applies them at once,
// in both directions under symmetric storage - its own sums to its first scratch row, K^T q_A to the partner's mirror row -
// so the downward pass folds exactly what the cached stream kernel would have left.  Saves 8 B per stored pair and kernel
// id (0.67 GB per id at N = 2^20) for ~12 evaluations per target and matvec.
// Lane map: TPS = min(32, block width) lanes share a source point; 32 / TPS source points per trip.
// ------------------------------------------------------------------------------------------------
template <bool D1>
__global__ void __launch_bounds__(256) p2p_onfly_kernel(DevCSR m, DevGeom G, SigmaGrid S, int id, int part, const double* __restrict__ q,
                                                        double* __restrict__ y, double sign) {
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (g >= m.groups) return;
    const int nr = m.nrow[g], toff = m.row_off[g];
    const int tps = nr >= 32 ? 32 : nr > 16 ? 32 : nr > 8 ? 16 : nr > 4 ? 8 : nr > 2 ? 4 : nr > 1 ? 2 : 1;  // power of two >= nr (<= 32)
    const int rp = 32 / tps, sub = lane / tps, tl = lane % tps;
    // up to two targets per lane (leaves hold at most 64 points on the symmetric path; ragged leaves loop below)
    double acc[2] = {0.0, 0.0};
    double tx[2], ty[2], qa[2];
    bool tok[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int t = tl + 32 * u;
        tok[u] = t < nr && (u == 0 || tps == 32);
        tx[u] = tok[u] ? G.px[toff + t] : 0.0;
        ty[u] = tok[u] ? G.py[toff + t] : 0.0;
        qa[u] = tok[u] ? q[toff + t] : 0.0;
    }
    const int e0 = m.ent_ptr[g], e1 = m.ent_ptr[g + 1];
    for (int e = e0; e < e1; ++e) {
        const int soff = m.col_off[e], nc = m.ncol[e];
        const int mir = m.symmetric ? m.mirror_row[e] : -1;
        for (int s0 = 0; s0 < nc; s0 += rp) {
            const int s = s0 + sub;
            const bool sok = s < nc;
            const double sx = sok ? G.px[soff + s] : 0.0, sy = sok ? G.py[soff + s] : 0.0, qs = sok ? q[soff + s] : 0.0;
            double z = 0.0;
#pragma unroll
            for (int u = 0; u < 2; ++u)
                if (sok && tok[u]) {
                    const double k = kernel_eval_t<D1, false>(S, id, part, sx, sy, tx[u], ty[u]);  // a = source, b = target (Q8)
                    acc[u] = fma(k, qs, acc[u]);
                    z = fma(k, qa[u], z);
                }
            if (mir >= 0) {  // transposed product for the partner: sum over the lanes that share this source point
                for (int off = tps >> 1; off > 0; off >>= 1) z += __shfl_xor_sync(0xffffffffu, z, off);
                if (tl == 0 && sok) m.scratch[(size_t)mir * m.uniform_block + s] = sign * z;
            }
        }
    }
    // the rp sub-groups hold partial sums of the same targets
    for (int off = tps; off < 32; off <<= 1) {
        acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], off);
        acc[1] += __shfl_xor_sync(0xffffffffu, acc[1], off);
    }
    if (sub == 0) {
        double* __restrict__ out = m.symmetric ? m.scratch + (size_t)m.srow_ptr[g] * m.uniform_block : y + toff;
#pragma unroll
        for (int u = 0; u < 2; ++u)
            if (tok[u]) out[tl + 32 * u] = acc[u];
    }
}

bool p2p_onfly_ok(const DevCSR& m, int max_leaf) { return max_leaf <= 64; }

void launch_p2p_onfly(const DevCSR& m, const DevGeom& G, const SigmaGrid& S, int id, int part, const double* q, double* y, double sign,
                      cudaStream_t st) {
    if (m.groups == 0) return;
    const unsigned ctas = (m.groups + 7) / 8;
    if (S.ty) p2p_onfly_kernel<true><<<ctas, 256, 0, st>>>(m, G, S, id, part, q, y, sign);
    else p2p_onfly_kernel<false><<<ctas, 256, 0, st>>>(m, G, S, id, part, q, y, sign);
}

__global__ void probe_line_integral_kernel(SigmaGrid S, int n, const double* __restrict__ p, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x0 = p[4 * i], y0 = p[4 * i + 1], x1 = p[4 * i + 2], y1 = p[4 * i + 3];
    const double len = sqrt((x1 - x0) * (x1 - x0) + (y1 - y0) * (y1 - y0));
    out[i] = (S.ty && len > 0.0) ? line_integral_d1<true>(S, x0, y0, x1, y1, len) : line_integral(S, x0, y0, x1, y1);
}
__global__ void probe_kernel_kernel(SigmaGrid S, int id, int part, int n, const double* __restrict__ p,
                                    double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = kernel_eval(S, id, part, p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]);
}
void launch_probe_line_integral(const SigmaGrid& S, int n, const double* pts4, double* out, cudaStream_t st) {
    if (n > 0) probe_line_integral_kernel<<<(n + 127) / 128, 128, 0, st>>>(S, n, pts4, out);
}
void launch_probe_kernel(const SigmaGrid& S, int id, int part, int n, const double* pts4, double* out, cudaStream_t st) {
    if (n > 0) probe_kernel_kernel<<<(n + 127) / 128, 128, 0, st>>>(S, id, part, n, pts4, out);
}

// ------------------------------------------------------------------------------------------------
// K7 + K8 + K9 corrections.  All three are linear in the (unscaled) source with support in the 3x3 cell
// neighbourhood of the target, so each is stored once per kernel id as a stencil
//     C[(cs*c2 + q)*nown + k],  cs = (dr+1)*3 + (dc+1),  k = target slot - slot0
// acting on f[(cell + dr*sz + dc)*c2 + q].  The reference instead keeps nearInteractions (N x 9 x 16 deg^2) and
// singInteractions (N x 8 sr^2) and redoes the small gemv / Legendre sums per target on every apply
// (KernelFactory.cpp:445-478, 550-609, 662-709, 752-788, 828-860).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double real_kernel(int id, double ax, double ay, double bx, double by) {
    const double ddx = ax - bx, ddy = ay - by;
    const double d2 = ddx * ddx + ddy * ddy;
    if (d2 == 0.0) return 0.0;
    double inv = rsqrt(d2);  // + one Newton step on the product: within an ulp of 1/sqrt(d2), 4x fewer instructions
    inv = fma(fma(-d2 * inv, inv, 1.0), 0.5 * inv, inv);
    if (id == 0) return inv;
    return cos_k_theta(id, ddx * inv, ddy * inv) * inv;
}

__device__ __forceinline__ double legendre_tr1_dev(int l, double x) {
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

__global__ void __launch_bounds__(128) build_stencils_kernel(DevGeom G, int id, int slot0, int nown, double* __restrict__ c7,
                                                             double* __restrict__ c8, double* __restrict__ c9) {
    const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int c2 = G.c2, sz = G.sz, deg = G.deg;
    const long total = (long)9 * c2 * nown;
    if (gid >= total) return;
    const int k = (int)(gid % nown);
    const int cq = (int)(gid / nown);
    const int q = cq % c2, cs = cq / c2;
    const int dr = cs / 3 - 1, dc = cs % 3 - 1;
    const int p = slot0 + k;
    const int t = G.perm[p];
    const int tc = t / c2, tq = t - tc * c2;
    const int ti = tc / sz, tj = tc - ti * sz;  // ti <-> x, tj <-> y
    double v7 = 0, v8 = 0;
    const int si = ti + dr, sj = tj + dc;
    if (si >= 0 && si < sz && sj >= 0 && sj < sz) {
        const double tx = G.px[p], ty = G.py[p], dx = G.dx;
        const int sslot = G.inv_perm[(si * sz + sj) * c2 + q];
        // K7 nearRemoval (KernelFactory.cpp:445-478): self cell included, the self pair evaluates to 0 (Q16).
        // Acts on the SCALED source (f .* weights), like the reference.
        v7 = -real_kernel(id, G.px[sslot], G.py[sslot], tx, ty);
        if (cs != 4) {
            // K8 refineAddOn (KernelFactory.cpp:583-601, 687-700): refined rule of the neighbour cell; acts on the
            // scaled source: old = scaled / sqrtWeights, new = nearMapping * old, ret += inter . new
            double acc = 0;
            for (int a = 0; a < G.rq; ++a) {
                const double ax = (0.5 + si) * dx + 0.5 * (G.rqx[a]) * dx;
                const double ay = (0.5 + sj) * dx + 0.5 * (G.rqy[a]) * dx;
                const double inter = real_kernel(id, ax, ay, tx, ty) * sqrt(G.rqw[a]);
                acc += inter * G.nearMap[(size_t)a * c2 + q];
            }
            v8 = acc / G.sqrtw[q];
        } else {
            // K9 singularAdd (KernelFactory.cpp:765-780, 840-856): Duffy points of the self cell; this thread owns the
            // Legendre mode nk = q.  Acts on the Legendre coefficients of the UNSCALED source, whose interpolant is
            // evaluated at GLOBAL coordinates (Q1).
            const double* __restrict__ sX = G.singX + (size_t)tq * G.nsing;
            const double* __restrict__ sY = G.singY + (size_t)tq * G.nsing;
            const double* __restrict__ sW = G.singW + (size_t)tq * G.nsing;
            const int n = q / deg, kk = q - n * deg;
            const double nrm = G.lnorm[q];
            const double xc = (0.5 + ti) * dx, yc = (0.5 + tj) * dx;
            double acc = 0;
            if (deg == 1) {  // P_0 = 1: the 8 sr^2 Duffy points only carry the kernel and the weight
                double acc1 = 0;
                int pp = 0;
                for (; pp + 2 <= G.nsing; pp += 2) {
                    const double x0 = xc + 0.5 * sX[pp] * dx, y0 = yc + 0.5 * sY[pp] * dx;
                    const double x1 = xc + 0.5 * sX[pp + 1] * dx, y1 = yc + 0.5 * sY[pp + 1] * dx;
                    acc += (1.0 / nrm) * (real_kernel(id, x0, y0, tx, ty) * (sW[pp] * (dx * dx) / 4.0));
                    acc1 += (1.0 / nrm) * (real_kernel(id, x1, y1, tx, ty) * (sW[pp + 1] * (dx * dx) / 4.0));
                }
                for (; pp < G.nsing; ++pp)
                    acc += (1.0 / nrm) * (real_kernel(id, xc + 0.5 * sX[pp] * dx, yc + 0.5 * sY[pp] * dx, tx, ty) * (sW[pp] * (dx * dx) / 4.0));
                acc += acc1;
            } else {
                for (int pp = 0; pp < G.nsing; ++pp) {
                    const double x = xc + 0.5 * (sX[pp]) * dx;
                    const double y = yc + 0.5 * (sY[pp]) * dx;
                    const double w = sW[pp] * (dx * dx) / 4.0;
                    const double inter = real_kernel(id, x, y, tx, ty) * w;
                    acc += (legendre_tr1_dev(n, x) * legendre_tr1_dev(kk, y) / nrm) * inter;
                }
            }
            c9[(size_t)q * nown + k] = acc;
        }
    }
    c7[gid] = v7;
    c8[gid] = v8;
}

// fused stencil acting on the UNSCALED source f:  (K7 + K8) .* weight  +  K9 o interpolation
__global__ void combine_stencils_kernel(DevGeom G, int slot0, int nown, const double* __restrict__ c7, const double* __restrict__ c8,
                                        const double* __restrict__ c9, double* __restrict__ call) {
    const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int c2 = G.c2, sz = G.sz;
    if (gid >= (long)9 * c2 * nown) return;
    const int k = (int)(gid % nown);
    const int cq = (int)(gid / nown);
    const int q = cq % c2, cs = cq / c2;
    const int t = G.perm[slot0 + k];
    const int tc = t / c2;
    const int ti = tc / sz, tj = tc - ti * sz;
    const int si = ti + cs / 3 - 1, sj = tj + cs % 3 - 1;
    double v = 0;
    if (si >= 0 && si < sz && sj >= 0 && sj < sz) {
        v = (c7[gid] + c8[gid]) * G.pw[G.inv_perm[(si * sz + sj) * c2 + q]];
        if (cs == 4) {
            double a = 0;
            for (int nk = 0; nk < c2; ++nk) a += c9[(size_t)nk * nown + k] * G.interp[(size_t)nk * c2 + q];
            v += a * G.sqrtw[q];
        }
    }
    call[gid] = v;
}

void launch_build_stencils(const DevGeom& G, int id, int slot0, int nown, double* c7, double* c8, double* c9, cudaStream_t st) {
    const long total = (long)9 * G.c2 * nown;
    if (total == 0) return;
    build_stencils_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(G, id, slot0, nown, c7, c8, c9);
}
void launch_combine_stencils(const DevGeom& G, int slot0, int nown, const double* c7, const double* c8, const double* c9,
                             double* call, cudaStream_t st) {
    const long total = (long)9 * G.c2 * nown;
    if (total == 0) return;
    combine_stencils_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(G, slot0, nown, c7, c8, c9, call);
}

// K9 stage form: acc[k] += sum_nk c9[nk][k] * coeff[cell(k)][nk]  (singularAddFast, KernelFactory.cpp:828-860)
__global__ void sing_apply_kernel(DevGeom G, int slot0, int nown, const double* __restrict__ c9, const double* __restrict__ coeff,
                                  double* __restrict__ acc_out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nown) return;
    const int c2 = G.c2;
    const int tc = G.perm[slot0 + k] / c2;
    double acc = 0;
    for (int nk = 0; nk < c2; ++nk) acc = fma(c9[(size_t)nk * nown + k], coeff[(size_t)tc * c2 + nk], acc);
    acc_out[k] += acc;
}
void launch_sing_apply(const DevGeom& G, int slot0, int nown, const double* c9, const double* coeff, double* acc, cudaStream_t st) {
    if (nown) sing_apply_kernel<<<(nown + 127) / 128, 128, 0, st>>>(G, slot0, nown, c9, coeff, acc);
}

__global__ void stencil_apply_kernel(DevGeom G, int slot0, int nown, const double* __restrict__ C, const double* __restrict__ f,
                                     double* __restrict__ acc_out, int accumulate) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nown) return;
    const int c2 = G.c2, sz = G.sz;
    const int t = G.perm[slot0 + k];
    const int tc = t / c2;
    const int ti = tc / sz, tj = tc - ti * sz;
    double acc = 0;
    for (int cs = 0; cs < 9; ++cs) {
        const int si = ti + cs / 3 - 1, sj = tj + cs % 3 - 1;
        if (si < 0 || si >= sz || sj < 0 || sj >= sz) continue;
        const int base = (si * sz + sj) * c2;
        for (int q = 0; q < c2; ++q)
            acc = fma(C[(size_t)(cs * c2 + q) * nown + k], f[G.inv_perm[base + q]], acc);
    }
    acc_out[k] = accumulate ? acc_out[k] + acc : acc;
}

void launch_stencil_apply(const DevGeom& G, int slot0, int nown, const double* C, const double* f, double* acc, int accumulate,
                          cudaStream_t st) {
    if (nown == 0) return;
    stencil_apply_kernel<<<(nown + 127) / 128, 128, 0, st>>>(G, slot0, nown, C, f, acc, accumulate);
}

__global__ void finalize_kernel(int n, const double* __restrict__ fmm, const double* __restrict__ corr, const double* __restrict__ x,
                                double* __restrict__ out, int subtract_from_x) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double v = (fmm[i] + (corr ? corr[i] : 0.0)) * (M_1_PI / 2.0);  // AnisoWrapper.cpp:129-130
    out[i] = subtract_from_x ? x[i] - v : v;                              // main.cpp:129-134
}
void launch_finalize(int n, const double* fmm, const double* corr, const double* x, double* out, int subtract_from_x,
                     cudaStream_t st) {
    if (n) finalize_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, fmm, corr, x, out, subtract_from_x);
}

__global__ void gather_kernel(int n, const double* __restrict__ src, const int* __restrict__ idx, double* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}
__global__ void scatter_kernel(int n, const double* __restrict__ src, const int* __restrict__ idx, double* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[idx[i]] = src[i];
}
void launch_scatter(int n, const double* src, const int* idx, double* dst, cudaStream_t st) {
    if (n) scatter_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, src, idx, dst);
}
void launch_gather(int n, const double* src, const int* idx, double* dst, cudaStream_t st) {
    if (n) gather_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, src, idx, dst);
}

__global__ void add3_kernel(long n, const double* __restrict__ a, const double* __restrict__ b, const double* __restrict__ c,
                            double* __restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = a[i] + b[i] + c[i];
}
void launch_add3(long n, const double* a, const double* b, const double* c, double* out, cudaStream_t st) {
    if (n) add3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, a, b, c, out);
}

// K10 interpolation (KernelFactory.cpp:212-227, 988-1005): per cell a c2 x c2 gemv; h is in REFERENCE node order
__global__ void interpolate_kernel(DevGeom G, const double* __restrict__ h, const double* __restrict__ lnorm,
                                   double* __restrict__ coeff) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= G.N) return;
    const int c2 = G.c2, c = i / c2, r = i - c * c2;
    double s = 0;
    for (int j = 0; j < c2; ++j) s += G.interp[(size_t)r * c2 + j] * (G.sqrtw[j] * h[(size_t)c * c2 + j]);
    coeff[i] = lnorm ? s / lnorm[r] : s;
}
void launch_interpolate(const DevGeom& G, const double* h_ref, const double* lnorm_or_null, double* coeff, cudaStream_t st) {
    interpolate_kernel<<<(G.N + 255) / 256, 256, 0, st>>>(G, h_ref, lnorm_or_null, coeff);
}

// Running integrals of the piecewise-constant sigma_t along every grid line (deg == 1): one thread per line.
__global__ void sigma_prefix_kernel(SigmaGrid S, double2* __restrict__ ty, double2* __restrict__ tx) {
    const int line = blockIdx.x * blockDim.x + threadIdx.x;
    const int sz = S.sz;
    if (line >= 2 * sz) return;
    double run = 0;
    if (line < sz) {  // column cx = line, along y
        for (int j = 0; j < sz; ++j) {
            const double sg = S.chat[(size_t)line * sz + j];
            ty[(size_t)line * sz + j] = make_double2(run, sg);
            run = fma(sg, S.h, run);
        }
    } else {  // row cy, along x
        const int cy = line - sz;
        for (int i = 0; i < sz; ++i) {
            const double sg = S.chat[(size_t)i * sz + cy];
            tx[(size_t)cy * sz + i] = make_double2(run, sg);
            run = fma(sg, S.h, run);
        }
    }
}
// Difference tables of neighbouring lines, D[c][j] = t[c][j] - t[c+1][j] (line sz-1 has no upper neighbour: zeros)
__global__ void sigma_diff_kernel(int sz, const double2* __restrict__ ty, const double2* __restrict__ tx, double2* __restrict__ dty,
                                  double2* __restrict__ dtx) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)sz * (sz - 1)) return;
    const double2 a = ty[i], b = ty[i + sz], c = tx[i], d = tx[i + sz];
    dty[i] = make_double2(a.x - b.x, a.y - b.y);
    dtx[i] = make_double2(c.x - d.x, c.y - d.y);
}
void launch_sigma_prefix(const SigmaGrid& S, double2* ty, double2* tx, cudaStream_t st) {
    sigma_prefix_kernel<<<(2 * S.sz + 63) / 64, 64, 0, st>>>(S, ty, tx);
    if (S.sz > 1) {
        const size_t n = (size_t)S.sz * (S.sz - 1);
        sigma_diff_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(S.sz, ty, tx, const_cast<double2*>(S.dty),
                                                                      const_cast<double2*>(S.dtx));
    }
}

}  // namespace an
