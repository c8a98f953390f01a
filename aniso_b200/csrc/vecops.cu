// K11: device-resident BLAS-1 pieces of the GMRES Arnoldi step (gmres.cpp:115-122, 12-24).  All scalars that
// feed the next kernel (Hessenberg entries, norms) stay in device memory, so one Arnoldi step needs a single
// host synchronisation (to read the new Hessenberg column for the Givens update).
#include <cstdint>

#include <algorithm>

#include "device.cuh"
#include "peer_dev.cuh"

namespace an {

static constexpr int kDotThreads = 256;

// Deterministic two-stage dot product: every block writes one partial, the last block to finish adds the
// partials in index order (fixed summation order => identical bits on every rank and every run).
__global__ void __launch_bounds__(kDotThreads) dot_kernel(int n, const double* __restrict__ a, const double* __restrict__ b,
                                                          double* __restrict__ result, double* __restrict__ partial,
                                                          unsigned int* __restrict__ counter) {
    __shared__ double warp_sum[kDotThreads / 32];
    __shared__ bool last;
    double acc = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) acc = fma(a[i], b[i], acc);
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int w = 0; w < kDotThreads / 32; ++w) s += warp_sum[w];
        partial[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // final pass by one block, fixed order: thread t sums partials t, t+256, ... then a tree over the block
    double s = 0;
    for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) s += ((volatile double*)partial)[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0;
        for (int w = 0; w < kDotThreads / 32; ++w) tot += warp_sum[w];
        *result = tot;
        *counter = 0;
    }
}

// One modified Gram-Schmidt step in a single pass (gmres.cpp:115-119 does a dot and an axpy per basis vector):
//   w -= h[0] * v          (h: the projection computed by the previous dot, still in device memory)
//   result = w . u         (u = the next basis vector, or u == nullptr: w . w for the final norm)
// Same deterministic two-stage reduction as dot_kernel.
template <bool VEC>
__global__ void __launch_bounds__(kDotThreads) axpy_dot_kernel(int n, const double* __restrict__ h, const double* __restrict__ v,
                                                               double* __restrict__ w, const double* __restrict__ u,
                                                               double* __restrict__ result, double* __restrict__ partial,
                                                               unsigned int* __restrict__ counter) {
    __shared__ double warp_sum[kDotThreads / 32];
    __shared__ bool last;
    const double alpha = -*h;
    double acc = 0;
    if (VEC) {  // n even, 16-byte aligned vectors: two double2 per thread and trip in flight
        const int n2 = n >> 1, stride = gridDim.x * blockDim.x;
        const double2* __restrict__ v2 = reinterpret_cast<const double2*>(v);
        const double2* __restrict__ u2 = reinterpret_cast<const double2*>(u);
        double2* __restrict__ w2 = reinterpret_cast<double2*>(w);
        double acc1 = 0;
        int i = blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + stride < n2; i += 2 * stride) {
            const double2 va = v2[i], vb = v2[i + stride];
            double2 wa = w2[i], wb = w2[i + stride];
            wa.x = fma(alpha, va.x, wa.x); wa.y = fma(alpha, va.y, wa.y);
            wb.x = fma(alpha, vb.x, wb.x); wb.y = fma(alpha, vb.y, wb.y);
            const double2 ua = u ? u2[i] : wa, ub = u ? u2[i + stride] : wb;
            w2[i] = wa;
            w2[i + stride] = wb;
            acc = fma(wa.x, ua.x, acc); acc = fma(wa.y, ua.y, acc);
            acc1 = fma(wb.x, ub.x, acc1); acc1 = fma(wb.y, ub.y, acc1);
        }
        if (i < n2) {
            const double2 va = v2[i];
            double2 wa = w2[i];
            wa.x = fma(alpha, va.x, wa.x); wa.y = fma(alpha, va.y, wa.y);
            const double2 ua = u ? u2[i] : wa;
            w2[i] = wa;
            acc = fma(wa.x, ua.x, acc); acc = fma(wa.y, ua.y, acc);
        }
        acc += acc1;
    } else
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double wi = fma(alpha, v[i], w[i]);
        w[i] = wi;
        acc = fma(wi, u ? u[i] : wi, acc);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int k = 0; k < kDotThreads / 32; ++k) s += warp_sum[k];
        partial[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double s = 0;
    for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) s += ((volatile double*)partial)[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0;
        for (int k = 0; k < kDotThreads / 32; ++k) tot += warp_sum[k];
        *result = tot;
        *counter = 0;
    }
}

void launch_axpy_dot(int n, const double* h, const double* v, double* w, const double* u, double* result, DotWork& wk,
                     cudaStream_t st) {
    int blocks = (n + kDotThreads * 4 - 1) / (kDotThreads * 4);
    if (blocks < 1) blocks = 1;
    if (blocks > wk.blocks) blocks = wk.blocks;
    const bool vec = (n & 1) == 0 && ((reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(u)) & 15) == 0;
    if (vec) axpy_dot_kernel<true><<<blocks, kDotThreads, 0, st>>>(n, h, v, w, u, result, wk.partial, wk.counter);
    else axpy_dot_kernel<false><<<blocks, kDotThreads, 0, st>>>(n, h, v, w, u, result, wk.partial, wk.counter);
}

void launch_dot(int n, const double* a, const double* b, double* result, DotWork& wk, cudaStream_t st) {
    int blocks = (n + kDotThreads * 4 - 1) / (kDotThreads * 4);
    if (blocks < 1) blocks = 1;
    if (blocks > wk.blocks) blocks = wk.blocks;
    dot_kernel<<<blocks, kDotThreads, 0, st>>>(n, a, b, result, wk.partial, wk.counter);
}

__global__ void axpy_neg_dev_kernel(int n, const double* __restrict__ h, const double* __restrict__ v, double* __restrict__ w) {
    const double alpha = -*h;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) w[i] = fma(alpha, v[i], w[i]);
}
void launch_axpy_neg_dev(int n, const double* h, const double* v, double* w, cudaStream_t st) {
    if (n) axpy_neg_dev_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, h, v, w);
}

__global__ void axpy_kernel(int n, double alpha, const double* __restrict__ x, double* __restrict__ y) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = fma(alpha, x[i], y[i]);
}
void launch_axpy(int n, double alpha, const double* x, double* y, cudaStream_t st) {
    if (n) axpy_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, alpha, x, y);
}

__global__ void scale_inv_dev_kernel(int n, const double* __restrict__ norm, const double* __restrict__ in, double* __restrict__ out,
                                     double* __restrict__ out2, int squared) {
    const double s = 1.0 / (squared ? sqrt(*norm) : *norm);  // gmres.cpp:122 / 100: dscal(1.0 / h, v)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const double v = in[i] * s;
        out[i] = v;
        if (out2) out2[i] = v;
    }
}
void launch_scale_inv_dev(int n, const double* norm, const double* in, double* out, int squared, cudaStream_t st, double* out2) {
    if (n) scale_inv_dev_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, norm, in, out, out2, squared);
}

// block operator: the inputs of (kernel id, output mode) pairs are combinations of at most two scaled input modes
__global__ void __launch_bounds__(256) lincomb_kernel(const __grid_constant__ LinList L, size_t len) {
    const int e = blockIdx.y;
    const double* __restrict__ a = L.a[e];
    const double* __restrict__ b = L.b[e];
    double* __restrict__ d = L.dst[e];
    const double ca = L.ca[e], cb = L.cb[e];
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += (size_t)gridDim.x * blockDim.x)
        d[i] = b ? fma(cb, b[i], ca * a[i]) : ca * a[i];
}
void launch_lincomb(const LinList& L, size_t len, cudaStream_t st) {
    if (L.n == 0 || len == 0) return;
    const unsigned gx = (unsigned)std::min<size_t>((len + 255) / 256, 1024);
    lincomb_kernel<<<dim3(gx, L.n), 256, 0, st>>>(L, len);
}

__global__ void sqrt1_kernel(double* v) { *v = sqrt(*v); }
void launch_sqrt1(double* v, cudaStream_t st) { sqrt1_kernel<<<1, 1, 0, st>>>(v); }

__global__ void sub_kernel(int n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = a[i] - b[i];
}
void launch_sub(int n, const double* a, const double* b, double* out, cudaStream_t st) {
    if (n) sub_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, a, b, out);
}

// ---- scalar part of an Arnoldi step on the device (gmres.cpp:123-144) ------------------------------------------------
// The host version reads the new Hessenberg column back and synchronises once per step; with the column, the rotations
// and the recurrence residual kept on the device the host can queue step i+1 before it learns how step i ended.
// One thread; the arithmetic is written with non-contracting intrinsics in the order of the host code (capi.cu
// plane_rotation / apply_rotation), so both paths produce the same bits.
__device__ __forceinline__ void dev_apply_rotation(double& dx, double& dy, double cs, double sn) {  // ApplyPlaneRotation
    const double t = __dadd_rn(__dmul_rn(cs, dx), __dmul_rn(sn, dy));
    dy = __dadd_rn(__dmul_rn(-sn, dx), __dmul_rn(cs, dy));
    dx = t;
}
__global__ void givens_step_kernel(int i, const double* __restrict__ h1, const double* __restrict__ h2, const double* __restrict__ norm2,
                                   const double* __restrict__ beta, double* __restrict__ H, int ldh, double* __restrict__ cs,
                                   double* __restrict__ sn, double* __restrict__ s, double normb, double tol, double* __restrict__ report) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double* __restrict__ col = H + (size_t)i * ldh;
    for (int k = 0; k <= i; ++k) col[k] = h2 ? __dadd_rn(h1[k], h2[k]) : h1[k];  // CGS2: H(:, i) = h1 + h2
    double n2;
    if (norm2) n2 = *norm2;
    else {  // two-reduction CGS2: w . w = |w'|^2 - |h2|^2, formed exactly as the last pass forms it
        double q = 0;
        for (int k = 0; k <= i; ++k) q = fma(h2[k], h2[k], q);
        n2 = h2[i + 1] - q;
    }
    col[i + 1] = sqrt(n2);
    if (i == 0) s[0] = *beta;
    s[i + 1] = 0.0;
    for (int k = 0; k < i; ++k) dev_apply_rotation(col[k], col[k + 1], cs[k], sn[k]);
    const double dx = col[i], dy = col[i + 1];
    double c, sv;  // GeneratePlaneRotation, gmres.cpp:26-39
    if (dy == 0.0) {
        c = 1.0;
        sv = 0.0;
    } else if (fabs(dy) > fabs(dx)) {
        const double t = __ddiv_rn(dx, dy);
        sv = __ddiv_rn(1.0, __dsqrt_rn(__dadd_rn(1.0, __dmul_rn(t, t))));
        c = __dmul_rn(t, sv);
    } else {
        const double t = __ddiv_rn(dy, dx);
        c = __ddiv_rn(1.0, __dsqrt_rn(__dadd_rn(1.0, __dmul_rn(t, t))));
        sv = __dmul_rn(t, c);
    }
    cs[i] = c;
    sn[i] = sv;
    dev_apply_rotation(col[i], col[i + 1], c, sv);
    dev_apply_rotation(s[i], s[i + 1], c, sv);
    const double resid = __ddiv_rn(fabs(s[i + 1]), normb);
    report[0] = resid;
    report[1] = resid < tol ? 1.0 : 0.0;  // gmres.cpp:135
}
void launch_givens_step(int i, const double* h1, const double* h2, const double* norm2, const double* beta, double* H, int ldh,
                        double* cs, double* sn, double* s, double normb, double tol, double* report, cudaStream_t st) {
    givens_step_kernel<<<1, 32, 0, st>>>(i, h1, h2, norm2, beta, H, ldh, cs, sn, s, normb, tol, report);
}

// Update, gmres.cpp:12-24: y = H(0:k, 0:k)^-1 s(0:k) by back substitution (same operation order as the host form, the inner
// loop spread over one warp); writes -y so that x += V y is one launch_multi_axpy (which subtracts)
__global__ void gm_solve_kernel(int k, const double* __restrict__ H, int ldh, const double* __restrict__ s, double* __restrict__ yneg) {
    extern __shared__ double y[];
    const int lane = threadIdx.x;
    for (int q = lane; q <= k; q += 32) y[q] = s[q];
    __syncwarp();
    for (int i = k; i >= 0; --i) {
        if (lane == 0) y[i] = __ddiv_rn(y[i], H[i + (size_t)i * ldh]);
        __syncwarp();
        const double yi = y[i];
        for (int j = i - 1 - lane; j >= 0; j -= 32) y[j] = __dadd_rn(y[j], -__dmul_rn(H[j + (size_t)i * ldh], yi));
        __syncwarp();
    }
    for (int q = lane; q <= k; q += 32) yneg[q] = -y[q];
}
void launch_gm_solve(int k, const double* H, int ldh, const double* s, double* yneg, cudaStream_t st) {
    gm_solve_kernel<<<1, 32, (size_t)(k + 1) * sizeof(double), st>>>(k, H, ldh, s, yneg);
}

// ---- classical Gram-Schmidt with re-orthogonalisation (CGS2): all projections of one pass in two launches ----
// multi_dot: partial[(k * gridDim.x) + b] = sum over block b's share of w . V_k, k = blockIdx.y * 8 + kk
static constexpr int kMD = 8;
__global__ void __launch_bounds__(kDotThreads) multi_dot_kernel(int n, int nv, const double* __restrict__ w, const double* __restrict__ V,
                                                                size_t ldv, double* __restrict__ partial) {
    __shared__ double warp_sum[kMD][kDotThreads / 32];
    const int k0 = blockIdx.y * kMD;
    double acc[kMD];
#pragma unroll
    for (int kk = 0; kk < kMD; ++kk) acc[kk] = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double wi = w[i];
#pragma unroll
        for (int kk = 0; kk < kMD; ++kk)
            if (k0 + kk < nv) acc[kk] = fma(wi, V[(size_t)(k0 + kk) * ldv + i], acc[kk]);
    }
#pragma unroll
    for (int kk = 0; kk < kMD; ++kk) {
        double a = acc[kk];
        for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
        if ((threadIdx.x & 31) == 0) warp_sum[kk][threadIdx.x >> 5] = a;
    }
    __syncthreads();
    if (threadIdx.x < kMD && k0 + threadIdx.x < nv) {
        double s = 0;
        for (int q = 0; q < kDotThreads / 32; ++q) s += warp_sum[threadIdx.x][q];
        partial[(size_t)(k0 + threadIdx.x) * gridDim.x + blockIdx.x] = s;
    }
}
// out[k] = sum_b partial[k][b] in a fixed order (one warp per k); loads in independent batches of eight
__global__ void reduce_partials_kernel(int nv, int nb, const double* __restrict__ partial, double* __restrict__ out) {
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (k >= nv) return;
    const double* __restrict__ row = partial + (size_t)k * nb;
    double s = 0;
    for (int b = lane; b < nb; b += 32 * 8) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = b + 32 * u < nb ? row[b + 32 * u] : 0.0;
        s += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) out[k] = s;
}
int multi_dot_blocks(int n, const DotWork& wk) {
    int blocks = (n + kDotThreads * 4 - 1) / (kDotThreads * 4);
    return blocks < 1 ? 1 : blocks > wk.blocks ? wk.blocks : blocks;
}
void launch_multi_dot(int n, int nv, const double* w, const double* V, size_t ldv, double* out, double* partial, DotWork& wk,
                      cudaStream_t st) {
    const int nb = multi_dot_blocks(n, wk);
    multi_dot_kernel<<<dim3(nb, (nv + kMD - 1) / kMD), kDotThreads, 0, st>>>(n, nv, w, V, ldv, partial);
    reduce_partials_kernel<<<(nv + 7) / 8, 256, 0, st>>>(nv, nb, partial, out);
}
// w -= sum_k h[k] V_k; optionally partial[b] = block b's share of w . w (new w) for the norm
__global__ void __launch_bounds__(kDotThreads) multi_axpy_kernel(int n, int nv, const double* __restrict__ h, const double* __restrict__ V,
                                                                 size_t ldv, double* __restrict__ w, double* __restrict__ partial) {
    __shared__ double warp_sum[kDotThreads / 32];
    double acc = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double a0 = w[i], a1 = 0;
        int k = 0;
        for (; k + 2 <= nv; k += 2) {
            a0 = fma(-__ldg(h + k), V[(size_t)k * ldv + i], a0);
            a1 = fma(-__ldg(h + k + 1), V[(size_t)(k + 1) * ldv + i], a1);
        }
        if (k < nv) a0 = fma(-__ldg(h + k), V[(size_t)k * ldv + i], a0);
        const double wi = a0 + a1;
        w[i] = wi;
        acc = fma(wi, wi, acc);
    }
    if (!partial) return;
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int q = 0; q < kDotThreads / 32; ++q) s += warp_sum[q];
        partial[blockIdx.x] = s;
    }
}
void launch_multi_axpy(int n, int nv, const double* h, const double* V, size_t ldv, double* w, double* norm2_out, double* partial,
                       DotWork& wk, cudaStream_t st) {
    const int nb = multi_dot_blocks(n, wk);
    multi_axpy_kernel<<<nb, kDotThreads, 0, st>>>(n, nv, h, V, ldv, w, norm2_out ? partial : nullptr);
    if (norm2_out) reduce_partials_kernel<<<1, 32, 0, st>>>(1, nb, partial, norm2_out);
}


// ---- CGS2 in three fused passes ---------------------------------------------------------------------------------------
// The two-launch form above makes four passes over the Krylov basis per Arnoldi step (dots, axpys, dots, axpys), re-reads
// w once per group of eight vectors and was measured at half the HBM roofline (148 us per step at N = 2^20, ~13 vectors).
// Here one kernel serves all three steps of the twice-applied classical Gram-Schmidt:
//   MODE 0   h1 = V^T w
//   MODE 1   w -= V h1, then h2 = V^T w for the NEW w: the dot is a sum over elements, so a block can form it for the
//            elements it has just updated; the basis is read once from HBM and once more while it is still in L2
//   MODE 2   w -= V h2 and |w|^2
// i.e. three passes over the basis instead of four.  A thread keeps eight elements of w in registers (four 16-byte
// loads, coalesced across the warp) and walks the basis vectors with eight 16-byte loads in flight; per vector the warp
// sum goes to a per-warp slot in shared memory, block sums to `partial`, and the last block to finish adds them in a fixed
// order (deterministic, identical on every rank).
// CGS_U 16-byte elements of w per thread, CGS_KU basis vectors per trip of the walk (CGS_U * CGS_KU loads in flight per
// thread), CGS_CTAS resident CTAs per SM
#ifndef CGS_U
#define CGS_U 2
#define CGS_KU 4
#define CGS_CTAS 4
#endif
static constexpr int kCgsKU = CGS_KU;
template <int MODE>
__global__ void __launch_bounds__(256, CGS_CTAS) cgs_pass_kernel(int n, int nv, const double* __restrict__ h_in, const double* __restrict__ V,
                                                          size_t ldv, double* __restrict__ w, double* __restrict__ partial,
                                                          double* __restrict__ out, unsigned int* __restrict__ counter,
                                                          const __grid_constant__ PeerRed pr, double* __restrict__ vnext,
                                                          double* __restrict__ vnext2) {
    extern __shared__ double cg_sm[];  // hs[nv] | wpart[nout][8]
    double* __restrict__ hs = cg_sm;
    double* __restrict__ wpart = cg_sm + nv;
    // MODE 3: MODE 1 plus one more output, |w|^2 of the updated w (out[nv]); MODE 4: the update only, and
    // out[0] = h_in[nv] - sum_k h_in[k]^2 = |w - V h|^2 for orthonormal V and h = V^T w (h_in[nv] = |w|^2 from MODE 3)
    const int nout = MODE == 2 ? 1 : MODE == 3 ? nv + 1 : MODE == 4 ? 0 : nv;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (MODE)
        for (int k = threadIdx.x; k < nv; k += blockDim.x) hs[k] = h_in[k];
    for (int k = threadIdx.x; k < nout * 8; k += blockDim.x) wpart[k] = 0.0;
    __syncthreads();
    double nrm2 = 0, inv = 1.0;
    if (MODE == 4) {  // |w - V h|^2 = |w|^2 - |h|^2 (see above); every thread forms it alike
        double s = 0;
        for (int k = 0; k < nv; ++k) s = fma(hs[k], hs[k], s);
        nrm2 = h_in[nv] - s;
        inv = 1.0 / sqrt(nrm2);  // gmres.cpp:122: dscal(1.0 / h, v)
    }
    const int npair = n >> 1;
    const int per = (npair + gridDim.x - 1) / gridDim.x;
    const int p0 = blockIdx.x * per, p1 = min(npair, p0 + per);
    double2* __restrict__ w2 = reinterpret_cast<double2*>(w);
    for (int c = p0; c < p1; c += 256 * CGS_U) {
        int idx[CGS_U];
        bool ok[CGS_U];
        double2 a[CGS_U];
#pragma unroll
        for (int u = 0; u < CGS_U; ++u) {
            idx[u] = c + u * 256 + threadIdx.x;
            ok[u] = idx[u] < p1;
            a[u] = ok[u] ? w2[idx[u]] : make_double2(0.0, 0.0);
        }
        if (MODE) {
#pragma unroll kCgsKU
            for (int k = 0; k < nv; ++k) {
                const double2* __restrict__ Vk = reinterpret_cast<const double2*>(V + (size_t)k * ldv);
                const double hk = -hs[k];
                double2 v[CGS_U];
#pragma unroll
                for (int u = 0; u < CGS_U; ++u) v[u] = ok[u] ? Vk[idx[u]] : make_double2(0.0, 0.0);
#pragma unroll
                for (int u = 0; u < CGS_U; ++u) {
                    a[u].x = fma(hk, v[u].x, a[u].x);
                    a[u].y = fma(hk, v[u].y, a[u].y);
                }
            }
            if (MODE == 4 && vnext) {  // the next basis vector, already normalised; w itself is not needed again
#pragma unroll
                for (int u = 0; u < CGS_U; ++u)
                    if (ok[u]) {
                        const double2 v = make_double2(a[u].x * inv, a[u].y * inv);
                        reinterpret_cast<double2*>(vnext)[idx[u]] = v;
                        if (vnext2) reinterpret_cast<double2*>(vnext2)[idx[u]] = v;
                    }
            } else {
#pragma unroll
                for (int u = 0; u < CGS_U; ++u)
                    if (ok[u]) w2[idx[u]] = a[u];
            }
        }
        if (MODE == 2 || MODE == 3) {
            double s = 0;
#pragma unroll
            for (int u = 0; u < CGS_U; ++u) s = fma(a[u].x, a[u].x, fma(a[u].y, a[u].y, s));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) wpart[(MODE == 3 ? nv * 8 : 0) + warp] += s;
        }
        if (MODE == 0 || MODE == 1 || MODE == 3) {
#pragma unroll kCgsKU
            for (int k = 0; k < nv; ++k) {
                const double2* __restrict__ Vk = reinterpret_cast<const double2*>(V + (size_t)k * ldv);
                double2 v[CGS_U];
#pragma unroll
                for (int u = 0; u < CGS_U; ++u) v[u] = ok[u] ? Vk[idx[u]] : make_double2(0.0, 0.0);
                double s0 = 0, s1 = 0;
#pragma unroll
                for (int u = 0; u < CGS_U; ++u) {
                    s0 = fma(a[u].x, v[u].x, s0);
                    s1 = fma(a[u].y, v[u].y, s1);
                }
                double s = s0 + s1;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (lane == 0) wpart[k * 8 + warp] += s;
            }
        }
    }
    if (MODE == 4) {  // no reduction: the norm of the result follows from the values of the previous pass
        if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = nrm2;
        return;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < nout; k += blockDim.x) {
        double s = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) s += wpart[k * 8 + q];
        partial[(size_t)k * gridDim.x + blockIdx.x] = s;
    }
    // the last block to finish adds the block sums in index order (as dot_kernel does): no separate reduction launch
    __shared__ bool last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(counter, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    for (int k = warp; k < nout; k += 8) {
        const double* row = partial + (size_t)k * gridDim.x;
        double s = 0;
        for (int b = lane; b < (int)gridDim.x; b += 32 * 4) {  // four independent loads per trip, fixed order
            double v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = b + 32 * u < (int)gridDim.x ? __ldcg(row + b + 32 * u) : 0.0;
            s += (v[0] + v[1]) + (v[2] + v[3]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) out[k] = s;
    }
    if (threadIdx.x == 0) *counter = 0;
    if (pr.P) {  // sums over the ranks, inside the same launch
        __syncthreads();
        peer_allreduce_block(out, nout, pr);
    }
}

bool cgs_pass_ok(int n, int nv, const double* V, size_t ldv, const double* w) {
    return n >= 2 && (n & 1) == 0 && (ldv & 1) == 0 && nv <= 512 && ((uintptr_t)V & 15) == 0 && ((uintptr_t)w & 15) == 0;
}
void launch_cgs_pass(int mode, int n, int nv, const double* h_in, const double* V, size_t ldv, double* w, double* out, double* partial,
                     DotWork& wk, cudaStream_t st, const PeerRed* prp, double* vnext, double* vnext2) {
    static const PeerRed none = PeerRed();
    const PeerRed& pr = prp ? *prp : none;
    const int npair = n >> 1;
    int nb = std::min(std::min(148 * CGS_CTAS, wk.blocks), (npair + 256 * CGS_U - 1) / (256 * CGS_U));  // one balanced wave of resident CTAs
    if (nb < 1) nb = 1;
    const int nout = mode == 2 ? 1 : mode == 3 ? nv + 1 : mode == 4 ? 0 : nv;
    const size_t smem = (size_t)(nv + 8 * nout) * sizeof(double);
    if (mode == 0) cgs_pass_kernel<0><<<nb, 256, smem, st>>>(n, nv, h_in, V, ldv, w, partial, out, wk.counter, pr, vnext, vnext2);
    else if (mode == 1) cgs_pass_kernel<1><<<nb, 256, smem, st>>>(n, nv, h_in, V, ldv, w, partial, out, wk.counter, pr, vnext, vnext2);
    else if (mode == 2) cgs_pass_kernel<2><<<nb, 256, smem, st>>>(n, nv, h_in, V, ldv, w, partial, out, wk.counter, pr, vnext, vnext2);
    else if (mode == 3) cgs_pass_kernel<3><<<nb, 256, smem, st>>>(n, nv, h_in, V, ldv, w, partial, out, wk.counter, pr, vnext, vnext2);
    else cgs_pass_kernel<4><<<nb, 256, smem, st>>>(n, nv, h_in, V, ldv, w, partial, out, wk.counter, pr, vnext, vnext2);
}

}  // namespace an
