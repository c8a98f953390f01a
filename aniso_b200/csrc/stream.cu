// K3/K4 cached apply: y[row] = sum_e sum_s A[blk_e + s*nrow + t] x[col_e + s]  (downPassFast, bbfmm.h:1051-1065 and
// 1081-1099).  HBM-bound: the operator blocks (5.9 GB at N = 2^20) are read exactly once per matvec and never reused,
// so every load of A bypasses L1 allocation and the kernels are organised around keeping enough bytes in flight.
//
//   bsr_apply_warp_kernel<R>   all blocks R x R with 64 % R == 0 (np = 4 -> R 16, np = 8 -> R 64; also the P2P tiles of
//                              uniform 16- or 64-point leaves).  The blocks of one target are contiguous, so a warp
//                              walks them as one linear stream of 16-byte loads; each lane always lands on the same
//                              target pair (t, t+1), accumulates in registers, and a few shuffles finish the sum.
//   bsr_apply_row_kernel       any block shape: one thread per output row, 8 independent 8-byte loads in flight,
//                              next entry's metadata prefetched.
#include <cstdlib>

#include "device.cuh"

namespace an {

// Streaming loads: no L1 allocation, and an L2 evict-first policy so that the 5.9 GB operator stream does not flush the
// ~100 MB of vectors, tree tables and multipole/local coefficients that every other kernel of the matvec re-reads
// (the L2 holds 126 MB).
__device__ __forceinline__ uint64_t stream_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ double ld_stream(const double* p, uint64_t pol) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ double2 ld_stream2(const double2* p, uint64_t pol) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;"
                 : "=d"(v.x), "=d"(v.y)
                 : "l"(p), "l"(pol));
    return v;
}

template <int R>
__global__ void __launch_bounds__(256) bsr_apply_warp_kernel(DevCSR m, const double* __restrict__ A,
                                                             const double* __restrict__ x, double* __restrict__ y) {
    constexpr int PAIRS = R * R / 2;  // double2 per block
    constexpr int ITERS = PAIRS / 32;
    static_assert(64 % R == 0 && PAIRS % 32 == 0, "lane -> target pair must be loop invariant");
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= m.groups) return;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const uint64_t pol = stream_policy();
    const double2* __restrict__ base = reinterpret_cast<const double2*>(A + m.blk_off[e0]);
    // source index of this lane inside a block for iteration it: s = (it*64 + 2*lane) / R
    const int s_lane = (2 * lane) / R;
    constexpr int S_STEP = 64 / R;
    constexpr int CH = ITERS < 4 ? ITERS : 4;  // iterations per chunk: 2*CH 16-byte loads in flight per lane
    double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    for (int eb = 0; eb < ne; eb += 32) {
        const int cnt = min(32, ne - eb);
        const int mycol = lane < cnt ? m.col_off[e0 + eb + lane] : 0;
        int j = 0;
        for (; j + 2 <= cnt; j += 2) {  // two blocks per trip
            const int c0 = __shfl_sync(0xffffffffu, mycol, j) + s_lane, c1 = __shfl_sync(0xffffffffu, mycol, j + 1) + s_lane;
            const double2* __restrict__ b0 = base + (size_t)(eb + j) * PAIRS + lane;
            const double2* __restrict__ b1 = b0 + PAIRS;
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH) {
                double2 a0[CH], a1[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                    a0[u] = ld_stream2(b0 + (it0 + u) * 32, pol);
                    a1[u] = ld_stream2(b1 + (it0 + u) * 32, pol);
                }
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                    const double x0 = __ldg(x + c0 + (it0 + u) * S_STEP), x1 = __ldg(x + c1 + (it0 + u) * S_STEP);
                    acc0 = fma(a0[u].x, x0, acc0);
                    acc1 = fma(a0[u].y, x0, acc1);
                    acc2 = fma(a1[u].x, x1, acc2);
                    acc3 = fma(a1[u].y, x1, acc3);
                }
            }
        }
        if (j < cnt) {
            const int c0 = __shfl_sync(0xffffffffu, mycol, j) + s_lane;
            const double2* __restrict__ b0 = base + (size_t)(eb + j) * PAIRS + lane;
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH) {
                double2 a0[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) a0[u] = ld_stream2(b0 + (it0 + u) * 32, pol);
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                    const double x0 = __ldg(x + c0 + (it0 + u) * S_STEP);
                    acc0 = fma(a0[u].x, x0, acc0);
                    acc1 = fma(a0[u].y, x0, acc1);
                }
            }
        }
    }
    acc0 += acc2;
    acc1 += acc3;
#pragma unroll
    for (int off = 16; off >= R / 2; off >>= 1) {  // lanes l, l + R/2, l + R, ... hold the same target pair
        acc0 += __shfl_down_sync(0xffffffffu, acc0, off);
        acc1 += __shfl_down_sync(0xffffffffu, acc1, off);
    }
    if (lane < R / 2) {
        double* __restrict__ out = y + m.row_off[g] + 2 * lane;
        out[0] = acc0;
        out[1] = acc1;
    }
}

// Swizzled storage of 16 x 16 blocks (DevCSR::swz): element (s, t) sits at s*16 + (t ^ 4*(s%4)), so that the 8 x 4 and
// 4 x 8 fragment reads of the DMMA kernel below are free of shared-memory bank conflicts.  For the row-streaming kernels
// this only renames the target pair a lane works on: lane (q = lane/8, c = lane%8) holds pair c ^ 2q in every iteration.
template <int R>
__device__ __forceinline__ int sym_t0(int lane, int swz) {
    return R == 16 && swz ? 2 * ((lane & 7) ^ (2 * (lane >> 3))) : (2 * lane) % R;
}
// before the final reduction over the lanes of a target pair: bring pair c back to lane position c of each group of 8
template <int R>
__device__ __forceinline__ double sym_unswizzle(double v, int lane, int swz) {
    if (R != 16 || !swz) return v;
    return __shfl_sync(0xffffffffu, v, (lane & 24) | ((lane & 7) ^ (2 * (lane >> 3))));
}

// One block's worth of work for the symmetric kernel, iterations [it0, it0+4) already loaded into a[].
// Forward part: acc += K x_B.  Transposed part (paired blocks): the four partial products of this chunk are reduced
// over the LG lanes that share a source index with a butterfly that halves the live values at every step (2 + 1
// shuffles over lane bits 0 and 1, then one shuffle per remaining lane bit), instead of 4 x log2(LG) full reductions.
template <int R>
__device__ __forceinline__ void sym_consume_x(const double2 (&a)[4], const double (&xs)[4], int it0, double xa0, double xa1, int mr,
                                              int lane, int s_lane, double sign, double& acc0, double& acc1,
                                              double* __restrict__ scratch) {
    constexpr int S_STEP = 64 / R, LG = R / 2 < 32 ? R / 2 : 32;
    static_assert(LG >= 4, "butterfly needs at least four lanes per source index");
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        acc0 = fma(a[u].x, xs[u], acc0);
        acc1 = fma(a[u].y, xs[u], acc1);
    }
    if (mr < 0) return;  // warp-uniform
    const double p0 = fma(a[0].x, xa0, a[0].y * xa1), p1 = fma(a[1].x, xa0, a[1].y * xa1);
    const double p2 = fma(a[2].x, xa0, a[2].y * xa1), p3 = fma(a[3].x, xa0, a[3].y * xa1);
    const bool b0 = lane & 1, b1 = lane & 2;
    // lane bit 0: keep iterations {2,3} if set, else {0,1}
    double k0 = b0 ? p2 : p0, k1 = b0 ? p3 : p1;
    k0 += __shfl_xor_sync(0xffffffffu, b0 ? p0 : p2, 1);
    k1 += __shfl_xor_sync(0xffffffffu, b0 ? p1 : p3, 1);
    // lane bit 1: keep the second of the pair if set
    double z = b1 ? k1 : k0;
    z += __shfl_xor_sync(0xffffffffu, b1 ? k0 : k1, 2);
#pragma unroll
    for (int off = 4; off < LG; off <<= 1) z += __shfl_xor_sync(0xffffffffu, z, off);
    // lanes with (lane % LG) < 4 hold the finished z[s] for iteration it0 + u, u = 2*b0 + b1
    if ((lane % LG) < 4) {
        const int u = (b0 ? 2 : 0) + (b1 ? 1 : 0);
        scratch[(size_t)mr * R + (it0 + u) * S_STEP + s_lane] = sign * z;
    }
}

template <int R>
__device__ __forceinline__ void sym_consume(const double2 (&a)[4], int it0, const double* __restrict__ xs_base, double xa0,
                                            double xa1, int mr, int lane, int s_lane, double sign, double& acc0, double& acc1,
                                            double* __restrict__ scratch) {
    constexpr int S_STEP = 64 / R;
    double xs[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) xs[u] = __ldg(xs_base + (it0 + u) * S_STEP);
    sym_consume_x<R>(a, xs, it0, xa0, xa1, mr, lane, s_lane, sign, acc0, acc1, scratch);
}

// Symmetric storage: only one block of each (A,B)/(B,A) pair is in memory.  While a warp streams the blocks of its
// target A it also forms, for every paired block, z = sign * K^T x_A (the contribution of that block to the partner B)
// and writes the R values into B's scratch row; its own partial sum goes into A's first scratch row.  A second, tiny
// kernel adds up the rows of each target in a fixed order, so the result is deterministic (no atomics).  HBM traffic
// per matvec: half the blocks + 2 x 8R bytes of scratch per dropped block.
template <int R>
__global__ void __launch_bounds__(256, 4) bsr_apply_sym_kernel(DevCSR m, const double* __restrict__ A,
                                                            const double* __restrict__ x, double sign) {
    constexpr int PAIRS = R * R / 2, ITERS = PAIRS / 32;
    static_assert(64 % R == 0 && PAIRS % 32 == 0, "lane -> target pair must be loop invariant");
    constexpr int CH = 4;  // per block; two blocks per trip => 8 16-byte loads in flight per lane
    static_assert(ITERS % CH == 0, "chunking");
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= m.groups) return;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const uint64_t pol = stream_policy();
    const double2* __restrict__ base = reinterpret_cast<const double2*>(A + (ne ? m.blk_off[e0] : 0));
    const int s_lane = (2 * lane) / R, t0 = sym_t0<R>(lane, m.swz);
    const int self = m.row_off[g];
    const double xa0 = __ldg(x + self + t0), xa1 = __ldg(x + self + t0 + 1);
    double* __restrict__ scratch = m.scratch;
    double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    for (int eb = 0; eb < ne; eb += 32) {
        const int cnt = min(32, ne - eb);
        const int mycol = lane < cnt ? m.col_off[e0 + eb + lane] : 0;
        const int mymir = lane < cnt ? m.mirror_row[e0 + eb + lane] : -1;
        int j = 0;
        for (; j + 2 <= cnt; j += 2) {
            const int c0 = __shfl_sync(0xffffffffu, mycol, j) + s_lane, c1 = __shfl_sync(0xffffffffu, mycol, j + 1) + s_lane;
            const int m0 = __shfl_sync(0xffffffffu, mymir, j), m1 = __shfl_sync(0xffffffffu, mymir, j + 1);
            const double2* __restrict__ b0 = base + (size_t)(eb + j) * PAIRS + lane;
            const double2* __restrict__ b1 = b0 + PAIRS;
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH) {
                double2 a0[CH], a1[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                    a0[u] = ld_stream2(b0 + (it0 + u) * 32, pol);
                    a1[u] = ld_stream2(b1 + (it0 + u) * 32, pol);
                }
                sym_consume<R>(a0, it0, x + c0, xa0, xa1, m0, lane, s_lane, sign, acc0, acc1, scratch);
                sym_consume<R>(a1, it0, x + c1, xa0, xa1, m1, lane, s_lane, sign, acc2, acc3, scratch);
            }
        }
        if (j < cnt) {
            const int c0 = __shfl_sync(0xffffffffu, mycol, j) + s_lane;
            const int m0 = __shfl_sync(0xffffffffu, mymir, j);
            const double2* __restrict__ b0 = base + (size_t)(eb + j) * PAIRS + lane;
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH) {
                double2 a0[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) a0[u] = ld_stream2(b0 + (it0 + u) * 32, pol);
                sym_consume<R>(a0, it0, x + c0, xa0, xa1, m0, lane, s_lane, sign, acc0, acc1, scratch);
            }
        }
    }
    acc0 = sym_unswizzle<R>(acc0 + acc2, lane, m.swz);
    acc1 = sym_unswizzle<R>(acc1 + acc3, lane, m.swz);
#pragma unroll
    for (int off = 16; off >= R / 2; off >>= 1) {
        acc0 += __shfl_down_sync(0xffffffffu, acc0, off);
        acc1 += __shfl_down_sync(0xffffffffu, acc1, off);
    }
    if (lane < R / 2) {
        double* __restrict__ out = scratch + (size_t)m.srow_ptr[g] * R + 2 * lane;
        out[0] = acc0;
        out[1] = acc1;
    }
}

// Multi-mode block operator (aniso.m:121-156): the same operator block applied to NRHS right-hand sides in one pass over
// the operator (x_r = x + r * xstride, scratch set r = scratch + r * sstride).  Same lane map and butterfly as
// bsr_apply_sym_kernel; per 16-byte load the FP64 work grows NRHS-fold, which moves the kernel from the HBM roofline
// towards the FP64 pipe (measured in profiles/).
template <int R, int NRHS>
__global__ void __launch_bounds__(256, NRHS <= 2 ? 3 : 2) bsr_apply_sym_multi_kernel(DevCSR m, const double* __restrict__ A,
                                                                                   const double* __restrict__ x, size_t xstride,
                                                                                   double* __restrict__ scr, size_t sstride,
                                                                                   double sign) {
    constexpr int PAIRS = R * R / 2, ITERS = PAIRS / 32, CH = 4;
    static_assert(64 % R == 0 && PAIRS % 32 == 0 && ITERS % CH == 0, "lane -> target pair must be loop invariant");
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= m.groups) return;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const uint64_t pol = stream_policy();
    const double2* __restrict__ base = reinterpret_cast<const double2*>(A + (ne ? m.blk_off[e0] : 0));
    const int s_lane = (2 * lane) / R, t0 = sym_t0<R>(lane, m.swz);
    const int self = m.row_off[g];
    double xa0[NRHS], xa1[NRHS], acc0[NRHS], acc1[NRHS];
#pragma unroll
    for (int r = 0; r < NRHS; ++r) {
        xa0[r] = __ldg(x + r * xstride + self + t0);
        xa1[r] = __ldg(x + r * xstride + self + t0 + 1);
        acc0[r] = acc1[r] = 0;
    }
    for (int eb = 0; eb < ne; eb += 32) {
        const int cnt = min(32, ne - eb);
        const int mycol = lane < cnt ? m.col_off[e0 + eb + lane] : 0;
        const int mymir = lane < cnt ? m.mirror_row[e0 + eb + lane] : -1;
        int j = 0;
        for (; j + 2 <= cnt; j += 2) {  // two blocks per trip: eight 16-byte loads in flight per lane, as in the single-RHS kernel
            const int c0 = __shfl_sync(0xffffffffu, mycol, j) + s_lane, c1 = __shfl_sync(0xffffffffu, mycol, j + 1) + s_lane;
            const int m0 = __shfl_sync(0xffffffffu, mymir, j), m1 = __shfl_sync(0xffffffffu, mymir, j + 1);
            const double2* __restrict__ b0 = base + (size_t)(eb + j) * PAIRS + lane;
            const double2* __restrict__ b1 = b0 + PAIRS;
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH) {
                double2 a0[CH], a1[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                    a0[u] = ld_stream2(b0 + (it0 + u) * 32, pol);
                    a1[u] = ld_stream2(b1 + (it0 + u) * 32, pol);
                }
#pragma unroll
                for (int r = 0; r < NRHS; ++r)
                    sym_consume<R>(a0, it0, x + r * xstride + c0, xa0[r], xa1[r], m0, lane, s_lane, sign, acc0[r], acc1[r],
                                   scr + r * sstride);
#pragma unroll
                for (int r = 0; r < NRHS; ++r)
                    sym_consume<R>(a1, it0, x + r * xstride + c1, xa0[r], xa1[r], m1, lane, s_lane, sign, acc0[r], acc1[r],
                                   scr + r * sstride);
            }
        }
        if (j < cnt) {
            const int c0 = __shfl_sync(0xffffffffu, mycol, j) + s_lane;
            const int m0 = __shfl_sync(0xffffffffu, mymir, j);
            const double2* __restrict__ b0 = base + (size_t)(eb + j) * PAIRS + lane;
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH) {
                double2 a0[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) a0[u] = ld_stream2(b0 + (it0 + u) * 32, pol);
#pragma unroll
                for (int r = 0; r < NRHS; ++r)
                    sym_consume<R>(a0, it0, x + r * xstride + c0, xa0[r], xa1[r], m0, lane, s_lane, sign, acc0[r], acc1[r],
                                   scr + r * sstride);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < NRHS; ++r) {
        double v0 = sym_unswizzle<R>(acc0[r], lane, m.swz), v1 = sym_unswizzle<R>(acc1[r], lane, m.swz);
#pragma unroll
        for (int off = 16; off >= R / 2; off >>= 1) {
            v0 += __shfl_down_sync(0xffffffffu, v0, off);
            v1 += __shfl_down_sync(0xffffffffu, v1, off);
        }
        if (lane < R / 2) {
            double* __restrict__ out = scr + r * sstride + (size_t)m.srow_ptr[g] * R + 2 * lane;
            out[0] = v0;
            out[1] = v1;
        }
    }
}

// ---- the same symmetric stream fed by the TMA engine ---------------------------------------------------------------
// The blocks of one target are one contiguous run of ne * R * R doubles, so a warp can have the copy engine fetch it in
// 2 KB pieces (cp.async.bulk global -> shared, completion on an mbarrier) into a private ring of STAGES buffers and only
// ever holds the four 16-byte values of the piece it is consuming in registers.  Bytes in flight per SM no longer depend
// on the register budget (warps x STAGES x 2 KB), which is what the multi-RHS form needs: its FP64 / shuffle work per
// byte is NRHS times that of the single-RHS kernel and it was latency-bound with register-held loads.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "l"(pol)
        : "memory");
}

__device__ __forceinline__ void bulk_g2s_keep(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}

constexpr int RING_PIECE = 2048;  // bytes of operator per stage: four iterations of 32 lanes x 16 bytes
constexpr int RING_COLS = 64;     // column offsets of a target's first blocks, kept in shared memory for the producer lane
template <int R, int NRHS>
__host__ __device__ constexpr int ring_stage_bytes() {  // operator piece + the 4 * (64 / R) source coefficients it multiplies, per right-hand side
    return RING_PIECE + NRHS * 4 * (64 / R) * 8;
}
template <int R, int NRHS, int STAGES>
constexpr int ring_smem_bytes() {
    return 8 * (STAGES * ring_stage_bytes<R, NRHS>() + STAGES * 8 + RING_COLS * 4);
}

// One stage of the ring is filled by 1 + NRHS bulk copies that complete on the same mbarrier: 2 KB of operator and, for
// every right-hand side, the slice of x_B that those 2 KB multiply.  The source coefficients travel with the operator
// because a separate gather (L1 hit rate ~30 %) queues behind the ~190 KB of operator that each SM keeps in flight and
// then costs more than the whole piece does.
template <int R, int NRHS, int STAGES>
__global__ void __launch_bounds__(256, 3) bsr_apply_sym_ring_kernel(DevCSR m, const double* __restrict__ A, const double* __restrict__ x,
                                                                    size_t xstride, double* __restrict__ scr, size_t sstride,
                                                                    double sign) {
    constexpr int PAIRS = R * R / 2, ITERS = PAIRS / 32, CH = 4, PPB = ITERS / CH;  // pieces per block
    constexpr int S_STEP = 64 / R, XB = CH * S_STEP * 8, STAGE = ring_stage_bytes<R, NRHS>();
    constexpr bool LAZY = NRHS >= 3;
    static_assert(64 % R == 0 && PAIRS % 32 == 0 && ITERS % CH == 0, "lane -> target pair must be loop invariant");
    static_assert(XB % 16 == 0 && STAGE % 16 == 0, "bulk copies move multiples of 16 bytes");
    extern __shared__ __align__(128) unsigned char ring_smem[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + wid;
    if (g >= m.groups) return;  // warps are independent: no CTA-wide barrier below
    unsigned char* ring = ring_smem + (size_t)wid * STAGES * STAGE;
    const uint32_t ring_a = smem_u32(ring);
    const uint32_t bar_a = smem_u32(ring_smem + 8 * STAGES * STAGE + wid * STAGES * 8);
    int* cols = reinterpret_cast<int*>(ring_smem + 8 * STAGES * STAGE + 8 * STAGES * 8) + wid * RING_COLS;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const uint64_t pol = stream_policy();
    const unsigned char* __restrict__ src = reinterpret_cast<const unsigned char*>(A + (ne ? m.blk_off[e0] : 0));
    const int npieces = ne * PPB;
#pragma unroll
    for (int k = 0; k < RING_COLS; k += 32)
        if (k + lane < ne) cols[k + lane] = m.col_off[e0 + k + lane];
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar_a + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    // producer (lane 0): piece q of this target -> stage q % STAGES
    auto fill = [&](int q, int stage) {
        const int e = q / PPB, it = (q - e * PPB) * CH;
        const int col = e < RING_COLS ? cols[e] : m.col_off[e0 + e];
        const uint32_t dst = ring_a + stage * STAGE, bar = bar_a + 8 * stage;
        mbar_expect_tx(bar, STAGE);
        bulk_g2s(dst, src + (size_t)q * RING_PIECE, RING_PIECE, bar, pol);
#pragma unroll
        for (int r = 0; r < NRHS; ++r) bulk_g2s_keep(dst + RING_PIECE + r * XB, x + r * xstride + col + it * S_STEP, XB, bar);
    };
    if (lane == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < STAGES; ++s)
            if (s < npieces) fill(s, s);
    }
    const int s_lane = (2 * lane) / R, t0 = sym_t0<R>(lane, m.swz);
    const int self = m.row_off[g];
    double xa0[NRHS], xa1[NRHS], acc0[NRHS], acc1[NRHS];
#pragma unroll
    for (int r = 0; r < NRHS; ++r) {
        xa0[r] = __ldg(x + r * xstride + self + t0);
        xa1[r] = __ldg(x + r * xstride + self + t0 + 1);
        acc0[r] = acc1[r] = 0;
    }
    int piece = 0, stage = 0;
    uint32_t parity = 0;
    for (int eb = 0; eb < ne; eb += 32) {
        const int cnt = min(32, ne - eb);
        const int mymir = lane < cnt ? m.mirror_row[e0 + eb + lane] : -1;
        for (int j = 0; j < cnt; ++j) {
            const int m0 = __shfl_sync(0xffffffffu, mymir, j);
#pragma unroll 1
            for (int it0 = 0; it0 < ITERS; it0 += CH, ++piece) {
                mbar_wait(bar_a + 8 * stage, parity);
                const unsigned char* sp = ring + stage * STAGE;
                double2 a0[CH];
#pragma unroll
                for (int u = 0; u < CH; ++u) a0[u] = reinterpret_cast<const double2*>(sp)[u * 32 + lane];
                auto refill = [&]() {
                    __syncwarp();  // every lane has read what it needs from this stage
                    if (lane == 0 && piece + STAGES < npieces) {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        fill(piece + STAGES, stage);
                    }
                };
                if (LAZY) {  // many right-hand sides: x_B is read from the stage as it is used, the stage is released afterwards
#pragma unroll
                    for (int r = 0; r < NRHS; ++r) {
                        double xs[CH];
#pragma unroll
                        for (int u = 0; u < CH; ++u) xs[u] = reinterpret_cast<const double*>(sp + RING_PIECE + r * XB)[u * S_STEP + s_lane];
                        sym_consume_x<R>(a0, xs, it0, xa0[r], xa1[r], m0, lane, s_lane, sign, acc0[r], acc1[r], scr + r * sstride);
                    }
                    refill();
                } else {
                    double xs[NRHS][CH];
#pragma unroll
                    for (int r = 0; r < NRHS; ++r)
#pragma unroll
                        for (int u = 0; u < CH; ++u)
                            xs[r][u] = reinterpret_cast<const double*>(sp + RING_PIECE + r * XB)[u * S_STEP + s_lane];
                    refill();
#pragma unroll
                    for (int r = 0; r < NRHS; ++r)
                        sym_consume_x<R>(a0, xs[r], it0, xa0[r], xa1[r], m0, lane, s_lane, sign, acc0[r], acc1[r], scr + r * sstride);
                }
                if (++stage == STAGES) {
                    stage = 0;
                    parity ^= 1;
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < NRHS; ++r) {
        double v0 = sym_unswizzle<R>(acc0[r], lane, m.swz), v1 = sym_unswizzle<R>(acc1[r], lane, m.swz);
#pragma unroll
        for (int off = 16; off >= R / 2; off >>= 1) {
            v0 += __shfl_down_sync(0xffffffffu, v0, off);
            v1 += __shfl_down_sync(0xffffffffu, v1, off);
        }
        if (lane < R / 2) {
            double* __restrict__ out = scr + r * sstride + (size_t)m.srow_ptr[g] * R + 2 * lane;
            out[0] = v0;
            out[1] = v1;
        }
    }
}

template <int R, int NRHS>
static void launch_ring(const DevCSR& m, const double* A, const double* x, size_t xstride, double* scr, size_t sstride, double sign,
                        cudaStream_t st) {
    constexpr int STAGES = 3;  // 3 CTAs x (8 warps x 3 stages x ~2.3 KB) per SM
    static PerDeviceOnce once;  // per instantiation and device
    if (once.first()) {
        cudaFuncSetAttribute(bsr_apply_sym_ring_kernel<R, NRHS, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             ring_smem_bytes<R, NRHS, STAGES>());
    }
    bsr_apply_sym_ring_kernel<R, NRHS, STAGES>
        <<<(m.groups + 7) / 8, 256, ring_smem_bytes<R, NRHS, STAGES>(), st>>>(m, A, x, xstride, scr, sstride, sign);
}

// ---- multi-RHS stream on the FP64 tensor cores (16 x 16 blocks, swizzled storage) -----------------------------------
// With up to eight right-hand sides the two products of a block are small GEMMs,
//   forward     Y[rhs][t] += sum_s X_B[rhs][s] * A[s][t]         (8 x 16) = (8 x 16)(16 x 16)
//   transposed  Z[s][rhs]  = sum_t A[s][t]     * X_A[rhs][t]     (16 x 8) = (16 x 16)(16 x 8)
// i.e. 8 + 8 mma.m8n8k4.f64 per 2 KB block whatever the number of right-hand sides, instead of ~55 instructions per block
// AND right-hand side in the DFMA / butterfly form above.  The block is read from the ring stage twice, once as the
// B operand (4 x 8 fragments: lane l holds A[s = 4 sb + l%4][t = 8 tb + l/4]) and once as the A operand (8 x 4 fragments:
// lane l holds A[s = 8 sg + l/4][t = 4 tq + l%4]); the storage swizzle t ^ 4 (s%4) makes both reads conflict-free.
// Results need no cross-lane reduction: the accumulator fragments are written straight to the scratch rows.
__device__ __forceinline__ void dmma884(double2& c, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(c.x), "+d"(c.y) : "d"(a), "d"(b));
}

constexpr int MMA_STAGE = RING_PIECE + 8 * 128;  // a block and the x_B slices of up to eight right-hand sides
#define ANISO_PIN(v) asm volatile("" : "+r"(v))  // keep a lane-derived value in its register instead of recomputing it per block
#define ANISO_PIN64(v) asm volatile("" : "+l"(v))

template <int STAGES>
__global__ void __launch_bounds__(256, 3) bsr_apply_sym_mma_kernel(DevCSR m, const double* __restrict__ A, const double* __restrict__ x,
                                                                   size_t xstride, int nrhs, double* __restrict__ scr, size_t sstride,
                                                                   double sign) {
    constexpr int R = 16, STAGE = MMA_STAGE;
    extern __shared__ __align__(128) unsigned char ring_smem[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + wid;
    if (g >= m.groups) return;  // warps are independent: no CTA-wide barrier below
    unsigned char* ring = ring_smem + (size_t)wid * STAGES * STAGE;
    uint32_t ring_a = smem_u32(ring);
    uint32_t bar_a = smem_u32(ring_smem + 8 * STAGES * STAGE + wid * STAGES * 8);
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    // x_B area of a stage: [s-block sb][rhs 0..7][4 doubles], so that lane l reads doubles sb*32 + l (no bank conflicts,
    // one base register).  Rows of right-hand sides >= nrhs are never written by a copy and stay zero.
#pragma unroll
    for (int s = 0; s < STAGES; ++s)
#pragma unroll
        for (int k = 0; k < 4; ++k) reinterpret_cast<double*>(ring + s * STAGE + RING_PIECE)[k * 32 + lane] = 0.0;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar_a + 8 * s, 33);  // lane 0's expect_tx + one cp.async arrival per lane
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    // Producers: lane 0 has the copy engine fetch the 2 KB block (evict-first: it is never read again); the x_B slices the
    // block multiplies (128 bytes per right-hand side) are fetched 16 bytes per lane with cp.async, whose completion arrives
    // on the same mbarrier.  (One bulk copy per right-hand side costs ~15 serialised instructions each: UBLKCP takes its
    // operands from uniform registers.)  A stage is refilled only after the DMMAs that consumed it were issued, i.e. after
    // its shared-memory reads have returned, so no proxy fence is needed between the reads and the refill.
    const unsigned char* __restrict__ src = reinterpret_cast<const unsigned char*>(A + (ne ? m.blk_off[e0] : 0));
    const uint64_t pol = stream_policy();
    // chunk c = lane (+ 32): rhs c / 8, doubles 2 (c % 8) .. + 1 of the slice, i.e. s-block (c % 8) / 2, half (c % 8) % 2
    const double* __restrict__ xc0 = x + (size_t)(lane >> 3) * xstride + 2 * (lane & 7);
    const double* __restrict__ xc1 = xc0 + 4 * xstride;
    uint32_t xdst = RING_PIECE + ((lane & 7) >> 1) * 256 + (lane >> 3) * 32 + (lane & 1) * 16;  // second chunk: + 4 * 32
    int c_ok = (lane < nrhs * 8 ? 1 : 0) | (lane + 32 < nrhs * 8 ? 2 : 0);
    ANISO_PIN64(xc0); ANISO_PIN64(xc1); ANISO_PIN(xdst); ANISO_PIN(c_ok); ANISO_PIN(ring_a); ANISO_PIN(bar_a);
    auto fill = [&](int q, int stage, int col) {  // block q of this target -> stage
        const uint32_t dst = ring_a + stage * STAGE, bar = bar_a + 8 * stage;
        if (lane == 0) {
            mbar_expect_tx(bar, RING_PIECE);
            bulk_g2s(dst, src + (size_t)q * RING_PIECE, RING_PIECE, bar, pol);
        }
        if (c_ok & 1) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + xdst), "l"(xc0 + col) : "memory");
        if (c_ok & 2) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + xdst + 128), "l"(xc1 + col) : "memory");
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
    };
    // column offsets / mirror rows of 32 blocks per lane-batch; the batch after the current one is loaded a batch early
    int mycol = lane < ne ? m.col_off[e0 + lane] : 0;
    int mycol_n = 32 + lane < ne ? m.col_off[e0 + 32 + lane] : 0;
    int mymir = lane < ne ? m.mirror_row[e0 + lane] : -1;
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
        const int col = __shfl_sync(0xffffffffu, mycol, s);
        if (s < ne) fill(s, s, col);
    }
    const int lr = lane >> 2, lc = lane & 3;  // fragment row / column of this lane
    const int self = m.row_off[g];
    double xa[4];  // B operand of the transposed product: sign * X_A[rhs = lr][t = 4 tq + lc]
#pragma unroll
    for (int tq = 0; tq < 4; ++tq) xa[tq] = lr < nrhs ? sign * __ldg(x + lr * xstride + self + 4 * tq + lc) : 0.0;
    double2 accf[2] = {make_double2(0, 0), make_double2(0, 0)};  // Y[rhs = lr][t = 8 tb + 2 lc + {0, 1}]
    // byte offsets of this lane's fragment elements inside a stage (storage swizzle: t ^ 4 (s % 4))
    uint32_t offb[2], offa[4];
#pragma unroll
    for (int tb = 0; tb < 2; ++tb) {  // + 512 sb: s = 4 sb + lc, t = 8 tb + lr
        offb[tb] = 8 * (lc * 16 + ((8 * tb + lr) ^ (4 * lc)));
        ANISO_PIN(offb[tb]);
    }
#pragma unroll
    for (int tq = 0; tq < 4; ++tq) {  // + 1024 sg: s = 8 sg + lr, t = 4 tq + lc
        offa[tq] = 8 * (lr * 16 + ((4 * tq + lc) ^ (4 * (lr & 3))));
        ANISO_PIN(offa[tq]);
    }
    uint32_t offx = RING_PIECE + 8 * lane;
    ANISO_PIN(offx);
    // transposed-product accumulator fragment: Z[s = 8 sg + lr][rhs = 2 lc + {0, 1}]
    double* z0 = scr + (size_t)(2 * lc) * sstride + lr;
    double* z1 = z0 + sstride;
    int z_ok = (2 * lc < nrhs ? 1 : 0) | (2 * lc + 1 < nrhs ? 2 : 0);
    ANISO_PIN64(z0); ANISO_PIN64(z1); ANISO_PIN(z_ok);
    int stage = 0;
    uint32_t parity = 0;
    for (int eb = 0; eb < ne; eb += 32) {
        const int cnt = min(32, ne - eb);
        if (eb) {
            mycol = mycol_n;
            mycol_n = eb + 32 + lane < ne ? m.col_off[e0 + eb + 32 + lane] : 0;
            mymir = lane < cnt ? m.mirror_row[e0 + eb + lane] : -1;
        }
        for (int j = 0; j < cnt; ++j) {
            const int q = eb + j;
            const int m0 = __shfl_sync(0xffffffffu, mymir, j);
            const int jn = j + STAGES;
            const int c_a = __shfl_sync(0xffffffffu, mycol, jn & 31), c_b = __shfl_sync(0xffffffffu, mycol_n, jn & 31);
            const int col = jn < 32 ? c_a : c_b;
            mbar_wait(bar_a + 8 * stage, parity);
            const unsigned char* __restrict__ sp = ring + stage * STAGE;
            double xb[4], fb[4][2];
#pragma unroll
            for (int sb = 0; sb < 4; ++sb) {
                xb[sb] = *reinterpret_cast<const double*>(sp + offx + 256 * sb);
#pragma unroll
                for (int tb = 0; tb < 2; ++tb) fb[sb][tb] = *reinterpret_cast<const double*>(sp + offb[tb] + 512 * sb);
            }
            if (m0 >= 0) {  // warp-uniform; four independent accumulator chains
                double za[4][2];
#pragma unroll
                for (int tq = 0; tq < 4; ++tq)
#pragma unroll
                    for (int sg = 0; sg < 2; ++sg) za[tq][sg] = *reinterpret_cast<const double*>(sp + offa[tq] + 1024 * sg);
                double2 z[2] = {make_double2(0, 0), make_double2(0, 0)};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    dmma884(accf[0], xb[k], fb[k][0]);
                    dmma884(accf[1], xb[k], fb[k][1]);
                    dmma884(z[0], za[k][0], xa[k]);
                    dmma884(z[1], za[k][1], xa[k]);
                }
                const size_t zo = (size_t)m0 * R;
                if (z_ok & 1) {
                    z0[zo] = z[0].x;
                    z0[zo + 8] = z[1].x;
                }
                if (z_ok & 2) {
                    z1[zo] = z[0].y;
                    z1[zo + 8] = z[1].y;
                }
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    dmma884(accf[0], xb[k], fb[k][0]);
                    dmma884(accf[1], xb[k], fb[k][1]);
                }
            }
            __syncwarp();  // every lane has read what it needs from this stage
            if (q + STAGES < ne) fill(q + STAGES, stage, col);
            if (++stage == STAGES) {
                stage = 0;
                parity ^= 1;
            }
        }
    }
    if (lr < nrhs) {
        double* __restrict__ out = scr + lr * sstride + (size_t)m.srow_ptr[g] * R + 2 * lc;
#pragma unroll
        for (int tb = 0; tb < 2; ++tb) *reinterpret_cast<double2*>(out + 8 * tb) = accf[tb];
    }
}

static void launch_mma(const DevCSR& m, const double* A, const double* x, size_t xstride, int nrhs, double* scr, size_t sstride,
                       double sign, cudaStream_t st) {
    constexpr int STAGES = 3, SMEM = 8 * (STAGES * MMA_STAGE + STAGES * 8);
    static PerDeviceOnce once;  // per instantiation and device
    if (once.first()) {
        cudaFuncSetAttribute(bsr_apply_sym_mma_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    }
    bsr_apply_sym_mma_kernel<STAGES><<<(m.groups + 7) / 8, 256, SMEM, st>>>(m, A, x, xstride, nrhs, scr, sstride, sign);
}

// ---- the same, for all kernel ids of a block apply in one launch ------------------------------------------------------
// out_i = sum_k K_k w^(i,k) (aniso.m:121-156 regrouped by kernel id): the warp of a target walks the operators of all ids
// one after the other (k-major), the forward accumulators live across the whole walk, and the transposed products of id k
// are added onto those of the ids before it in the partner's scratch row: plain stores for the first id, red.add for the
// others.  Each scratch element has exactly one writing thread, which issues its additions in id order, so the sums are
// deterministic.  One scratch set per output mode instead of one per (id, input mode): K(2K-1) -> K sets written and read.
template <int STAGES>
__global__ void __launch_bounds__(256, 3) bsr_apply_sym_mma_ids_kernel(DevCSR m, const __grid_constant__ IdList ids, size_t xstride,
                                                                       int nrhs, double* __restrict__ scr, size_t sstride) {
    constexpr int R = 16, STAGE = MMA_STAGE, NCOL = 32;
    extern __shared__ __align__(128) unsigned char ring_smem[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + wid;
    if (g >= m.groups) return;  // warps are independent: no CTA-wide barrier below
    unsigned char* ring = ring_smem + (size_t)wid * STAGES * STAGE;
    uint32_t ring_a = smem_u32(ring);
    uint32_t bar_a = smem_u32(ring_smem + 8 * STAGES * STAGE + wid * STAGES * 8);
    int* cols = reinterpret_cast<int*>(ring_smem + 8 * STAGES * STAGE + 8 * STAGES * 8) + wid * NCOL;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const int nid = ids.n, npieces = nid * ne;
#pragma unroll
    for (int s = 0; s < STAGES; ++s)
#pragma unroll
        for (int k = 0; k < 4; ++k) reinterpret_cast<double*>(ring + s * STAGE + RING_PIECE)[k * 32 + lane] = 0.0;
    if (lane < ne) cols[lane] = m.col_off[e0 + lane];
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar_a + 8 * s, 33);  // lane 0's expect_tx + one cp.async arrival per lane
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    const size_t blk0 = ne ? (size_t)m.blk_off[e0] : 0;
    const uint64_t pol = stream_policy();
    size_t xo0 = (size_t)(lane >> 3) * xstride + 2 * (lane & 7);  // chunk `lane` of the x_B slices (see the kernel above)
    uint32_t xdst = RING_PIECE + ((lane & 7) >> 1) * 256 + (lane >> 3) * 32 + (lane & 1) * 16;
    int c_ok = (lane < nrhs * 8 ? 1 : 0) | (lane + 32 < nrhs * 8 ? 2 : 0);
    ANISO_PIN64(xo0); ANISO_PIN(xdst); ANISO_PIN(c_ok); ANISO_PIN(ring_a); ANISO_PIN(bar_a);
    int fk = 0, fq = 0, fstage = 0;  // fill cursor: next piece to request = block fq of id fk, into stage fstage
    auto fill = [&]() {
        const int col = fq < NCOL ? cols[fq] : m.col_off[e0 + fq];
        const uint32_t dst = ring_a + fstage * STAGE, bar = bar_a + 8 * fstage;
        if (lane == 0) {
            mbar_expect_tx(bar, RING_PIECE);
            bulk_g2s(dst, reinterpret_cast<const unsigned char*>(ids.A[fk] + blk0) + (size_t)fq * RING_PIECE, RING_PIECE, bar, pol);
        }
        const double* __restrict__ xs = ids.x[fk] + xo0 + col;
        if (c_ok & 1) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + xdst), "l"(xs) : "memory");
        if (c_ok & 2) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + xdst + 128), "l"(xs + 4 * xstride) : "memory");
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
        if (++fq == ne) {
            fq = 0;
            ++fk;
        }
        if (++fstage == STAGES) fstage = 0;
    };
    for (int s = 0; s < STAGES && s < npieces; ++s) fill();
    const int lr = lane >> 2, lc = lane & 3;  // fragment row / column of this lane
    const int self = m.row_off[g];
    double2 accf[2] = {make_double2(0, 0), make_double2(0, 0)};  // Y[rhs = lr][t = 8 tb + 2 lc + {0, 1}]
    uint32_t offb[2], offa[4];
#pragma unroll
    for (int tb = 0; tb < 2; ++tb) {  // + 512 sb: s = 4 sb + lc, t = 8 tb + lr
        offb[tb] = 8 * (lc * 16 + ((8 * tb + lr) ^ (4 * lc)));
        ANISO_PIN(offb[tb]);
    }
#pragma unroll
    for (int tq = 0; tq < 4; ++tq) {  // + 1024 sg: s = 8 sg + lr, t = 4 tq + lc
        offa[tq] = 8 * (lr * 16 + ((4 * tq + lc) ^ (4 * (lr & 3))));
        ANISO_PIN(offa[tq]);
    }
    uint32_t offx = RING_PIECE + 8 * lane;
    ANISO_PIN(offx);
    double* z0 = scr + (size_t)(2 * lc) * sstride + lr;  // transposed-product fragment: Z[s = 8 sg + lr][rhs = 2 lc + {0, 1}]
    double* z1 = z0 + sstride;
    int z_ok = (2 * lc < nrhs ? 1 : 0) | (2 * lc + 1 < nrhs ? 2 : 0);
    ANISO_PIN64(z0); ANISO_PIN64(z1); ANISO_PIN(z_ok);
    const size_t xa_off = (size_t)lr * xstride + self + lc;
    double xa[4];  // B operand of the transposed product of the current id: sign_k * X_A[rhs = lr][t = 4 tq + lc]
#pragma unroll
    for (int tq = 0; tq < 4; ++tq) xa[tq] = lr < nrhs && nid ? ids.sign[0] * __ldg(ids.x[0] + xa_off + 4 * tq) : 0.0;
    int mymir = lane < ne ? m.mirror_row[e0 + lane] : -1;
    int stage = 0, piece = 0;
    uint32_t parity = 0;
    for (int k = 0; k < nid; ++k) {
        double xan[4] = {0, 0, 0, 0};  // the next id's, fetched while this id streams
        if (k + 1 < nid && lr < nrhs) {
#pragma unroll
            for (int tq = 0; tq < 4; ++tq) xan[tq] = ids.sign[k + 1] * __ldg(ids.x[k + 1] + xa_off + 4 * tq);
        }
        for (int eb = 0; eb < ne; eb += 32) {
            const int cnt = min(32, ne - eb);
            if (ne > 32) mymir = lane < cnt ? m.mirror_row[e0 + eb + lane] : -1;
            for (int j = 0; j < cnt; ++j, ++piece) {
                const int m0 = __shfl_sync(0xffffffffu, mymir, j);
                mbar_wait(bar_a + 8 * stage, parity);
                const unsigned char* __restrict__ sp = ring + stage * STAGE;
                double xb[4], fb[4][2];
#pragma unroll
                for (int sb = 0; sb < 4; ++sb) {
                    xb[sb] = *reinterpret_cast<const double*>(sp + offx + 256 * sb);
#pragma unroll
                    for (int tb = 0; tb < 2; ++tb) fb[sb][tb] = *reinterpret_cast<const double*>(sp + offb[tb] + 512 * sb);
                }
                if (m0 >= 0) {  // warp-uniform; four independent accumulator chains
                    double za[4][2];
#pragma unroll
                    for (int tq = 0; tq < 4; ++tq)
#pragma unroll
                        for (int sg = 0; sg < 2; ++sg) za[tq][sg] = *reinterpret_cast<const double*>(sp + offa[tq] + 1024 * sg);
                    double2 z[2] = {make_double2(0, 0), make_double2(0, 0)};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        dmma884(accf[0], xb[q], fb[q][0]);
                        dmma884(accf[1], xb[q], fb[q][1]);
                        dmma884(z[0], za[q][0], xa[q]);
                        dmma884(z[1], za[q][1], xa[q]);
                    }
                    const size_t zo = (size_t)m0 * R;
                    if (k == 0) {
                        if (z_ok & 1) {
                            z0[zo] = z[0].x;
                            z0[zo + 8] = z[1].x;
                        }
                        if (z_ok & 2) {
                            z1[zo] = z[0].y;
                            z1[zo + 8] = z[1].y;
                        }
                    } else {
                        if (z_ok & 1) {
                            asm volatile("red.global.add.f64 [%0], %1;" ::"l"(z0 + zo), "d"(z[0].x) : "memory");
                            asm volatile("red.global.add.f64 [%0], %1;" ::"l"(z0 + zo + 8), "d"(z[1].x) : "memory");
                        }
                        if (z_ok & 2) {
                            asm volatile("red.global.add.f64 [%0], %1;" ::"l"(z1 + zo), "d"(z[0].y) : "memory");
                            asm volatile("red.global.add.f64 [%0], %1;" ::"l"(z1 + zo + 8), "d"(z[1].y) : "memory");
                        }
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        dmma884(accf[0], xb[q], fb[q][0]);
                        dmma884(accf[1], xb[q], fb[q][1]);
                    }
                }
                __syncwarp();  // every lane has read what it needs from this stage
                if (piece + STAGES < npieces) fill();
                if (++stage == STAGES) {
                    stage = 0;
                    parity ^= 1;
                }
            }
        }
#pragma unroll
        for (int tq = 0; tq < 4; ++tq) xa[tq] = xan[tq];
    }
    if (lr < nrhs) {
        double* __restrict__ out = scr + lr * sstride + (size_t)m.srow_ptr[g] * R + 2 * lc;
#pragma unroll
        for (int tb = 0; tb < 2; ++tb) *reinterpret_cast<double2*>(out + 8 * tb) = accf[tb];
    }
}

void launch_bsr_apply_ids(const DevCSR& m, const IdList& ids, size_t xstride, int nrhs, double* scr, size_t sstride, cudaStream_t st) {
    if (m.groups == 0 || nrhs == 0 || ids.n == 0) return;
    constexpr int STAGES = 3, SMEM = 8 * (STAGES * MMA_STAGE + STAGES * 8 + 32 * 4);
    static PerDeviceOnce once;  // per instantiation and device
    if (once.first()) {
        cudaFuncSetAttribute(bsr_apply_sym_mma_ids_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    }
    bsr_apply_sym_mma_ids_kernel<STAGES><<<(m.groups + 7) / 8, 256, SMEM, st>>>(m, ids, xstride, nrhs, scr, sstride);
}

template <int R>
static void launch_multi_R(const DevCSR& m, const double* A, const double* x, size_t xstride, int nrhs, double* scr, size_t sstride,
                           double sign, cudaStream_t st) {
    const unsigned ctas = (m.groups + 7) / 8;
    int r = 0;
    while (r < nrhs) {  // at most four right-hand sides per pass over the operator
        const int c = nrhs - r >= 4 ? 4 : nrhs - r == 3 ? 3 : nrhs - r == 2 ? 2 : 1;
        const double* xr = x + (size_t)r * xstride;
        double* sr = scr + (size_t)r * sstride;
        const bool regs = getenv("ANISO_STREAM_REGS") != nullptr;  // A/B switch: register-held loads instead of the TMA ring
        if (!regs) {
            if (c == 4) launch_ring<R, 4>(m, A, xr, xstride, sr, sstride, sign, st);
            else if (c == 3) launch_ring<R, 3>(m, A, xr, xstride, sr, sstride, sign, st);
            else if (c == 2) launch_ring<R, 2>(m, A, xr, xstride, sr, sstride, sign, st);
            else launch_ring<R, 1>(m, A, xr, xstride, sr, sstride, sign, st);
        } else if (c == 4) bsr_apply_sym_multi_kernel<R, 4><<<ctas, 256, 0, st>>>(m, A, xr, xstride, sr, sstride, sign);
        else if (c == 3) bsr_apply_sym_multi_kernel<R, 3><<<ctas, 256, 0, st>>>(m, A, xr, xstride, sr, sstride, sign);
        else if (c == 2) bsr_apply_sym_multi_kernel<R, 2><<<ctas, 256, 0, st>>>(m, A, xr, xstride, sr, sstride, sign);
        else bsr_apply_sym_multi_kernel<R, 1><<<ctas, 256, 0, st>>>(m, A, xr, xstride, sr, sstride, sign);
        r += c;
    }
}
// number of kernel launches is returned; requires m.symmetric and uniform 16 x 16 or 64 x 64 blocks
int launch_bsr_apply_multi(const DevCSR& m, const double* A, const double* x, size_t xstride, int nrhs, double* scr, size_t sstride,
                           double sign, cudaStream_t st) {
    if (m.groups == 0 || nrhs == 0) return 0;
    const bool no_mma = getenv("ANISO_NO_DMMA") != nullptr;  // A/B switch: DFMA / butterfly form
    if (m.uniform_block == 16 && m.swz && !no_mma) {  // tensor-core form: up to eight right-hand sides per pass
        int launches = 0;
        for (int r = 0; r < nrhs; r += 8, ++launches) {
            const int c = nrhs - r < 8 ? nrhs - r : 8;
            const double* xr = x + (size_t)r * xstride;
            double* sr = scr + (size_t)r * sstride;
            launch_mma(m, A, xr, xstride, c, sr, sstride, sign, st);
        }
        return launches;
    }
    if (m.uniform_block == 16) launch_multi_R<16>(m, A, x, xstride, nrhs, scr, sstride, sign, st);
    else launch_multi_R<64>(m, A, x, xstride, nrhs, scr, sstride, sign, st);
    return (nrhs + 3) / 4;
}

__global__ void __launch_bounds__(256) bsr_reduce_kernel(DevCSR m, double* __restrict__ y) {
    const int R = m.uniform_block;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int g = i / R, t = i - g * R;
    if (g >= m.groups) return;
    const int r0 = m.srow_ptr[g], r1 = m.srow_ptr[g + 1];
    double s0 = 0, s1 = 0;
    int r = r0;
    for (; r + 2 <= r1; r += 2) {
        s0 += m.scratch[(size_t)r * R + t];
        s1 += m.scratch[(size_t)(r + 1) * R + t];
    }
    if (r < r1) s0 += m.scratch[(size_t)r * R + t];
    y[m.row_off[g] + t] = s0 + s1;
}

// Symmetric storage for any uniform block size R <= 64 (np = 5, 6, 7: R = 25, 36, 49).  One warp per target; lane l
// owns targets l and l + 32, so a block row (R consecutive doubles) is one or two coalesced 8-byte loads per lane and
// eight rows are in flight per trip.  Forward product: acc[t] += A[s][t] x_B[s].  Transposed product for the partner:
// z[s] = sign * sum_t A[s][t] x_A[t]; the eight per-lane partials of a trip are reduced across the warp with a butterfly
// that halves the live values at each of its first three steps (4 + 2 + 1 + 1 + 1 shuffles instead of 8 x 5).
// One trip of eight rows [s0, s0+8) of a block (FULL: all eight exist).
template <int RC, bool FULL>
__device__ __forceinline__ void symg_trip(const double* __restrict__ rowp, const double* __restrict__ xs, int s0, int R, int lane,
                                          bool has0, bool has1, double xa0, double xa1, int mr, double sign, uint64_t pol,
                                          double& acc0, double& acc1, double* __restrict__ scratch) {
    constexpr bool TWO = RC == 0 || RC > 32;
    const int Rr = RC ? RC : R;
    double a0[8], a1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const bool ok = FULL || s0 + j < Rr;
        a0[j] = ok && has0 ? ld_stream(rowp + j * Rr, pol) : 0.0;
        a1[j] = TWO && ok && has1 ? ld_stream(rowp + j * Rr + 32, pol) : 0.0;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const double xv = (FULL || s0 + j < Rr) ? __ldg(xs + j) : 0.0;
        acc0 = fma(a0[j], xv, acc0);
        if (TWO) acc1 = fma(a1[j], xv, acc1);
    }
    if (mr < 0) return;  // warp-uniform
    double p[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) p[j] = TWO ? fma(a0[j], xa0, a1[j] * xa1) : a0[j] * xa0;
    const bool b0 = lane & 1, b1 = lane & 2, b2 = lane & 4;
    double q[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)  // lane bit 0 keeps rows {4..7} if set, else {0..3}
        q[j] = (b0 ? p[4 + j] : p[j]) + __shfl_xor_sync(0xffffffffu, b0 ? p[j] : p[4 + j], 1);
    const double w0 = (b1 ? q[2] : q[0]) + __shfl_xor_sync(0xffffffffu, b1 ? q[0] : q[2], 2);
    const double w1 = (b1 ? q[3] : q[1]) + __shfl_xor_sync(0xffffffffu, b1 ? q[1] : q[3], 2);
    double z = (b2 ? w1 : w0) + __shfl_xor_sync(0xffffffffu, b2 ? w0 : w1, 4);
    z += __shfl_xor_sync(0xffffffffu, z, 8);
    z += __shfl_xor_sync(0xffffffffu, z, 16);
    // lane (low three bits b2 b1 b0 as read above) holds row 4*b0 + 2*b1 + b2 of this trip
    const int row = s0 + (b0 ? 4 : 0) + (b1 ? 2 : 0) + (b2 ? 1 : 0);
    if (lane < 8 && (FULL || row < Rr)) scratch[(size_t)mr * Rr + row] = sign * z;
}

template <int RC>  // RC: compile-time block size (25, 36, 49), or 0 = run-time m.uniform_block
__global__ void __launch_bounds__(256, 3) bsr_apply_symg_kernel(DevCSR m, const double* __restrict__ A, const double* __restrict__ x,
                                                                double sign) {
    const int R = RC ? RC : m.uniform_block;
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= m.groups) return;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const uint64_t pol = stream_policy();
    const double* __restrict__ blk = A + (ne ? m.blk_off[e0] : 0) + lane;
    const int self = m.row_off[g];
    const bool has0 = lane < R, has1 = lane + 32 < R;
    const double xa0 = has0 ? __ldg(x + self + lane) : 0.0, xa1 = has1 ? __ldg(x + self + lane + 32) : 0.0;
    double* __restrict__ scratch = m.scratch;
    double acc0 = 0, acc1 = 0;
    const int full = R & ~7;
    for (int e = 0; e < ne; ++e, blk += R * R) {
        const int mr = m.mirror_row[e0 + e];
        const double* __restrict__ xs = x + m.col_off[e0 + e];
        int s0 = 0;
        for (; s0 < full; s0 += 8)
            symg_trip<RC, true>(blk + s0 * R, xs + s0, s0, R, lane, has0, has1, xa0, xa1, mr, sign, pol, acc0, acc1, scratch);
        if (s0 < R) symg_trip<RC, false>(blk + s0 * R, xs + s0, s0, R, lane, has0, has1, xa0, xa1, mr, sign, pol, acc0, acc1, scratch);
    }
    double* __restrict__ out = scratch + (size_t)m.srow_ptr[g] * R;
    if (has0) out[lane] = acc0;
    if (has1) out[lane + 32] = acc1;
}

// ---- the same, fed by the TMA engine -------------------------------------------------------------------------------
// The register-load form above holds 8 rows per warp in flight (24 warps x 8 x 8R bytes = 38 - 75 KB per SM) and ncu
// shows it waiting on those loads (long-scoreboard stalls, 4.0 - 5.2 TB/s).  As for 16 x 16 blocks, the blocks of one
// target are one contiguous run, here of ne * R rows of R doubles; the copy engine fetches it in pieces of EIGHT ROWS
// (64 R bytes, a multiple of 16 for every R) into a private ring, so the bytes in flight (warps x STAGES x 64 R) no
// longer depend on the register budget.  A piece may straddle two blocks: every row of a piece carries its own (block,
// source index), i.e. its own x_B value and its own partner scratch row; the forward accumulators and x_A do not change
// across the blocks of a target.  For odd R a target's run may start on an odd double: the ring copies from the 16-byte
// boundary below it (`skew`), each stage holds 16 bytes more than a piece, and the operator allocation is padded by 16
// bytes (capi.cu ensure_ops).
template <int R, int STAGES, int MINB>
__global__ void __launch_bounds__(256, MINB) bsr_apply_symg_ring_kernel(DevCSR m, const double* __restrict__ A,
                                                                        const double* __restrict__ x, double sign) {
    constexpr bool TWO = R > 32;
    constexpr int PIECE = 64 * R, STAGE = PIECE + 16;
    extern __shared__ __align__(128) unsigned char ring_smem[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + wid;
    if (g >= m.groups) return;  // warps are independent: no CTA-wide barrier below
    unsigned char* ring = ring_smem + (size_t)wid * STAGES * STAGE;
    const uint32_t ring_a = smem_u32(ring);
    const uint32_t bar_a = smem_u32(ring_smem + 8 * STAGES * STAGE + wid * STAGES * 8);
    double* __restrict__ xsh = reinterpret_cast<double*>(ring_smem + 8 * STAGES * STAGE + 8 * STAGES * 8) + wid * 8;
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const int64_t off0 = ne ? m.blk_off[e0] : 0;
    const int skew = (int)(off0 & 1);
    const unsigned char* __restrict__ src = reinterpret_cast<const unsigned char*>(A + (off0 - skew));
    const int nrows = ne * R, npieces = (nrows + 7) >> 3;
    const uint64_t pol = stream_policy();
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar_a + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto fill = [&](int q, int stage) {  // producer (lane 0): piece q -> stage
        const int rows = min(8, nrows - 8 * q);
        const uint32_t bytes = (uint32_t)(8 * (skew + rows * R) + 15) & ~15u;
        const uint32_t bar = bar_a + 8 * stage;
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring_a + stage * STAGE, src + (size_t)q * PIECE, bytes, bar, pol);
    };
    if (lane == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < STAGES; ++s)
            if (s < npieces) fill(s, s);
    }
    const int self = m.row_off[g];
    const bool has0 = lane < R, has1 = TWO && lane + 32 < R;
    const double xa0 = has0 ? __ldg(x + self + lane) : 0.0, xa1 = has1 ? __ldg(x + self + lane + 32) : 0.0;
    double* __restrict__ scratch = m.scratch;
    const bool b0 = lane & 1, b1 = lane & 2, b2 = lane & 4;
    const int l8 = lane & 7, rperm = (b0 ? 4 : 0) + (b1 ? 2 : 0) + (b2 ? 1 : 0);  // row of a piece whose z this lane ends up with
    // per-row metadata of piece q, fetched one piece ahead: x_B of row l8 (natural order, published through xsh) and the
    // partner scratch position of row rperm
    auto meta = [&](int q, double& xv, int& mrow, int& srow) {
        const int i = 8 * q + l8, ir = 8 * q + rperm;
        xv = 0.0;
        mrow = -1;
        srow = 0;
        if (i < nrows) {
            const int e = i / R;
            xv = __ldg(x + m.col_off[e0 + e] + (i - e * R));
        }
        if (ir < nrows) {
            const int e = ir / R;
            srow = ir - e * R;
            mrow = m.mirror_row[e0 + e];
        }
    };
    double xv_c, acc0 = 0, acc1 = 0;
    int mr_c, sr_c;
    meta(0, xv_c, mr_c, sr_c);
    int stage = 0;
    uint32_t parity = 0;
#pragma unroll 1
    for (int q = 0; q < npieces; ++q) {
        double xv_n;
        int mr_n, sr_n;
        meta(q + 1, xv_n, mr_n, sr_n);
        if (lane < 8) xsh[lane] = xv_c;
        const int rows = min(8, nrows - 8 * q);
        mbar_wait(bar_a + 8 * stage, parity);
        __syncwarp();  // xsh
        const double* __restrict__ sp = reinterpret_cast<const double*>(ring + stage * STAGE) + skew + lane;
        double a0[8], a1[8], xs[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            a0[j] = has0 && j < rows ? sp[j * R] : 0.0;
            a1[j] = has1 && j < rows ? sp[j * R + 32] : 0.0;
        }
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
            const double2 v = reinterpret_cast<const double2*>(xsh)[j >> 1];
            xs[j] = v.x;
            xs[j + 1] = v.y;
        }
        __syncwarp();  // every lane has read what it needs from this stage (and from xsh)
        if (lane == 0 && q + STAGES < npieces) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            fill(q + STAGES, stage);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc0 = fma(a0[j], xs[j], acc0);
            if (TWO) acc1 = fma(a1[j], xs[j], acc1);
        }
        if (__any_sync(0xffffffffu, mr_c >= 0)) {
            double p[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) p[j] = TWO ? fma(a0[j], xa0, a1[j] * xa1) : a0[j] * xa0;
            double qq[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) qq[j] = (b0 ? p[4 + j] : p[j]) + __shfl_xor_sync(0xffffffffu, b0 ? p[j] : p[4 + j], 1);
            const double w0 = (b1 ? qq[2] : qq[0]) + __shfl_xor_sync(0xffffffffu, b1 ? qq[0] : qq[2], 2);
            const double w1 = (b1 ? qq[3] : qq[1]) + __shfl_xor_sync(0xffffffffu, b1 ? qq[1] : qq[3], 2);
            double z = (b2 ? w1 : w0) + __shfl_xor_sync(0xffffffffu, b2 ? w0 : w1, 4);
            z += __shfl_xor_sync(0xffffffffu, z, 8);
            z += __shfl_xor_sync(0xffffffffu, z, 16);
            if (lane < 8 && mr_c >= 0) scratch[(size_t)mr_c * R + sr_c] = sign * z;
        }
        xv_c = xv_n;
        mr_c = mr_n;
        sr_c = sr_n;
        if (++stage == STAGES) {
            stage = 0;
            parity ^= 1;
        }
    }
    double* __restrict__ out = scratch + (size_t)m.srow_ptr[g] * R;
    if (has0) out[lane] = acc0;
    if (has1) out[lane + 32] = acc1;
}

template <int R, int STAGES, int MINB>
static void launch_symg_ring(const DevCSR& m, const double* A, const double* x, double sign, cudaStream_t st) {
    constexpr int SMEM = 8 * STAGES * (64 * R + 16) + 8 * STAGES * 8 + 8 * 8 * 8;
    static PerDeviceOnce once;  // per instantiation and device
    if (once.first()) {
        cudaFuncSetAttribute(bsr_apply_symg_ring_kernel<R, STAGES, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    }
    bsr_apply_symg_ring_kernel<R, STAGES, MINB><<<(m.groups + 7) / 8, 256, SMEM, st>>>(m, A, x, sign);
}

// ---- eight-row pieces with the transposed product on the FP64 tensor cores -----------------------------------------
// ncu of the kernel above at R = 25 / 36: issue slots 77 / 73 % busy, ~270 / 335 instructions per piece, a third of them
// the butterfly (18 SHFL + 28 FSEL + DADD / DMUL) and much of the rest address arithmetic re-derived under register
// pressure.  The transposed product of a piece, z[row] = sum_t P[row][t] x_A[t], is an (8 x R)(R x 1) product: as
// ceil(R / 4) mma.m8n8k4.f64 with P read from the stage as the A operand (lane l: P[row l/4][4 kt + l%4]) and x_A
// replicated over the eight columns of the B operand, lane 4 row of the accumulator fragment holds z[row] with no
// cross-lane reduction at all.  The forward product stays on DFMA (lane = target; one LDS + one DFMA per row).  Rows are
// tracked as (block, source index) pairs advanced by eight per piece instead of divided out.  The ring is zeroed before
// the first copy so that the rows behind the end of a target's run (xs = 0, no partner row) multiply finite values.
__device__ __forceinline__ double lds_f64(uint32_t a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ double2 lds_v2f64(uint32_t a) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}

template <int R, int STAGES, int MINB>
__global__ void __launch_bounds__(256, MINB) bsr_apply_symg_mma_kernel(DevCSR m, const double* __restrict__ A,
                                                                       const double* __restrict__ x, double sign) {
    constexpr bool TWO = R > 32;
    constexpr int PIECE = 64 * R, STAGE = PIECE + 16, KT = (R + 3) / 4;
    extern __shared__ __align__(128) unsigned char ring_smem[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + wid;
    if (g >= m.groups) return;  // warps are independent: no CTA-wide barrier below
    unsigned char* ring = ring_smem + (size_t)wid * STAGES * STAGE;
    uint32_t ring_a = smem_u32(ring);
    uint32_t bar_a = smem_u32(ring_smem + 8 * STAGES * STAGE + wid * STAGES * 8);
    uint32_t xsh_a = smem_u32(ring_smem + 8 * STAGES * STAGE + 8 * STAGES * 8 + wid * 64);
    for (int o = lane * 16; o < STAGES * STAGE; o += 512) *reinterpret_cast<double2*>(ring + o) = make_double2(0.0, 0.0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const int e0 = m.ent_ptr[g], ne = m.ent_ptr[g + 1] - e0;
    const int64_t off0 = ne ? m.blk_off[e0] : 0;
    const int skew = (int)(off0 & 1);
    const unsigned char* __restrict__ src = reinterpret_cast<const unsigned char*>(A + (off0 - skew));
    const int nrows = ne * R, npieces = (nrows + 7) >> 3;
    const uint64_t pol = stream_policy();
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar_a + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    // producer (lane 0): piece q -> the stage at shared address sa with barrier ba; every piece but the last has 8 rows
    const uint32_t full_bytes = (uint32_t)(8 * (skew + 8 * R) + 15) & ~15u;
    const uint32_t last_bytes = (uint32_t)(8 * (skew + (nrows - 8 * (npieces - 1)) * R) + 15) & ~15u;
    auto fill = [&](int q, uint32_t sa, uint32_t ba) {
        const uint32_t bytes = q == npieces - 1 ? last_bytes : full_bytes;
        mbar_expect_tx(ba, bytes);
        bulk_g2s(sa, src + (size_t)q * PIECE, bytes, ba, pol);
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s)
            if (s < npieces) fill(s, ring_a + s * STAGE, bar_a + 8 * s);
    }
    const int self = m.row_off[g];
    const int lr = lane >> 2, lc = lane & 3;
    const bool has0 = lane < R, has1 = TWO && lane + 32 < R;
    double xak[KT];  // B operand of the transposed product: sign * x_A[4 kt + lc], the same in all eight columns
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) xak[kt] = 4 * kt + lc < R ? sign * __ldg(x + self + 4 * kt + lc) : 0.0;
    double* __restrict__ scratch = m.scratch;
    // Byte offsets of this lane inside a stage.  Lanes without a target / past the end of a row read a valid location
    // instead of being predicated off (their products are multiplied by zero or never stored), so every load of the loop
    // is base register + immediate.
    uint32_t offf = 8 * (skew + min(lane, R - 1));                       // forward rows: lane = target
    uint32_t offf1 = 8 * (skew + 32 + min(lane, TWO ? R - 33 : 0));      //   second half of a row (R > 32)
    uint32_t offa = 8 * (skew + lr * R + lc);                            // A-operand fragment: row lr, column 4 kt + lc
    uint32_t offa_last = 4 * (KT - 1) + lc < R ? offa + 32 * (KT - 1) : offa;  // R % 4 != 0: last k-step, lanes past the row
    ANISO_PIN(offf); ANISO_PIN(offf1); ANISO_PIN(offa); ANISO_PIN(offa_last); ANISO_PIN(ring_a); ANISO_PIN(bar_a); ANISO_PIN(xsh_a);
    // (block, source index) of the row whose x_B this lane gathers (row lane % 8 of the piece; lanes 0-7 publish it) and of
    // the row whose z it ends up with (row lr; lanes lc == 0 write it)
    int ex = 0, sx = lane & 7, ez = 0, sz = lr;
    const int* __restrict__ colp = m.col_off + e0;
    const int* __restrict__ mirp = m.mirror_row + e0;
    double xv_c = ex < ne ? __ldg(x + colp[ex] + sx) : 0.0;
    int mr_c = ez < ne ? mirp[ez] : -1;
    double acc0 = 0, acc1 = 0;
    uint32_t sa = ring_a, ba = bar_a, parity = 0;
    int stage = 0;
#pragma unroll 1
    for (int q = 0; q < npieces; ++q) {
        const int sz_c = sz;
        sx += 8;
        if (sx >= R) {
            sx -= R;
            ++ex;
        }
        sz += 8;
        if (sz >= R) {
            sz -= R;
            ++ez;
        }
        const double xv_n = ex < ne ? __ldg(x + colp[ex] + sx) : 0.0;
        const int mr_n = ez < ne ? mirp[ez] : -1;
        if (lane < 8) asm volatile("st.shared.f64 [%0], %1;" ::"r"(xsh_a + 8 * lane), "d"(xv_c) : "memory");
        mbar_wait(ba, parity);
        __syncwarp();  // xsh
        {
            double a0[8], a1[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                a0[j] = lds_f64(sa + offf + 8 * j * R);
                if (TWO) a1[j] = lds_f64(sa + offf1 + 8 * j * R);
            }
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                const double2 v = lds_v2f64(xsh_a + 8 * j);
                acc0 = fma(a0[j], v.x, acc0);
                acc0 = fma(a0[j + 1], v.y, acc0);
                if (TWO) {
                    acc1 = fma(a1[j], v.x, acc1);
                    acc1 = fma(a1[j + 1], v.y, acc1);
                }
            }
        }
        if (__any_sync(0xffffffffu, mr_c >= 0)) {
            double2 z0 = make_double2(0.0, 0.0), z1 = make_double2(0.0, 0.0);
#pragma unroll
            for (int kt = 0; kt < KT; ++kt) {
                const double a = lds_f64(kt == KT - 1 && (R & 3) ? sa + offa_last : sa + offa + 32 * kt);
                if (kt & 1) dmma884(z1, a, xak[kt]);
                else dmma884(z0, a, xak[kt]);
            }
            if (lc == 0 && mr_c >= 0) scratch[(size_t)mr_c * R + sz_c] = z0.x + z1.x;
        }
        __syncwarp();  // every lane's reads of this stage (and of xsh) have returned: the DFMA / DMMA above consumed them
        if (lane == 0 && q + STAGES < npieces) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            fill(q + STAGES, sa, ba);
        }
        xv_c = xv_n;
        mr_c = mr_n;
        sa += STAGE;
        ba += 8;
        if (++stage == STAGES) {
            stage = 0;
            sa = ring_a;
            ba = bar_a;
            parity ^= 1;
        }
    }
    double* __restrict__ out = scratch + (size_t)m.srow_ptr[g] * R;
    if (has0) out[lane] = acc0;
    if (has1) out[lane + 32] = acc1;
}

template <int R, int STAGES, int MINB>
static void launch_symg_mma(const DevCSR& m, const double* A, const double* x, double sign, cudaStream_t st) {
    constexpr int SMEM = 8 * STAGES * (64 * R + 16) + 8 * STAGES * 8 + 8 * 8 * 8;
    static PerDeviceOnce once;  // per instantiation and device
    if (once.first()) {
        cudaFuncSetAttribute(bsr_apply_symg_mma_kernel<R, STAGES, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    }
    bsr_apply_symg_mma_kernel<R, STAGES, MINB><<<(m.groups + 7) / 8, 256, SMEM, st>>>(m, A, x, sign);
}

__global__ void __launch_bounds__(128) bsr_apply_row_kernel(DevCSR m, const double* __restrict__ A, const double* __restrict__ x,
                                                            double* __restrict__ y) {
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= m.nrows_total) return;
    int g, t, nrow;
    if (m.uniform_nrow) {
        nrow = m.uniform_nrow;
        g = row / nrow;
        t = row - g * nrow;
    } else {
        g = m.row_group[row];
        t = row - m.grow_ptr[g];
        nrow = m.nrow[g];
    }
    const int e0 = m.ent_ptr[g], e1 = m.ent_ptr[g + 1];
    const uint64_t pol = stream_policy();
    double acc[4] = {0, 0, 0, 0};
    // software pipeline over the entry metadata: the loads of entry e+1 are issued before the block of e is consumed
    int64_t boff = e0 < e1 ? m.blk_off[e0] : 0;
    int coff = e0 < e1 ? m.col_off[e0] : 0, nc = e0 < e1 ? m.ncol[e0] : 0;
    for (int e = e0; e < e1; ++e) {
        const double* __restrict__ blk = A + boff + t;
        const double* __restrict__ xs = x + coff;
        const int ncur = nc;
        if (e + 1 < e1) {
            boff = m.blk_off[e + 1];
            coff = m.col_off[e + 1];
            nc = m.ncol[e + 1];
        }
        int s = 0;
        for (; s + 8 <= ncur; s += 8) {
            double a[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) a[u] = ld_stream(blk + (size_t)(s + u) * nrow, pol);
#pragma unroll
            for (int u = 0; u < 8; ++u) acc[u & 3] = fma(a[u], __ldg(xs + s + u), acc[u & 3]);
        }
        for (; s < ncur; ++s) acc[0] = fma(ld_stream(blk + (size_t)s * nrow, pol), __ldg(xs + s), acc[0]);
    }
    y[m.row_off[g] + t] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

void launch_bsr_reduce(const DevCSR& m, double* y, cudaStream_t st) {
    if (!m.symmetric || m.groups == 0) return;
    const unsigned rb = ((unsigned)m.groups * m.uniform_block + 255) / 256;
    bsr_reduce_kernel<<<rb, 256, 0, st>>>(m, y);
}

void launch_bsr_apply(const DevCSR& m, const double* A, const double* x, double* y, double sign, bool reduce, cudaStream_t st) {
    if (m.nrows_total == 0) return;
    const unsigned warps_per_cta = 8, ctas = (m.groups + warps_per_cta - 1) / warps_per_cta;
    if (m.symmetric) {
        const bool ring = getenv("ANISO_STREAM_REGS") == nullptr;  // A/B switch: register-held loads instead of the TMA ring
        if (ring && m.uniform_block == 16) launch_ring<16, 1>(m, A, x, 0, m.scratch, 0, sign, st);
        else if (ring && m.uniform_block == 64) launch_ring<64, 1>(m, A, x, 0, m.scratch, 0, sign, st);
        else if (m.uniform_block == 16) bsr_apply_sym_kernel<16><<<ctas, 256, 0, st>>>(m, A, x, sign);
        else if (m.uniform_block == 64) bsr_apply_sym_kernel<64><<<ctas, 256, 0, st>>>(m, A, x, sign);
        else if (ring && m.uniform_block == 25 && !getenv("ANISO_SYMG_NO_DMMA")) launch_symg_mma<25, 4, 4>(m, A, x, sign, st);
        else if (ring && m.uniform_block == 36 && !getenv("ANISO_SYMG_NO_DMMA")) launch_symg_mma<36, 3, 3>(m, A, x, sign, st);
        else if (ring && m.uniform_block == 25) launch_symg_ring<25, 4, 4>(m, A, x, sign, st);
        else if (ring && m.uniform_block == 36) launch_symg_ring<36, 3, 3>(m, A, x, sign, st);
        else if (ring && m.uniform_block == 49) launch_symg_ring<49, 4, 2>(m, A, x, sign, st);
        else if (m.uniform_block == 25) bsr_apply_symg_kernel<25><<<ctas, 256, 0, st>>>(m, A, x, sign);
        else if (m.uniform_block == 36) bsr_apply_symg_kernel<36><<<ctas, 256, 0, st>>>(m, A, x, sign);
        else if (m.uniform_block == 49) bsr_apply_symg_kernel<49><<<ctas, 256, 0, st>>>(m, A, x, sign);
        else bsr_apply_symg_kernel<0><<<ctas, 256, 0, st>>>(m, A, x, sign);
        if (reduce) launch_bsr_reduce(m, y, st);
        return;
    }
    static const bool force_row = getenv("ANISO_BSR_ROW") != nullptr;  // A/B switch for profiling only
    if (force_row)
        bsr_apply_row_kernel<<<(m.nrows_total + 127) / 128, 128, 0, st>>>(m, A, x, y);
    else if (m.uniform_block == 16)
        bsr_apply_warp_kernel<16><<<ctas, 256, 0, st>>>(m, A, x, y);
    else if (m.uniform_block == 64)
        bsr_apply_warp_kernel<64><<<ctas, 256, 0, st>>>(m, A, x, y);
    else
        bsr_apply_row_kernel<<<(m.nrows_total + 127) / 128, 128, 0, st>>>(m, A, x, y);
}

}  // namespace an
