// Collectives of the sharded matvec over NVLink / NVSwitch peer memory (one process per GPU, CUDA IPC).
//
// The exchange steps of the sharded path move little data - 8 MB of source values and 11 MB of multipoles per
// operator application at N = 2^20, a few hundred bytes per Gram-Schmidt reduction - so they are bound by latency, and
// NCCL's ring / tree schedules cost 45 - 90 us per all-gather and ~20 us per all-reduce on 8 GPUs (measured, profiles/).
// On an NVSwitch box every GPU can store into every other GPU's memory directly, so here each collective is ONE push:
//   all-gather      every rank copies its part into the same place of a buffer on EVERY rank (plain 16-byte stores through
//                   the IPC mappings), fences, and then raises its flag on every rank; a rank's consumers wait until all
//                   flags show the epoch of this exchange
//   small all-reduce  every rank pushes its partial sums to a slot of its own on every rank; every rank then adds the slots in
//                   rank order, so all ranks hold bit-identical results (GMRES stays in lock-step, as with ncclAllReduce)
// Buffers are double-buffered by the parity of an epoch counter every rank advances alike (the entry points are SPMD): a rank
// can be at most one exchange ahead of the slowest rank, because finishing exchange e needs every rank's push of e.
// Every wait carries a time-out (error flag in pinned host memory) instead of spinning for ever on a lost peer.
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#include "device.cuh"
#include "peer.hpp"

namespace an {

namespace {

__device__ __forceinline__ unsigned long long ld_flag(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_flag(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// waits until flags[0..P) >= epoch; ~2 s time-out at 1.9 GHz
__device__ __forceinline__ bool wait_all(const unsigned long long* flags, int P, unsigned long long epoch, int* err) {
    const long long t0 = clock64();
    for (int r = 0; r < P; ++r) {
        while (ld_flag(flags + r) < epoch) {
            if (clock64() - t0 > 4000000000ll) {
                *err = 1;
                return false;
            }
        }
    }
    return true;
}

// All-gather by push: src (n doubles, n even, 16-byte aligned) -> dst[r] + 0..n on every rank r (dst[r] already points at
// this rank's place in rank r's buffer).  The last block to finish raises this rank's flag on every rank.
__global__ void __launch_bounds__(256) peer_push_kernel(const double* __restrict__ src, size_t n, PeerPtrs dst, PeerFlags flags, int P,
                                                        int me, unsigned long long epoch, unsigned int* __restrict__ counter) {
    const size_t n2 = n >> 1;
    const double2* __restrict__ s2 = reinterpret_cast<const double2*>(src);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) {
        const double2 v = s2[i];
#pragma unroll 1
        for (int r = 0; r < P; ++r) reinterpret_cast<double2*>(dst.p[r])[i] = v;
    }
    __threadfence_system();
    __shared__ bool last;
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(counter, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence_system();
    if (threadIdx.x < P) st_flag(flags.p[threadIdx.x] + me, epoch);
    if (threadIdx.x == 0) *counter = 0;
}

__global__ void peer_wait_kernel(const unsigned long long* __restrict__ flags, int P, unsigned long long epoch, int* __restrict__ err) {
    if (threadIdx.x == 0) wait_all(flags, P, epoch, err);
}

// Small all-reduce (sum) of vals[0..n): push to slot `me` of every rank, wait for all slots, add in rank order.
// One block; red[r] = base of rank r's slot array (P slots of `cap` doubles), flags as above.
__global__ void __launch_bounds__(256) peer_allreduce_kernel(double* __restrict__ vals, int n, int cap, PeerPtrs red, PeerFlags flags,
                                                             const double* __restrict__ my_red, const unsigned long long* __restrict__ my_flags,
                                                             int P, int me, unsigned long long epoch, int* __restrict__ err) {
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        const double v = vals[k];
        for (int r = 0; r < P; ++r) red.p[r][(size_t)me * cap + k] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < P) st_flag(flags.p[threadIdx.x] + me, epoch);
    __shared__ bool ok;
    if (threadIdx.x == 0) ok = wait_all(my_flags, P, epoch, err);
    __syncthreads();
    if (!ok) return;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        double s = 0;
        for (int r = 0; r < P; ++r) s += __ldcg(my_red + (size_t)r * cap + k);
        vals[k] = s;
    }
}

}  // namespace

// ---- host side --------------------------------------------------------------------------------------------------
#define PCK(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) throw std::runtime_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// Layout of the symmetric region (the same on every rank), in bytes from its base:
//   flags   3 kinds x 2 parities x kMaxPeers u64        (written by the peers)
//   red     2 parities x P slots x red_cap doubles
//   x       2 parities x nx doubles                     gathered source vector (slot order)
//   mc      2 parities x nmc doubles                    gathered multipoles (rank-major, see DevTree::pack_*)
void peer_layout(PeerComm& pc, size_t nx, size_t nmc, int red_cap) {
    pc.nx = nx;
    pc.nmc = nmc;
    pc.red_cap = red_cap;
    size_t off = 0;
    pc.off_flags = off;
    off += sizeof(unsigned long long) * 3 * 2 * kMaxPeers;
    off = (off + 255) & ~size_t(255);
    pc.off_red = off;
    off += sizeof(double) * 2 * pc.P * red_cap;
    off = (off + 255) & ~size_t(255);
    pc.off_x = off;
    off += sizeof(double) * 2 * nx;
    off = (off + 255) & ~size_t(255);
    pc.off_mc = off;
    off += sizeof(double) * 2 * nmc;
    pc.bytes = (off + 255) & ~size_t(255);
}

// exchange(handle bytes in, all ranks' handle bytes out): supplied by the caller (an NCCL all-gather in capi.cu)
void peer_open(PeerComm& pc, const std::function<void(const void*, void*, size_t)>& exchange) {
    PCK(cudaMalloc(&pc.local, pc.bytes));
    PCK(cudaMemset(pc.local, 0, pc.bytes));
    PCK(cudaMalloc(&pc.counter, sizeof(unsigned int)));
    PCK(cudaMemset(pc.counter, 0, sizeof(unsigned int)));
    PCK(cudaHostAlloc(&pc.err, sizeof(int), cudaHostAllocMapped));
    *pc.err = 0;
    PCK(cudaHostGetDevicePointer(&pc.err_dev, pc.err, 0));
    PCK(cudaDeviceSynchronize());
    cudaIpcMemHandle_t mine;
    PCK(cudaIpcGetMemHandle(&mine, pc.local));
    std::vector<cudaIpcMemHandle_t> all(pc.P);
    exchange(&mine, all.data(), sizeof(mine));
    for (int r = 0; r < pc.P; ++r) {
        if (r == pc.me) {
            pc.remote[r] = pc.local;
            continue;
        }
        void* p = nullptr;
        PCK(cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess));
        pc.remote[r] = static_cast<char*>(p);
    }
    pc.ready = true;
}

void peer_close(PeerComm& pc) {
    if (!pc.local) return;
    for (int r = 0; r < pc.P; ++r)
        if (r != pc.me && pc.remote[r]) cudaIpcCloseMemHandle(pc.remote[r]);
    cudaFree(pc.local);
    cudaFree(pc.counter);
    if (pc.err) cudaFreeHost(pc.err);
    pc = PeerComm();
}

static PeerFlags flag_ptrs(const PeerComm& pc, int kind, int par) {
    PeerFlags f;
    for (int r = 0; r < pc.P; ++r)
        f.p[r] = reinterpret_cast<unsigned long long*>(pc.remote[r] + pc.off_flags) + ((size_t)kind * 2 + par) * kMaxPeers;
    return f;
}

double* peer_x(const PeerComm& pc, int par) { return reinterpret_cast<double*>(pc.local + pc.off_x) + (size_t)par * pc.nx; }
double* peer_mc(const PeerComm& pc, int par) { return reinterpret_cast<double*>(pc.local + pc.off_mc) + (size_t)par * pc.nmc; }

// kind 0: source vector (own part = [own_off, own_off + n) of the x buffer), kind 1: multipoles (same of the mc buffer).
// src may already be the own part of this rank's buffer (multipoles written in place by the upward pass): that copy is
// skipped by pointing the local destination at itself - the store is idempotent.
void peer_push(PeerComm& pc, int kind, int par, unsigned long long epoch, const double* src, size_t own_off, size_t n, cudaStream_t st) {
    PeerPtrs d;
    const size_t base = kind == 0 ? pc.off_x + sizeof(double) * par * pc.nx : pc.off_mc + sizeof(double) * par * pc.nmc;
    for (int r = 0; r < pc.P; ++r) d.p[r] = reinterpret_cast<double*>(pc.remote[r] + base) + own_off;
    const int blocks = (int)std::min<size_t>(296, std::max<size_t>(1, (n / 2 + 255) / 256));
    peer_push_kernel<<<blocks, 256, 0, st>>>(src, n, d, flag_ptrs(pc, kind, par), pc.P, pc.me, epoch, pc.counter);
}

void peer_wait(PeerComm& pc, int kind, int par, unsigned long long epoch, cudaStream_t st) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(pc.local + pc.off_flags) + ((size_t)kind * 2 + par) * kMaxPeers;
    peer_wait_kernel<<<1, 32, 0, st>>>(f, pc.P, epoch, pc.err_dev);
}

void peer_allreduce(PeerComm& pc, double* vals, int n, cudaStream_t st) {
    if (n > pc.red_cap) throw std::runtime_error("peer_allreduce: too many values");
    const unsigned long long epoch = ++pc.epoch_red;
    const int par = (int)(epoch & 1);
    PeerPtrs red;
    for (int r = 0; r < pc.P; ++r) red.p[r] = reinterpret_cast<double*>(pc.remote[r] + pc.off_red) + (size_t)par * pc.P * pc.red_cap;
    const double* my_red = reinterpret_cast<const double*>(pc.local + pc.off_red) + (size_t)par * pc.P * pc.red_cap;
    const unsigned long long* my_flags = reinterpret_cast<const unsigned long long*>(pc.local + pc.off_flags) + ((size_t)2 * 2 + par) * kMaxPeers;
    peer_allreduce_kernel<<<1, 256, 0, st>>>(vals, n, pc.red_cap, red, flag_ptrs(pc, 2, par), my_red, my_flags, pc.P, pc.me, epoch, pc.err_dev);
}

}  // namespace an
