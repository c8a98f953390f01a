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
// Every wait carries a time-out (~20 s; error flag in pinned host memory) instead of spinning for ever on a lost peer.
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#include "device.cuh"
#include "peer.hpp"
#include "peer_dev.cuh"

namespace an {

namespace {

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
    if (threadIdx.x < P) peer_st_flag(flags.p[threadIdx.x] + me, epoch);
    if (threadIdx.x == 0) *counter = 0;
}

__global__ void peer_wait_kernel(const unsigned long long* __restrict__ flags, int P, unsigned long long epoch, int* __restrict__ err) {
    if (threadIdx.x == 0) peer_wait_all(flags, P, epoch, err);
}

// Small all-reduce (sum) of vals[0..n) as a kernel of its own (one block); the reduction kernels of the GMRES passes carry the
// same exchange in their epilogue (PeerRed in device.cuh).
__global__ void __launch_bounds__(256) peer_allreduce_kernel(double* __restrict__ vals, int n, const __grid_constant__ PeerRed pr) {
    peer_allreduce_block(vals, n, pr);
}

}  // namespace

// ---- host side --------------------------------------------------------------------------------------------------
#define PCK(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) throw std::runtime_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// Layout of the symmetric region (the same on every rank), in bytes from its base:
//   flags   4 kinds (source, multipoles, small reductions, leaf-level multipoles) x 2 parities x kMaxPeers u64   (written by the peers)
//   red     2 parities x P slots x red_cap doubles
//   x       2 parities x nx doubles                     gathered source vector (slot order)
//   mc      2 parities x nmc doubles                    gathered multipoles (rank-major, see DevTree::pack_*)
void peer_layout(PeerComm& pc, size_t nx, size_t nmc, int red_cap) {
    pc.nx = nx;
    pc.nmc = nmc;
    pc.red_cap = red_cap;
    size_t off = 0;
    pc.off_flags = off;
    off += sizeof(unsigned long long) * 4 * 2 * kMaxPeers;
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
    PCK(cudaMalloc(&pc.counter, 4 * sizeof(unsigned int)));  // one per kind: pushes of different kinds run on different streams
    PCK(cudaMemset(pc.counter, 0, 4 * sizeof(unsigned int)));
    PCK(cudaHostAlloc(&pc.err, sizeof(int), cudaHostAllocMapped));
    *pc.err = 0;
    PCK(cudaHostGetDevicePointer(&pc.err_dev, pc.err, 0));
    PCK(cudaDeviceSynchronize());
    // what a rank publishes: an IPC handle for ranks in other processes, the raw pointer and its device for ranks that are
    // threads of this process (CUDA IPC cannot map a buffer into the process that owns it; unified addressing plus peer
    // access makes the pointer itself usable there)
    struct Item {
        cudaIpcMemHandle_t ipc;
        void* ptr;
        int dev;
    };
    Item mine;
    std::memset(&mine, 0, sizeof(mine));
    if (!pc.in_process) PCK(cudaIpcGetMemHandle(&mine.ipc, pc.local));
    mine.ptr = pc.local;
    mine.dev = pc.device;
    std::vector<Item> all(pc.P);
    exchange(&mine, all.data(), sizeof(mine));
    for (int r = 0; r < pc.P; ++r) {
        if (r == pc.me) {
            pc.remote[r] = pc.local;
            continue;
        }
        if (pc.in_process) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(all[r].dev, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) throw std::runtime_error(std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
            cudaGetLastError();
            pc.remote[r] = static_cast<char*>(all[r].ptr);
            continue;
        }
        void* p = nullptr;
        PCK(cudaIpcOpenMemHandle(&p, all[r].ipc, cudaIpcMemLazyEnablePeerAccess));
        pc.remote[r] = static_cast<char*>(p);
    }
    pc.ready = true;
}

void peer_close(PeerComm& pc) {
    if (!pc.local) return;
    for (int r = 0; r < pc.P; ++r)
        if (r != pc.me && pc.remote[r] && !pc.in_process) cudaIpcCloseMemHandle(pc.remote[r]);
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
    const size_t base = kind == 0 ? pc.off_x + sizeof(double) * par * pc.nx : pc.off_mc + sizeof(double) * par * pc.nmc;  // kinds 1, 3: multipoles
    for (int r = 0; r < pc.P; ++r) d.p[r] = reinterpret_cast<double*>(pc.remote[r] + base) + own_off;
    const int blocks = (int)std::min<size_t>(296, std::max<size_t>(1, (n / 2 + 255) / 256));
    peer_push_kernel<<<blocks, 256, 0, st>>>(src, n, d, flag_ptrs(pc, kind, par), pc.P, pc.me, epoch, pc.counter + kind);
}

void peer_wait(PeerComm& pc, int kind, int par, unsigned long long epoch, cudaStream_t st) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(pc.local + pc.off_flags) + ((size_t)kind * 2 + par) * kMaxPeers;
    peer_wait_kernel<<<1, 32, 0, st>>>(f, pc.P, epoch, pc.err_dev);
}

// arguments of the next small all-reduce (advances the epoch: call once per exchange, on every rank alike)
void peer_red_args(PeerComm& pc, PeerRed& pr) {
    const unsigned long long epoch = ++pc.epoch_red;
    const int par = (int)(epoch & 1);
    pr.P = pc.P;
    pr.me = pc.me;
    pr.cap = pc.red_cap;
    const PeerFlags f = flag_ptrs(pc, 2, par);
    for (int r = 0; r < pc.P; ++r) {
        pr.red[r] = reinterpret_cast<double*>(pc.remote[r] + pc.off_red) + (size_t)par * pc.P * pc.red_cap;
        pr.flags[r] = f.p[r];
    }
    pr.my_red = reinterpret_cast<const double*>(pc.local + pc.off_red) + (size_t)par * pc.P * pc.red_cap;
    pr.my_flags = reinterpret_cast<const unsigned long long*>(pc.local + pc.off_flags) + ((size_t)2 * 2 + par) * kMaxPeers;
    pr.epoch = epoch;
    pr.err = pc.err_dev;
}

void peer_allreduce(PeerComm& pc, double* vals, int n, cudaStream_t st) {
    if (n > pc.red_cap) throw std::runtime_error("peer_allreduce: too many values");
    PeerRed pr;
    peer_red_args(pc, pr);
    peer_allreduce_kernel<<<1, 256, 0, st>>>(vals, n, pr);
}

}  // namespace an
