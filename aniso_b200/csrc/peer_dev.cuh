// device side of the peer-memory collectives (see peer.cu)
#pragma once
#include "device.cuh"

namespace an {

__device__ __forceinline__ unsigned long long peer_ld_flag(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void peer_st_flag(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// waits until flags[0..P) >= epoch.  Time-out ~20 s at 1.9 GHz: ranks of a solver loop are microseconds apart, but callers
// may let ranks drift by seconds between collective calls (NCCL would simply wait); a peer that never arrives must still
// not hold the GPU for ever
__device__ __forceinline__ bool peer_wait_all(const unsigned long long* flags, int P, unsigned long long epoch, int* err) {
    const long long t0 = clock64();
    for (int r = 0; r < P; ++r) {
        while (peer_ld_flag(flags + r) < epoch) {
            if (clock64() - t0 > 40000000000ll) {
                *err = 1;
                return false;
            }
        }
    }
    return true;
}
// all-reduce (sum) of vals[0..n) by the calling block (all of its threads): push to slot `me` on every rank, wait for all
// slots, add in rank order (the same bits on every rank)
__device__ __forceinline__ void peer_allreduce_block(double* vals, int n, const PeerRed& pr) {
    __shared__ bool peer_ok;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        const double v = vals[k];
        for (int r = 0; r < pr.P; ++r) pr.red[r][(size_t)pr.me * pr.cap + k] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < pr.P) peer_st_flag(pr.flags[threadIdx.x] + pr.me, pr.epoch);
    if (threadIdx.x == 0) peer_ok = peer_wait_all(pr.my_flags, pr.P, pr.epoch, pr.err);
    __syncthreads();
    if (!peer_ok) return;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        double s = 0;
        for (int r = 0; r < pr.P; ++r) s += __ldcg(pr.my_red + (size_t)r * pr.cap + k);
        vals[k] = s;
    }
}

}  // namespace an
