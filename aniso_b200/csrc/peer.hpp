// Peer-memory collectives of the sharded handle (peer.cu): CUDA IPC mappings of one symmetric region per rank.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <functional>
#include <vector>

namespace an {

constexpr int kMaxPeers = 16;

struct PeerPtrs {
    double* p[kMaxPeers];
};
struct PeerFlags {
    unsigned long long* p[kMaxPeers];
};

struct PeerComm {
    int P = 0, me = 0;
    bool ready = false;
    bool in_process = false;  // all ranks live in this process (aniso_create_multi): raw pointers + peer access instead of CUDA IPC
    int device = 0;
    char* local = nullptr;
    char* remote[kMaxPeers] = {};
    size_t bytes = 0, off_flags = 0, off_red = 0, off_x = 0, off_mc = 0, nx = 0, nmc = 0;
    int red_cap = 0;
    unsigned long long epoch_x = 0, epoch_mc = 0, epoch_red = 0;  // advanced alike on every rank (SPMD call sequence)
    unsigned int* counter = nullptr;
    int* err = nullptr;      // pinned host flag: a wait timed out
    int* err_dev = nullptr;
};

void peer_layout(PeerComm& pc, size_t nx, size_t nmc, int red_cap);
void peer_open(PeerComm& pc, const std::function<void(const void*, void*, size_t)>& exchange);
void peer_close(PeerComm& pc);
double* peer_x(const PeerComm& pc, int par);
double* peer_mc(const PeerComm& pc, int par);
void peer_push(PeerComm& pc, int kind, int par, unsigned long long epoch, const double* src, size_t own_off, size_t n, cudaStream_t st);
void peer_wait(PeerComm& pc, int kind, int par, unsigned long long epoch, cudaStream_t st);
void peer_allreduce(PeerComm& pc, double* vals, int n, cudaStream_t st);
struct PeerRed;
void peer_red_args(PeerComm& pc, PeerRed& pr);

}  // namespace an
