// C-ABI of the B200-native aniso hot path (include/aniso_b200.h): owns all device state of one handle and
// sequences the kernels of kernels.cu / vecops.cu.  No CPU fallback anywhere on this path.
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <stdexcept>
#include <thread>
#include <string>
#include <utility>
#include <vector>

#include "../../include/aniso_b200.h"
#include "device.cuh"
#include "nccl_dyn.hpp"
#include "peer.hpp"
#include "plan.hpp"

using namespace an;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
// group handles (aniso_create_multi) --------------------------------------------------------------
// fn(rank index, rank handle) on every rank at once, one thread each; the first failing rank's code and message are returned.
// Like any SPMD program the ranks must fail or succeed together: the arguments are the same on all of them, so argument and
// state errors are; a resource failure on one rank only (out of memory) would leave the others waiting in a collective.
template <class F>
static int multi_all(aniso_handle* h, F&& fn);
#define MULTI_ALL(h, call)                                                                        \
    if ((h) && !(h)->multi.empty()) return multi_all((h), [&](int rank_, aniso_t* r_) -> int {    \
        (void)rank_;                                                                              \
        return call;                                                                              \
    })
#define MULTI_R0(h) \
    if ((h) && !(h)->multi.empty()) h = (h)->multi[0]
#define MULTI_REJECT(h, what)                                                                                         \
    if ((h) && !(h)->multi.empty())                                                                                    \
    return fail(ANISO_EINVAL, what " takes device pointers of one rank: not available on a single-process multi-GPU handle")

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            throw std::runtime_error(std::string(#call) + ": " + cudaGetErrorString(e_));                \
    } while (0)

struct OperatorCache {  // Cache[node][slot] of bbfmm.h:466, flattened
    double* p2p = nullptr;
    double* m2l = nullptr;
    double sign = 1.0;  // (-1)^id
    int id = 0, part = PART_BOTH;  // what the blocks hold (needed when the P2P tiles are evaluated on the fly: p2p == nullptr)
    bool valid = true;             // false after setCoeff: the allocation is kept and refilled in place by the next use
};
struct StencilCache {
    double *c7 = nullptr, *c8 = nullptr, *c9 = nullptr, *call = nullptr;
};

struct aniso_handle {
    HostPlan plan;
    int nranks = 1, rank = 0, device = -1;
    int leaf_begin = 0, leaf_end = 0, slot0 = 0, nown = 0;
    std::vector<int> all_slot_begin, all_slot_end;  // ownership of every rank
    bool dev_ready = false, coeff_set = false;
    cudaStream_t st = nullptr;
    // second stream of the matvec: the two big operator streams (P2P, deepest-level M2L) run here while the
    // latency-bound tree passes run on `st`; forked and joined with events inside every run_fmm
    cudaStream_t st_bulk = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_leaf = nullptr, ev_join = nullptr, ev_own = nullptr, ev_p2p = nullptr;
    bool split_tail = false;  // ANISO_SPLIT_TAIL=1 (one rank, complete tree): second half of the P2P stream behind the deep M2L stream, leaf pass under it
    cudaEvent_t ev_chunk[kDeepChunks] = {};  // chunked tail: the deepest-level M2L rows of chunk c have been streamed
    bool overlap = true;
    // sharded upward pass (nranks > 1, complete uniform tree, equal shares): every rank runs P2M / M2M on its own sub-trees
    // only and the multipoles of all levels are all-gathered (one grouped NCCL call) instead of a replicated upward pass
    // Single-process multi-GPU (aniso_create_multi): this handle is a GROUP that owns one sharded rank handle per device; it
    // has no plan and no device state of its own.  Collective entry points run on all ranks at once, one host thread each
    // (the rank handles talk NCCL to each other exactly as ranks in separate processes do); host-only getters go to rank 0.
    std::vector<aniso_handle*> multi;
    bool in_group = false;  // this is a rank handle of a group: its peers are threads of the same process
    bool p2p_onfly = false;  // aniso_set_option("p2p", "onfly"): near-field tiles evaluated during the apply, not cached
    bool up_sharded = false;
    // CUDA graphs of the operator y = x - mapping_0(sigma_s x), keyed by its two device pointers: the second application
    // with a pointer pair is captured (both streams, the NCCL calls and the fork / join events become one graph), later ones
    // replay it - one launch per matvec instead of ~12 kernel launches, ~8 event calls and 2 NCCL enqueues, which the host
    // cannot issue fast enough for a 0.1 - 0.2 ms sharded matvec.  Dropped whenever the operators change (setCoeff).
    struct FwdGraph {
        int seen = 0;
        cudaGraphExec_t exec = nullptr;
        int64_t kernels = 0;
    };
    std::map<std::pair<const void*, void*>, FwdGraph> fwd_graphs;
    bool use_graphs = true;
    std::vector<cudaEvent_t> copy_ev;  // per-chunk events of the device -> host copy
    // ANISO_PHASE_TIMES=1: events between the phases of every matvec on `st`, averages printed by aniso_destroy (development aid)
    struct PhaseTimer {
        bool on = false;
        std::vector<cudaEvent_t> ev;
        std::vector<const char*> name;
        std::vector<double> sum;
        int cur = 0, reps = 0;
        bool closed = true;
    } pt;
    int m2l_top_groups = 0;  // M2L groups of the first level that has V lists (a prefix: level-major numbering)
    int m2l_deep_g0 = 0;     // M2L groups [m2l_deep_g0, groups) belong to the deepest level and read only leaf multipoles
    bool subtree_ok = false;  // complete uniform tree: the sub-tree passes of subtree.cu replace the level-by-level ones
    SubtreePlan sub;
    std::vector<void*> allocs;
    DevTree T;
    DevGeom G;
    DevCSR m2l, p2p;
    SigmaGrid S;
    std::vector<double> sigma_s, sigma_t, coeff_t, coeff_s;  // host copies (reference order / cell-major)
    double *d_sig_s = nullptr, *d_chat = nullptr, *d_lnorm = nullptr;
    int* d_leaf_of_slot = nullptr;
    int first_m2l_level = 0;
    // work vectors
    double *d_f = nullptr, *d_q = nullptr, *d_mcharge = nullptr, *d_mpot = nullptr, *d_pot = nullptr;
    double *d_fmm = nullptr, *d_corr = nullptr, *d_a = nullptr, *d_b = nullptr, *d_xfull = nullptr, *d_stage = nullptr;
    double *d_blk_full = nullptr, *d_blk_tmp = nullptr;  // block operator work space
    double *d_blk_q = nullptr, *d_blk_f = nullptr, *d_blk_mc = nullptr, *d_blk_sm = nullptr, *d_blk_sp = nullptr;  // batched form
    double *d_blk_wq = nullptr, *d_blk_wf = nullptr, *d_blk_wm = nullptr;  // regrouped form: inputs of the (kernel id, output mode) pairs
    double *d_gm_b = nullptr, *d_gm_x = nullptr, *d_blk_gm = nullptr, *d_blk_io = nullptr;  // device copies of host right-hand sides / iterates
    double* gm_buf = nullptr;   // Krylov work space, kept between solves ((m+3) vectors + CGS2 partials)
    size_t gm_cap = 0;
    double* pinned_small = nullptr;  // Hessenberg column / step report read-back (2m + 8 doubles)
    size_t pinned_small_cap = 0;
    cudaEvent_t gm_ev[2] = {nullptr, nullptr};  // pipelined Arnoldi: report of step i has arrived
    cudaStream_t st_aux = nullptr;              // scalar recurrence of the pipelined Arnoldi (Givens step + report copy)
    cudaEvent_t ev_h = nullptr;                 // projections of the current step are complete
    double* pinned = nullptr;  // 2N doubles
    // sharded host path (nranks > 1): every rank moves 1/nranks of a host vector over PCIe, the rest travels over NVLink
    double* d_stage_pad = nullptr;     // nranks * stage_chunk doubles: in-place all-gather of the uploaded pieces
    int stage_chunk = 0;
    int* d_own_sorted = nullptr;       // owned rows (relative slots) sorted by reference node index
    struct Copy2D { size_t ref_start, stage_off, run, pitch, nruns; };
    std::vector<Copy2D> own_copies;    // strided device -> host copies that put the owned rows at their reference positions
    std::map<std::pair<int, int>, OperatorCache> ops;
    std::map<int, StencilCache> stencils;
    int64_t launches = 0;
    int64_t fill_counts[2] = {0, 0};
    float fill_ms[3] = {0, 0, 0};  // last P2P fill, M2L fill, stencil build (CUDA events)
    DotWork dot;
    double* d_scal = nullptr;  // Hessenberg column / scalars on device
    NcclApi* nccl = nullptr;
    ncclComm_t comm = nullptr;
    PeerComm peer;  // peer-memory collectives of the sharded matvec and of the GMRES reductions (peer.cu); NCCL otherwise
    unsigned char uid[128];
};

namespace {

// ANISO_PHASE_TIMES=1 (development aid): a time line of the matvec.  Every mark records an event on the stream given
// (default: the latency stream `st`); when the next matvec starts, the marks of the previous one are turned into times
// since its start and accumulated.  name == nullptr: start of a matvec (only aniso_forward_dev starts one).
void phase_mark(aniso_handle* h, const char* name, cudaStream_t stream = nullptr) {
    auto& t = h->pt;
    if (!t.on) return;
    if (name && (t.cur == 0 || t.cur >= 24 || t.closed)) return;  // only inside a matvec started by aniso_forward_dev
    if (!name) {
        t.closed = false;
        if (t.cur > 1) {  // fold in the previous matvec (its events have completed by now or are waited for)
            if (t.sum.size() < (size_t)t.cur) t.sum.resize(t.cur, 0.0);
            for (int i = 1; i < t.cur; ++i) {
                cudaEventSynchronize(t.ev[i]);
                float ms = 0;
                cudaEventElapsedTime(&ms, t.ev[0], t.ev[i]);
                t.sum[i] += ms;
            }
            ++t.reps;
        }
        t.cur = 0;
    }
    if ((int)t.ev.size() <= t.cur) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        t.ev.push_back(e);
        t.name.push_back(name ? name : "start");
    }
    cudaEventRecord(t.ev[t.cur++], stream ? stream : h->st);
    if (name && std::string(name) == "down_leaf") t.closed = true;
}

void drop_graphs(aniso_handle* h) {
    for (auto& kv : h->fwd_graphs)
        if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    h->fwd_graphs.clear();
}

template <class T>
T* dalloc(aniso_handle* h, size_t n) {
    T* p = nullptr;
    CK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    h->allocs.push_back(p);
    return p;
}
template <class T>
T* upload(aniso_handle* h, const std::vector<T>& v) {
    T* p = dalloc<T>(h, v.size());
    if (!v.empty()) CK(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, h->st));
    return p;
}
void dfree(aniso_handle* h, void* p) {
    if (!p) return;
    auto it = std::find(h->allocs.begin(), h->allocs.end(), p);
    if (it != h->allocs.end()) h->allocs.erase(it);
    cudaFree(p);
}

DevCSR upload_csr(aniso_handle* h, const BlockCSR& b) {
    DevCSR d;
    {
        std::vector<int> gof(h->plan.nodes.size(), -1);
        for (size_t g = 0; g < b.group_node.size(); ++g) gof[b.group_node[g]] = (int)g;
        d.group_of_node = upload(h, gof);
    }
    d.groups = (int)b.nrow.size();
    d.entries = (int)b.col_off.size();
    d.nrows_total = b.nrows_total;
    d.uniform_nrow = b.uniform_nrow;
    d.uniform_block = b.uniform_block;
    d.nnz = b.nnz;
    d.row_off = upload(h, b.row_off);
    d.nrow = upload(h, b.nrow);
    d.ent_ptr = upload(h, b.ent_ptr);
    d.col_off = upload(h, b.col_off);
    d.ncol = upload(h, b.ncol);
    d.blk_off = upload(h, b.blk_off);
    d.tgt_node = upload(h, b.tgt_node);
    d.src_node = upload(h, b.src_node);
    d.grow_ptr = upload(h, b.grow_ptr);
    d.row_group = upload(h, b.row_group);
    d.symmetric = b.symmetric ? 1 : 0;
    const bool no_swz = getenv("ANISO_NO_SWIZZLE") != nullptr;  // A/B switch
    d.swz = b.symmetric && b.uniform_block == 16 && !no_swz;
    d.scratch_rows = b.scratch_rows;
    if (b.symmetric) {
        d.mirror_row = upload(h, b.mirror_row);
        d.srow_ptr = upload(h, b.srow_ptr);
        d.scratch = dalloc<double>(h, (size_t)b.scratch_rows * b.uniform_block);
        std::vector<int2> rows(h->plan.nodes.size(), make_int2(0, 0));
        for (size_t g = 0; g < b.group_node.size(); ++g) rows[b.group_node[g]] = make_int2(b.srow_ptr[g], b.srow_ptr[g + 1]);
        d.node_rows = upload(h, rows);
    }
    return d;
}

void ensure_device(aniso_handle* h) {
    if (h->dev_ready) return;
    if (h->st || !h->allocs.empty())  // an earlier attempt failed half-way (e.g. in ncclCommInitRank): do not build on its leftovers
        throw std::logic_error("device set-up of this handle failed earlier; destroy it and create a new one");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        throw std::domain_error("no CUDA device available (aniso_b200 has no CPU fallback)");
    }
    if (h->device < 0) CK(cudaGetDevice(&h->device));
    CK(cudaSetDevice(h->device));
    {
        // `st` carries the latency-bound kernels and gets the higher priority: its few CTAs are dispatched as soon as an
        // SM slot frees up, instead of queueing behind the thousands of CTAs of a bulk operator stream on `st_bulk`
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CK(cudaStreamCreateWithPriority(&h->st, cudaStreamNonBlocking, hi));
        CK(cudaStreamCreateWithPriority(&h->st_bulk, cudaStreamNonBlocking, lo));
    }
    for (cudaEvent_t* e : {&h->ev_fork, &h->ev_leaf, &h->ev_join, &h->ev_own, &h->ev_p2p}) CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    for (auto& e : h->ev_chunk) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->overlap = getenv("ANISO_NO_OVERLAP") == nullptr;
    h->pt.on = getenv("ANISO_PHASE_TIMES") != nullptr;
    h->use_graphs = getenv("ANISO_GRAPH") != nullptr && !h->pt.on;
    h->split_tail = getenv("ANISO_SPLIT_TAIL") != nullptr;  // opt-in: measured slower than stream launches
    const HostPlan& p = h->plan;
    const int N = p.N, nn = (int)p.nodes.size(), r = p.rank;
    // geometry in slot order
    std::vector<double> sx(N), sy(N), sw(N);
    for (int s = 0; s < N; ++s) {
        sx[s] = p.nx[p.perm[s]];
        sy[s] = p.ny[p.perm[s]];
        sw[s] = p.w[p.perm[s]];
    }
    DevGeom& G = h->G;
    G.sz = p.sz; G.deg = p.deg; G.c2 = p.c2; G.rq = p.rq; G.nsing = p.nsing; G.N = N; G.dx = p.dx;
    G.px = upload(h, sx); G.py = upload(h, sy); G.pw = upload(h, sw);
    G.perm = upload(h, p.perm); G.inv_perm = upload(h, p.inv_perm);
    {
        std::vector<double> sxn(N), syn(N);
        for (int s = 0; s < N; ++s) {  // (x - center)/radius exactly as bbfmm.h:737-740
            const TreeNode& lf = p.nodes[p.leaf_of_slot[s]];
            sxn[s] = (sx[s] - lf.cx) / lf.rx;
            syn[s] = (sy[s] - lf.cy) / lf.ry;
        }
        G.sxn = upload(h, sxn); G.syn = upload(h, syn);
    }
    G.sqrtw = upload(h, p.sqrtw); G.interp = upload(h, p.interp); G.lnorm = upload(h, p.lnorm);
    G.nearMap = upload(h, p.nearMap); G.rqx = upload(h, p.rqx); G.rqy = upload(h, p.rqy); G.rqw = upload(h, p.rqw);
    G.singX = upload(h, p.singX); G.singY = upload(h, p.singY); G.singW = upload(h, p.singW);
    h->d_lnorm = const_cast<double*>(G.lnorm);
    // tree
    std::vector<double> cx(nn), cy(nn), rx(nn), ry(nn);
    std::vector<int> parent(nn), index(nn), child(4 * nn), pt_off(nn), npts(nn);
    std::vector<unsigned char> flags(nn);
    for (int i = 0; i < nn; ++i) {
        const TreeNode& n = p.nodes[i];
        cx[i] = n.cx; cy[i] = n.cy; rx[i] = n.rx; ry[i] = n.ry;
        parent[i] = n.parent; index[i] = n.index; pt_off[i] = n.pt_off; npts[i] = n.npts;
        for (int c = 0; c < 4; ++c) child[4 * i + c] = n.child[c];
        flags[i] = (n.leaf ? 1 : 0) | (n.empty ? 2 : 0);
    }
    DevTree& T = h->T;
    T.nnodes = nn; T.np = p.np; T.r = r; T.nlevels = (int)p.level_ptr.size() - 1;
    T.cx = upload(h, cx); T.cy = upload(h, cy); T.rx = upload(h, rx); T.ry = upload(h, ry);
    T.parent = upload(h, parent); T.index = upload(h, index); T.child = upload(h, child);
    T.pt_off = upload(h, pt_off); T.npts = upload(h, npts); T.flags = upload(h, flags);
    T.cheb = upload(h, p.cheb); T.tnode = upload(h, p.tnode); T.R = upload(h, p.R);
    {
        std::vector<double> RT(p.R.size());
        for (int c = 0; c < 4; ++c)
            for (int a = 0; a < r; ++a)
                for (int b = 0; b < r; ++b) RT[((size_t)c * r + b) * r + a] = p.R[((size_t)c * r + a) * r + b];
        T.RT = upload(h, RT);
    }
    T.leaves = upload(h, p.leaves); T.nleaves = (int)p.leaves.size();
    {
        std::vector<int4> info(p.leaves.size());
        std::vector<int> child0(nn);
        T.max_leaf = 0;
        for (size_t k = 0; k < p.leaves.size(); ++k) {
            const TreeNode& n = p.nodes[p.leaves[k]];
            info[k] = make_int4(p.leaves[k], n.pt_off, n.npts, 0);
            T.max_leaf = std::max(T.max_leaf, n.npts);
        }
        for (int i = 0; i < nn; ++i) {
            const TreeNode& n = p.nodes[i];
            child0[i] = n.leaf ? -1 : n.child[0];
            if (!n.leaf)
                for (int c = 1; c < 4; ++c)
                    if (n.child[c] != n.child[0] + c) throw std::runtime_error("internal: children not consecutive");
        }
        T.uniform_npts = p.leaves.empty() ? 0 : p.nodes[p.leaves[0]].npts;
        for (size_t k = 0; k < p.leaves.size(); ++k)
            if (p.nodes[p.leaves[k]].npts != T.uniform_npts || p.nodes[p.leaves[k]].pt_off != (int)k * T.uniform_npts) T.uniform_npts = 0;
        T.leaf_info = upload(h, info); T.child0 = upload(h, child0);
        T.level_ptr = upload(h, p.level_ptr);
    }
    h->d_leaf_of_slot = upload(h, p.leaf_of_slot);
    h->first_m2l_level = T.nlevels;
    for (int g : p.m2l.group_node) h->first_m2l_level = std::min(h->first_m2l_level, p.nodes[g].level);
    {
        // complete uniform quadtree?  (every node non-empty, all leaves on the deepest level, 4^L nodes on level L)
        const int deepest = T.nlevels - 1;
        bool ok = getenv("ANISO_NO_SUBTREE") == nullptr && !p.generic && !p.m2l.group_node.empty() && r <= 64 && (int)p.leaves.size() == (1 << (2 * deepest));
        for (int L = 0; ok && L <= deepest; ++L) ok = p.level_ptr[L + 1] - p.level_ptr[L] == (1 << (2 * L));
        for (int i = 0; ok && i < nn; ++i) ok = !p.nodes[i].empty && p.nodes[i].leaf == (p.nodes[i].level == deepest);
        for (size_t k = 0; ok && k < p.leaves.size(); ++k) ok = p.leaves[k] == p.level_ptr[deepest] + (int)k;
        h->subtree_ok = ok;
        if (ok) h->sub = make_subtree_plan(h->first_m2l_level, deepest, r);
        bool equal = h->nranks > 1 && h->overlap && !getenv("ANISO_UP_REPLICATED");
        for (int q = 0; equal && q < h->nranks; ++q)
            equal = h->all_slot_end[q] - h->all_slot_begin[q] == N / h->nranks && h->all_slot_begin[q] == q * (N / h->nranks);
        h->up_sharded = ok && equal && p.m2l.symmetric && N % h->nranks == 0 && (h->nranks & (h->nranks - 1)) == 0 &&
                        (1 << (2 * h->first_m2l_level)) % h->nranks == 0;
    }
    if (h->up_sharded) {
        // rank-major multipole storage (DevTree::pack_*): one all-gather moves the multipoles of every level
        int logP = 0;
        while ((1 << logP) < h->nranks) ++logP;
        T.pack_logP = logP;
        T.pack_L0 = h->first_m2l_level;
        T.pack_tot = ((((1 << (2 * T.nlevels)) - 1) / 3) - (((1 << (2 * T.pack_L0)) - 1) / 3)) >> logP;
        BlockCSR packed = p.m2l;  // sources of the M2L blocks are read at the packed positions
        for (size_t e = 0; e < packed.col_off.size(); ++e) {
            const int s = packed.src_node[e], L = p.nodes[s].level;
            packed.col_off[e] = mc_node(T, L, s - p.level_ptr[L]) * r;
        }
        // symmetric storage: a group's row offset is only used to read the target's OWN multipole (the transposed product;
        // results go to scratch rows), so it follows the multipole numbering too
        for (size_t g = 0; g < packed.group_node.size(); ++g) {
            const int t = packed.group_node[g], L = p.nodes[t].level;
            packed.row_off[g] = mc_node(T, L, t - p.level_ptr[L]) * r;
        }
        h->m2l = upload_csr(h, packed);
    } else
        h->m2l = upload_csr(h, p.m2l);
    h->p2p = upload_csr(h, p.p2p);
    h->m2l_top_groups = 0;
    for (int g : p.m2l.group_node)
        if (p.nodes[g].level == h->first_m2l_level) ++h->m2l_top_groups;
    {
        // suffix of groups at the deepest level whose sources are all leaves (their multipoles come straight from P2M)
        const int deepest = T.nlevels - 1, ng = (int)p.m2l.group_node.size();
        int g0 = ng;
        while (g0 > 0 && p.nodes[p.m2l.group_node[g0 - 1]].level == deepest) --g0;
        bool leaf_sources = true;
        for (int e = g0 < ng ? p.m2l.ent_ptr[g0] : (int)p.m2l.src_node.size(); e < (int)p.m2l.src_node.size(); ++e)
            leaf_sources = leaf_sources && p.nodes[p.m2l.src_node[e]].leaf;
        h->m2l_deep_g0 = (leaf_sources && h->first_m2l_level < deepest) ? g0 : ng;
    }
    // sigma grid
    h->d_chat = dalloc<double>(h, N);
    h->d_sig_s = dalloc<double>(h, N);
    SigmaGrid& S = h->S;
    S.sz = p.sz; S.deg = p.deg; S.c2 = p.c2; S.h = p.dx;  // cell edge k sits at k*dx (KernelFactory.cpp:71-72: k / (1/dx))
    S.chat = h->d_chat; S.glx = upload(h, p.glx); S.glw = upload(h, p.glw);
    if (p.deg == 1 && !getenv("ANISO_CELL_WALK")) {  // line-integral tables of the deg-1 fast path, filled by setCoeff
        S.ty = dalloc<double2>(h, (size_t)N + 1);  // + 1: see magic_floor_inner
        S.tx = dalloc<double2>(h, (size_t)N + 1);
        CK(cudaMemsetAsync(const_cast<double2*>(S.ty), 0, sizeof(double2) * ((size_t)N + 1), h->st));
        CK(cudaMemsetAsync(const_cast<double2*>(S.tx), 0, sizeof(double2) * ((size_t)N + 1), h->st));
        S.dty = dalloc<double2>(h, (size_t)N + 1);
        S.dtx = dalloc<double2>(h, (size_t)N + 1);
        CK(cudaMemsetAsync(const_cast<double2*>(S.dty), 0, sizeof(double2) * ((size_t)N + 1), h->st));
        CK(cudaMemsetAsync(const_cast<double2*>(S.dtx), 0, sizeof(double2) * ((size_t)N + 1), h->st));
    }
    // work space
    h->d_f = dalloc<double>(h, N); h->d_q = dalloc<double>(h, N); h->d_pot = dalloc<double>(h, N);
    h->d_fmm = dalloc<double>(h, N); h->d_corr = dalloc<double>(h, N); h->d_a = dalloc<double>(h, N);
    h->d_b = dalloc<double>(h, N); h->d_xfull = dalloc<double>(h, N); h->d_stage = dalloc<double>(h, N);
    h->d_mcharge = dalloc<double>(h, (size_t)nn * r);
    h->d_mpot = dalloc<double>(h, (size_t)nn * r);
    CK(cudaMemsetAsync(h->d_mcharge, 0, (size_t)nn * r * sizeof(double), h->st));
    CK(cudaMemsetAsync(h->d_mpot, 0, (size_t)nn * r * sizeof(double), h->st));
    CK(cudaMemsetAsync(h->d_pot, 0, (size_t)N * sizeof(double), h->st));
    CK(cudaMallocHost(&h->pinned, (size_t)2 * N * sizeof(double)));
    h->dot.blocks = 1184;  // 148 SMs x 8
    h->dot.partial = dalloc<double>(h, h->dot.blocks);
    h->dot.counter = dalloc<unsigned int>(h, 1);
    CK(cudaMemsetAsync(h->dot.counter, 0, sizeof(unsigned int), h->st));
    h->d_scal = dalloc<double>(h, 1024);
    if (h->nranks > 1) {
        h->nccl = load_nccl();
        if (!h->nccl) throw std::runtime_error("libnccl.so.2 could not be loaded");
        ncclUniqueId id;
        static_assert(sizeof(id) == 128, "ncclUniqueId size");
        std::memcpy(&id, h->uid, 128);
        ncclResult_t rc = h->nccl->CommInitRank(&h->comm, h->nranks, id, h->rank);
        if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclCommInitRank: ") + h->nccl->GetErrorString(rc));
        if (h->up_sharded && h->nranks <= kMaxPeers && !getenv("ANISO_NO_PEER")) {
            // peer-memory collectives: one symmetric region per rank, mapped into every rank through CUDA IPC (the handles
            // travel through an NCCL all-gather); every rank must succeed, else all fall back to NCCL
            PeerComm& pc = h->peer;
            pc.P = h->nranks;
            pc.me = h->rank;
            pc.in_process = h->in_group;
            pc.device = h->device;
            peer_layout(pc, (size_t)N, (size_t)T.pack_tot * r * h->nranks, 1024);
            int ok = 1;
            try {
                peer_open(pc, [&](const void* mine, void* all, size_t bytes) {
                    char* d = dalloc<char>(h, bytes * h->nranks);
                    CK(cudaMemcpyAsync(d + bytes * h->rank, mine, bytes, cudaMemcpyHostToDevice, h->st));
                    ncclResult_t e = h->nccl->AllGather(d + bytes * h->rank, d, bytes, ncclChar, h->comm, h->st);
                    if (e != ncclSuccess) throw std::runtime_error(h->nccl->GetErrorString(e));
                    CK(cudaMemcpyAsync(all, d, bytes * h->nranks, cudaMemcpyDeviceToHost, h->st));
                    CK(cudaStreamSynchronize(h->st));
                    dfree(h, d);
                });
            } catch (const std::exception& e) {
                fprintf(stderr, "[aniso rank %d] peer-memory collectives unavailable (%s): using NCCL\n", h->rank, e.what());
                cudaGetLastError();
                ok = 0;
            }
            int* d_ok = dalloc<int>(h, 1);
            CK(cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, h->st));
            rc = h->nccl->AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, h->comm, h->st);
            if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclAllReduce: ") + h->nccl->GetErrorString(rc));
            CK(cudaMemcpyAsync(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, h->st));
            CK(cudaStreamSynchronize(h->st));
            dfree(h, d_ok);
            if (!ok) {
                pc.ready = false;
                peer_close(pc);
            }
        }
    }
    if (h->nranks > 1) {
        // host path of the sharded handle: padded staging vector for the in-place all-gather of uploaded pieces, and the
        // strided copies that return the owned rows to their places in a reference-order host vector
        h->stage_chunk = (N + h->nranks - 1) / h->nranks;
        h->d_stage_pad = dalloc<double>(h, (size_t)h->stage_chunk * h->nranks);
        std::vector<std::pair<int, int>> own(h->nown);  // (reference index, relative slot)
        for (int k = 0; k < h->nown; ++k) own[k] = {p.perm[h->slot0 + k], k};
        std::sort(own.begin(), own.end());
        std::vector<int> rel(h->nown);
        for (int k = 0; k < h->nown; ++k) rel[k] = own[k].second;
        h->d_own_sorted = upload(h, rel);
        std::vector<std::pair<size_t, size_t>> runs;  // (reference start, length)
        for (int k = 0; k < h->nown;) {
            int e = k + 1;
            while (e < h->nown && own[e].first == own[e - 1].first + 1) ++e;
            runs.push_back({(size_t)own[k].first, (size_t)(e - k)});
            k = e;
        }
        size_t off = 0;
        for (size_t i = 0; i < runs.size();) {  // runs of one length at a constant pitch form one 2-D copy
            size_t j = i + 1;
            const size_t pitch = j < runs.size() ? runs[j].first - runs[i].first : runs[i].second;
            while (j < runs.size() && runs[j].second == runs[i].second && runs[j].first - runs[j - 1].first == pitch) ++j;
            h->own_copies.push_back({runs[i].first, off, runs[i].second, std::max(pitch, runs[i].second), j - i});
            off += runs[i].second * (j - i);
            i = j;
        }
    }
    CK(cudaStreamSynchronize(h->st));
    h->dev_ready = true;
}

}  // namespace

#define API_BEGIN try {
#define API_END                                                    \
    }                                                              \
    catch (const std::domain_error& e) {                           \
        return fail(ANISO_ENODEVICE, e.what());                    \
    }                                                              \
    catch (const std::bad_alloc&) {                                \
        return fail(ANISO_ENOMEM, "out of host memory");           \
    }                                                              \
    catch (const std::invalid_argument& e) {                       \
        return fail(ANISO_EINVAL, e.what());                       \
    }                                                              \
    catch (const std::logic_error& e) {                            \
        return fail(ANISO_ESTATE, e.what());                       \
    }                                                              \
    catch (const std::exception& e) {                              \
        return fail(ANISO_ECUDA, e.what());                        \
    }
#define REQUIRE(cond, msg) \
    if (!(cond)) throw std::invalid_argument(msg)
#define LAUNCH(h, call) \
    do {                \
        call;           \
        ++(h)->launches; \
    } while (0)

extern "C" {

const char* aniso_last_error(void) { return g_err.c_str(); }
const char* aniso_version(void) { return "aniso_b200 0.1 (sm_100a)"; }

int aniso_gl_table(int degree, double* x, double* w) {
    if (!x || !w) return fail(ANISO_EINVAL, "null pointer");
    if (!gl_table(degree, x, w)) return fail(ANISO_EINVAL, "The degree is not implemented.");
    return ANISO_OK;
}

int aniso_nccl_unique_id(void* out128) {
    API_BEGIN
    REQUIRE(out128, "null pointer");
    NcclApi* api = load_nccl();
    if (!api) return fail(ANISO_ENCCL, "libnccl.so.2 could not be loaded");
    ncclUniqueId id;
    ncclResult_t rc = api->GetUniqueId(&id);
    if (rc != ncclSuccess) return fail(ANISO_ENCCL, api->GetErrorString(rc));
    std::memcpy(out128, &id, 128);
    return ANISO_OK;
    API_END
}

int aniso_create_sharded(int sz, int deg, int K, double g, int sr, int np, int maxLevel, int nranks, int rank,
                         const void* uid, int device, aniso_t** out) {
    API_BEGIN
    REQUIRE(out, "null output pointer");
    REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
    REQUIRE(nranks == 1 || uid, "nccl_unique_id required for nranks > 1");
    aniso_handle* h = new aniso_handle();
    try {
        build_plan(h->plan, sz, deg, K, g, sr, np, maxLevel, nranks == 1);
    } catch (const std::exception& e) {
        delete h;
        throw std::invalid_argument(e.what());
    }
    h->nranks = nranks;
    h->rank = rank;
    h->device = device;
    if (uid) std::memcpy(h->uid, uid, 128);
    std::vector<int> lb, le;
    partition_leaves(h->plan, nranks, lb, le);
    h->all_slot_begin.resize(nranks);
    h->all_slot_end.resize(nranks);
    const HostPlan& p = h->plan;
    for (int r = 0; r < nranks; ++r) {
        h->all_slot_begin[r] = lb[r] < (int)p.leaves.size() ? p.nodes[p.leaves[lb[r]]].pt_off : p.N;
        h->all_slot_end[r] = le[r] > lb[r] ? p.nodes[p.leaves[le[r] - 1]].pt_off + p.nodes[p.leaves[le[r] - 1]].npts
                                           : h->all_slot_begin[r];
    }
    h->leaf_begin = lb[rank];
    h->leaf_end = le[rank];
    h->slot0 = h->all_slot_begin[rank];
    h->nown = h->all_slot_end[rank] - h->slot0;
    if (nranks > 1) flatten_lists(h->plan, h->leaf_begin, h->leaf_end);
    *out = h;
    return ANISO_OK;
    API_END
}

int aniso_create(int sz, int deg, int K, double g, int sr, int np, int maxLevel, aniso_t** out) {
    return aniso_create_sharded(sz, deg, K, g, sr, np, maxLevel, 1, 0, nullptr, -1, out);
}

int aniso_destroy(aniso_t* h) {
    if (h && !h->multi.empty()) {  // every rank leaves its communicator at the same time
        multi_all(h, [&](int, aniso_t* r_) -> int { return aniso_destroy(r_); });
        delete h;
        return ANISO_OK;
    }
    if (!h) return ANISO_OK;
    // also after a device set-up that failed half-way (dev_ready still false): whatever exists is released
    if (h->dev_ready || h->st || !h->allocs.empty()) {
        if (h->device >= 0) cudaSetDevice(h->device);
        if (h->st) cudaStreamSynchronize(h->st);
        if (h->dev_ready && h->pt.on && h->pt.reps > 0) {
            phase_mark(h, nullptr);
            fprintf(stderr, "[aniso rank %d] time line over %d matvecs (us after the start; [b] = bulk stream):", h->rank, h->pt.reps);
            for (size_t i = 1; i < h->pt.sum.size(); ++i) fprintf(stderr, " %s %.1f", h->pt.name[i], 1e3 * h->pt.sum[i] / h->pt.reps);
            fprintf(stderr, "\n");
        }
        drop_graphs(h);
        if (h->peer.local) {
            // no rank may unmap or free its region while a peer can still store into it: a barrier first
            if (h->comm && h->nccl && h->d_scal && h->st) {
                h->nccl->AllReduce(h->d_scal, h->d_scal, 1, ncclDouble, ncclSum, h->comm, h->st);
                cudaStreamSynchronize(h->st);
            }
            peer_close(h->peer);
        }
        if (h->comm && h->nccl) h->nccl->CommDestroy(h->comm);
        for (void* p : h->allocs) cudaFree(p);
        if (h->pinned) cudaFreeHost(h->pinned);
        if (h->pinned_small) cudaFreeHost(h->pinned_small);
        if (h->st_bulk) cudaStreamSynchronize(h->st_bulk);
        for (cudaEvent_t e : {h->ev_fork, h->ev_leaf, h->ev_join, h->ev_own, h->ev_p2p})
            if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : h->copy_ev) cudaEventDestroy(e);
        for (cudaEvent_t e : h->gm_ev)
            if (e) cudaEventDestroy(e);
        if (h->ev_h) cudaEventDestroy(h->ev_h);
        if (h->st_aux) cudaStreamDestroy(h->st_aux);
        for (cudaEvent_t e : h->ev_chunk)
            if (e) cudaEventDestroy(e);
        if (h->st_bulk) cudaStreamDestroy(h->st_bulk);
        if (h->st) cudaStreamDestroy(h->st);
        cudaGetLastError();
    }
    delete h;
    return ANISO_OK;
}

int aniso_num_nodes(const aniso_t* h) {
    MULTI_R0(h);
    return h ? h->plan.N : fail(ANISO_EINVAL, "null handle");
}
int aniso_kernel_size(const aniso_t* h) {
    MULTI_R0(h);
    return h ? h->plan.kernelSize : fail(ANISO_EINVAL, "null handle");
}

int aniso_get_nodes(const aniso_t* h, double* x, double* y, double* w) {
    MULTI_R0(h);
    if (!h) return fail(ANISO_EINVAL, "null handle");
    const size_t b = sizeof(double) * h->plan.N;
    if (x) std::memcpy(x, h->plan.nx.data(), b);
    if (y) std::memcpy(y, h->plan.ny.data(), b);
    if (w) std::memcpy(w, h->plan.w.data(), b);
    return ANISO_OK;
}

int aniso_geometry_tables(const aniso_t* h, double* interp, double* nearMap, double* lnorm, double* rx, double* ry,
                          double* rw) {
    MULTI_R0(h);
    if (!h) return fail(ANISO_EINVAL, "null handle");
    const HostPlan& p = h->plan;
    auto cp = [](double* d, const std::vector<double>& v) {
        if (d) std::memcpy(d, v.data(), v.size() * sizeof(double));
    };
    cp(interp, p.interp); cp(nearMap, p.nearMap); cp(lnorm, p.lnorm); cp(rx, p.rqx); cp(ry, p.rqy); cp(rw, p.rqw);
    return ANISO_OK;
}

int aniso_sing_tables(const aniso_t* h, double* sx, double* sy, double* sw) {
    MULTI_R0(h);
    if (!h) return fail(ANISO_EINVAL, "null handle");
    const HostPlan& p = h->plan;
    if (sx) std::memcpy(sx, p.singX.data(), p.singX.size() * sizeof(double));
    if (sy) std::memcpy(sy, p.singY.data(), p.singY.size() * sizeof(double));
    if (sw) std::memcpy(sw, p.singW.data(), p.singW.size() * sizeof(double));
    return ANISO_OK;
}

int aniso_tree_stats(const aniso_t* h, int64_t out[9]) {
    MULTI_R0(h);
    if (!h || !out) return fail(ANISO_EINVAL, "null pointer");
    for (int i = 0; i < 9; ++i) out[i] = h->plan.stats[i];
    return ANISO_OK;
}

int aniso_tree_nodes(const aniso_t* h, int cap, double* geom, int* info) {
    MULTI_R0(h);
    if (!h) return fail(ANISO_EINVAL, "null handle");
    const HostPlan& p = h->plan;
    const int nn = (int)p.nodes.size();
    for (int i = 0; i < nn && i < cap; ++i) {
        const TreeNode& n = p.nodes[i];
        if (geom) { geom[4 * i] = n.cx; geom[4 * i + 1] = n.cy; geom[4 * i + 2] = n.rx; geom[4 * i + 3] = n.ry; }
        if (info) { info[4 * i] = n.level; info[4 * i + 1] = n.leaf; info[4 * i + 2] = n.npts; info[4 * i + 3] = n.parent; }
    }
    return nn;
}

int aniso_node_lists(const aniso_t* h, int node, int cap, int* ids, int* kinds) {
    MULTI_R0(h);
    if (!h || node < 0 || node >= (int)h->plan.nodes.size()) return fail(ANISO_EINVAL, "bad node");
    const TreeNode& n = h->plan.nodes[node];
    int c = 0;
    const std::vector<int>* L[4] = {&n.V, &n.X, &n.U, &n.W};
    for (int kind = 0; kind < 4; ++kind)
        for (int e : *L[kind]) {
            if (c < cap) {
                if (ids) ids[c] = e;
                if (kinds) kinds[c] = kind;
            }
            ++c;
        }
    return c;
}

int aniso_slot_permutation(const aniso_t* h, int* perm) {
    MULTI_R0(h);
    if (!h || !perm) return fail(ANISO_EINVAL, "null pointer");
    std::memcpy(perm, h->plan.perm.data(), sizeof(int) * h->plan.N);
    return ANISO_OK;
}

int aniso_partition(const aniso_t* h, int nranks, int rank, int* slot_begin, int* slot_end, int* nleaves) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h && nranks >= 1 && rank >= 0 && rank < nranks, "bad arguments");
    std::vector<int> lb, le;
    partition_leaves(h->plan, nranks, lb, le);
    const HostPlan& p = h->plan;
    const int b = lb[rank] < (int)p.leaves.size() ? p.nodes[p.leaves[lb[rank]]].pt_off : p.N;
    const int e = le[rank] > lb[rank] ? p.nodes[p.leaves[le[rank] - 1]].pt_off + p.nodes[p.leaves[le[rank] - 1]].npts : b;
    if (slot_begin) *slot_begin = b;
    if (slot_end) *slot_end = e;
    if (nleaves) *nleaves = le[rank] - lb[rank];
    return ANISO_OK;
    API_END
}

int aniso_owned_range(const aniso_t* h, int* slot_begin, int* slot_end) {
    if (h && !h->multi.empty()) {  // one process holds every row
        if (slot_begin) *slot_begin = 0;
        if (slot_end) *slot_end = h->multi[0]->plan.N;
        return ANISO_OK;
    }
    if (!h) return fail(ANISO_EINVAL, "null handle");
    if (slot_begin) *slot_begin = h->slot0;
    if (slot_end) *slot_end = h->slot0 + h->nown;
    return ANISO_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// compute path
// ---------------------------------------------------------------------------------------------------
namespace {

// the same block-CSR restricted to groups [g0, g0+count): used to launch the M2L of one level range
DevCSR sub_range(const DevCSR& m, int g0, int count) {
    DevCSR s = m;
    s.groups = count;
    s.row_off += g0;
    s.nrow += g0;
    s.ent_ptr += g0;
    if (s.srow_ptr) s.srow_ptr += g0;
    // rows per group: r for the M2L operator (uniform_nrow), the block size for a uniform P2P operator (symmetric storage)
    s.nrows_total = count * (m.uniform_nrow ? m.uniform_nrow : m.uniform_block);
    return s;
}

// setCoeff: the sigma-dependent operators become stale, their memory stays.  An inverse-problem loop calls setCoeff every
// iteration; returning 3 GB to the driver and asking for it again cost more than evaluating the operators (measured 21 ms
// from setCoeff to the first mapping result at N = 2^20, of which 10 ms were cudaFree / cudaMalloc).  The real-part
// operators (cos(k theta) / r, KernelFactory.cpp:243-252) and the correction stencils (K7 / K8 / K9 act on the real part
// only) do not depend on sigma and stay valid.
void invalidate_caches(aniso_handle* h) {
    drop_graphs(h);
    for (auto& kv : h->ops)
        if (kv.second.part != PART_REAL) kv.second.valid = false;
}

void free_caches(aniso_handle* h) {
    drop_graphs(h);
    for (auto& kv : h->ops) {
        dfree(h, kv.second.p2p);
        dfree(h, kv.second.m2l);
    }
    h->ops.clear();
    for (auto& kv : h->stencils) {
        dfree(h, kv.second.c7); dfree(h, kv.second.c8); dfree(h, kv.second.c9); dfree(h, kv.second.call);
    }
    h->stencils.clear();
}

// The probes evaluate the kernel family between arbitrary points.  Every point the matvec itself uses (nodes, Chebyshev
// nodes of tree boxes) lies inside the unit square, where the table line integral needs no clamp of its crossing
// indices.  The reference does not guard its cell look-ups at all (getRow/getCol, KernelFactory.cpp:392-401, are a bare
// floor: a point outside the square indexes outside the coefficient grid); here such points are rejected, and a coordinate
// of exactly 1 is counted to the last cell.
void check_probe_points(int n, const double* pts4) {
    for (int i = 0; i < 4 * n; ++i)
        if (!(pts4[i] >= 0.0 && pts4[i] <= 1.0)) throw std::invalid_argument("probe point outside the unit square");
}

void check_id(const aniso_handle* h, int id) {
    if (id < 0 || id >= h->plan.kernelSize) throw std::invalid_argument("kernel id out of range");
}

// runKernelsCache / runKernelsCacheSing (KernelFactory.cpp:323-334, 279-288): evaluate every operator block
OperatorCache& ensure_ops(aniso_handle* h, int id, int part) {
    auto key = std::make_pair(id, part);
    auto it = h->ops.find(key);
    if (it != h->ops.end() && it->second.valid) return it->second;
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede cache/mapping");
    const bool refill = it != h->ops.end();  // stale after setCoeff: same allocation, new values
    OperatorCache c;
    if (refill) c = it->second;
    c.sign = (id & 1) ? -1.0 : 1.0;
    c.id = id;
    c.part = part;
    c.valid = true;
    const bool onfly = h->p2p_onfly && p2p_onfly_ok(h->p2p, h->T.max_leaf);
    if (!refill) {
        if (!onfly) c.p2p = dalloc<double>(h, (size_t)h->p2p.nnz + 2);  // + 16 bytes: the last bulk copy of a stream that starts on an odd
        c.m2l = dalloc<double>(h, (size_t)h->m2l.nnz + 2);  //   double rounds up to the next 16-byte boundary (stream.cu symg ring)
    }
    cudaEvent_t ev[3];
    for (auto& e : ev) CK(cudaEventCreate(&e));
    // the optical depth is the same for every kernel id: ids > 0 of the fused operators are derived from id 0
    const OperatorCache* base = nullptr;
    if (id > 0 && part == PART_BOTH && !getenv("ANISO_NO_SHARED_TAU")) base = &ensure_ops(h, 0, PART_BOTH);
    CK(cudaEventRecord(ev[0], h->st));
    if (onfly) {
        // nothing to fill
    } else if (base && base->p2p) LAUNCH(h, launch_fill_from_base(h->p2p, h->T, h->G, id, 0, base->p2p, c.p2p, h->st));
    else LAUNCH(h, launch_fill(h->p2p, h->T, h->G, h->S, id, part, 0, c.p2p, h->st));
    CK(cudaEventRecord(ev[1], h->st));
    if (base) LAUNCH(h, launch_fill_from_base(h->m2l, h->T, h->G, id, 1, base->m2l, c.m2l, h->st));
    else LAUNCH(h, launch_fill(h->m2l, h->T, h->G, h->S, id, part, 1, c.m2l, h->st));
    CK(cudaEventRecord(ev[2], h->st));
    CK(cudaGetLastError());
    CK(cudaEventSynchronize(ev[2]));
    CK(cudaEventElapsedTime(&h->fill_ms[0], ev[0], ev[1]));
    CK(cudaEventElapsedTime(&h->fill_ms[1], ev[1], ev[2]));
    for (auto& e : ev) cudaEventDestroy(e);
    h->fill_counts[0] = onfly ? 0 : h->p2p.nnz;
    h->fill_counts[1] = h->m2l.nnz;
    return h->ops[key] = c;
}

// refineAddOnCache + singularAddCache (KernelFactory.cpp:550-609, 752-788), plus the K7 table
StencilCache& ensure_stencils(aniso_handle* h, int id) {
    if (h->plan.generic) throw std::logic_error("this handle is a bbfmm::kernel point-set handle (aniso_fmm_create), not an Aniso object");
    auto it = h->stencils.find(id);
    if (it != h->stencils.end()) return it->second;
    const size_t n9 = (size_t)9 * h->G.c2 * h->nown;
    StencilCache s;
    s.c7 = dalloc<double>(h, n9);
    s.c8 = dalloc<double>(h, n9);
    s.call = dalloc<double>(h, n9);
    s.c9 = dalloc<double>(h, (size_t)h->G.c2 * h->nown);
    cudaEvent_t ev[2];
    for (auto& e : ev) CK(cudaEventCreate(&e));
    CK(cudaEventRecord(ev[0], h->st));
    LAUNCH(h, launch_build_stencils(h->G, id, h->slot0, h->nown, s.c7, s.c8, s.c9, h->st));
    LAUNCH(h, launch_combine_stencils(h->G, h->slot0, h->nown, s.c7, s.c8, s.c9, s.call, h->st));
    CK(cudaEventRecord(ev[1], h->st));
    CK(cudaGetLastError());
    CK(cudaEventSynchronize(ev[1]));
    CK(cudaEventElapsedTime(&h->fill_ms[2], ev[0], ev[1]));
    for (auto& e : ev) cudaEventDestroy(e);
    return h->stencils[id] = s;
}

// The source of a sharded operator application when it exists only as the owned shards of the ranks (GMRES vectors): the
// all-gather and the scaling are then part of the two-stream schedule of run_fmm instead of preceding it.
struct ShardSource {
    const double* shard;  // owned rows, slot order
    const double* sig_s;  // full vector (slot order) or nullptr
};

void allgather_slots(aniso_handle* h, const double* shard, double* full, cudaStream_t st);

// all-gather of the multipoles of every level that has M2L lists: rank-major storage (DevTree::pack_*, mc_node) puts the
// nodes a rank owns on all levels back to back, so this is one in-place ncclAllGather of pack_tot * r doubles per rank
void allgather_multipoles(aniso_handle* h) {
    NcclApi* a = h->nccl;
    const size_t cnt = (size_t)h->T.pack_tot * h->T.r;
    ncclResult_t rc = a->AllGather(h->d_mcharge + cnt * h->rank, h->d_mcharge, cnt, ncclDouble, h->comm, h->st);
    if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclAllGather: ") + a->GetErrorString(rc));
}

// kernel::runFast (bbfmm.h:565-592): upPass + downPassFast from the cached operators.
// q: scaled source, full vector in slot order.  out: owned range.
// src (sharded handles, sub-tree path): the source is still spread over the ranks; q / f are produced here (h->d_q, h->d_f).
void run_fmm(aniso_handle* h, const OperatorCache& c, const double* q, double* out, int mode = 0, const double* stencil = nullptr,
             const double* f = nullptr, const double* x_own = nullptr, const ShardSource* src = nullptr) {
    const HostPlan& p = h->plan;
    const DevTree& T = h->T;
    const int L0 = h->first_m2l_level, deepest = T.nlevels - 1;
    const bool sym_m = h->m2l.symmetric, sym_p = h->p2p.symmetric;
    auto count = [&](int L) { return p.level_ptr[L + 1] - p.level_ptr[L]; };
    // levels [L0, Lt] hold at most 256 nodes each (a single CTA is slower than a launch beyond that): one single-CTA launch walks them (top of the tree)
    int Lt = L0 - 1;
    while (Lt + 1 <= deepest && count(Lt + 1) <= 256) ++Lt;
    // Two streams.  `st` carries the latency-bound chain (P2M, M2M, shallow M2L, L2L), `st_bulk` the two big operator
    // streams that depend on little: P2P needs only q, the deepest-level M2L only the leaf multipoles from P2M.
    const int ng = h->m2l.groups, gd = h->overlap ? h->m2l_deep_g0 : ng;
    cudaStream_t sb = h->overlap ? h->st_bulk : h->st;
    double* mch = h->d_mcharge;  // multipoles (the peer-memory exchange uses its own double-buffered arrays)
    // split tail: cached symmetric P2P tiles in at least two groups, one rank, both streams, deep M2L rows on the bulk stream
    const bool split = h->split_tail && !src && h->nranks == 1 && h->overlap && h->subtree_ok && c.p2p && h->p2p.symmetric &&
                       h->p2p.groups >= 2 && h->m2l_deep_g0 < h->m2l.groups && !(p.deep_chunks > 1);
    bool deep_gathered_on_bulk = false;  // peer exchange: the bulk stream itself has waited for the leaf multipoles
    const bool m2l_deep_leaf_only = h->m2l_deep_g0 < h->m2l.groups;  // the deep rows exist and read leaf multipoles only
    if (!sym_m) CK(cudaMemsetAsync(h->d_mpot, 0, (size_t)T.nnodes * T.r * sizeof(double), h->st));
    // The launches of the latency chain are issued to `st` BEFORE the bulk stream's operator streams: a stream kernel fills
    // every SM, and sub-tree CTAs launched behind it would wait for its CTAs to retire (measured: upward pass 88 us instead of
    // 55 at N = 2^20).  Issued first, they take their slots at once and the stream kernel fills the rest of the machine.
    auto p2p_on_bulk = [&] {
        if (c.p2p) LAUNCH(h, launch_bsr_apply(h->p2p, c.p2p, q, h->d_pot, c.sign, false, sb));
        else LAUNCH(h, launch_p2p_onfly(h->p2p, h->G, h->S, c.id, c.part, q, h->d_pot, c.sign, sb));  // cache-free near field
        phase_mark(h, "[b]p2p", sb);
    };
    if (src) {
        // Sharded source (requires up_sharded): the all-gather of the shards starts at once on `st_bulk`; meanwhile the owned
        // part is scaled on `st` and this rank's upward pass runs from it.  The other parts are scaled behind the gather.
        const int N = p.N, s0 = h->slot0, n = h->nown;
        CK(cudaEventRecord(h->ev_fork, h->st));
        CK(cudaStreamWaitEvent(sb, h->ev_fork, 0));
        PeerComm& pc = h->peer;
        const bool peer = pc.ready && (n & 1) == 0 && (s0 & 1) == 0;
        const double* xfull = h->d_xfull;
        unsigned long long ex = 0, em = 0;
        if (peer) {
            // all-gather by push over peer memory: this rank's shard goes straight into the gathered vector of every rank
            ex = ++pc.epoch_x;
            em = ++pc.epoch_mc;
            xfull = peer_x(pc, (int)(ex & 1));
            mch = peer_mc(pc, (int)(em & 1));
            const double* shard = src->shard;
            if ((uintptr_t)shard & 15) {  // 16-byte stores need an aligned source: go through a local copy
                CK(cudaMemcpyAsync(h->d_a + s0, shard, sizeof(double) * n, cudaMemcpyDeviceToDevice, sb));
                shard = h->d_a + s0;
            }
            LAUNCH(h, peer_push(pc, 0, (int)(ex & 1), ex, shard, (size_t)s0, (size_t)n, sb));
        } else
            allgather_slots(h, src->shard, h->d_xfull, sb);
        LAUNCH(h, launch_scale(n, src->shard, src->sig_s ? src->sig_s + s0 : nullptr, h->G.pw + s0, h->G.perm + s0, h->d_f, h->d_q + s0, h->st));
        CK(cudaEventRecord(h->ev_own, h->st));
        phase_mark(h, "scale_own");
        q = h->d_q;
        // leaf level of this rank's part of the rank-major multipole array: its last cnt_leaf nodes
        const size_t tot = (size_t)T.pack_tot * T.r, cnt_leaf = (((size_t)1 << (2 * deepest)) >> T.pack_logP) * T.r;
        const size_t own0 = tot * h->rank, leaf0 = own0 + tot - cnt_leaf;
        const std::function<void()> push_leaf = [&] {
            // the deepest-level M2L rows only read leaf multipoles: they travel as soon as P2M has produced them
            LAUNCH(h, peer_push(pc, 3, (int)(em & 1), em, mch + leaf0, leaf0, cnt_leaf, h->st));
        };
        h->launches += launch_up_tree(T, h->G, h->sub, q, mch, h->st, nullptr, h->rank, h->nranks, peer ? &push_leaf : nullptr);
        phase_mark(h, "up_own");
        if (peer) LAUNCH(h, peer_wait(pc, 0, (int)(ex & 1), ex, sb));
        LAUNCH(h, launch_scale(s0, xfull, src->sig_s, h->G.pw, h->G.perm, h->d_f, h->d_q, sb));
        const int e0 = s0 + n;
        LAUNCH(h, launch_scale(N - e0, xfull + e0, src->sig_s ? src->sig_s + e0 : nullptr, h->G.pw + e0, h->G.perm + e0, h->d_f,
                               h->d_q + e0, sb));
        CK(cudaStreamWaitEvent(sb, h->ev_own, 0));  // P2P reads the whole of q
        phase_mark(h, "[b]x_gathered+scaled", sb);
        p2p_on_bulk();
        if (peer) {
            if (tot > cnt_leaf) {  // levels above the leaves (none when the first level with M2L lists is the leaf level)
                LAUNCH(h, peer_push(pc, 1, (int)(em & 1), em, mch + own0, own0, tot - cnt_leaf, h->st));
                LAUNCH(h, peer_wait(pc, 1, (int)(em & 1), em, h->st));  // the shallow M2L rows on `st` read them
            }
            LAUNCH(h, peer_wait(pc, 3, (int)(em & 1), em, sb));      // leaf level: the deepest M2L rows on `st_bulk`
            deep_gathered_on_bulk = m2l_deep_leaf_only;
            if (!deep_gathered_on_bulk) LAUNCH(h, peer_wait(pc, 3, (int)(em & 1), em, h->st));
        } else
            allgather_multipoles(h);
        CK(cudaEventRecord(h->ev_leaf, h->st));
        phase_mark(h, "gather_multipoles");
    } else if (h->overlap) {
        CK(cudaEventRecord(h->ev_fork, h->st));
        CK(cudaStreamWaitEvent(sb, h->ev_fork, 0));
    }
    if (!h->subtree_ok || !h->overlap) p2p_on_bulk();
    if (h->subtree_ok) {
        // complete uniform tree: each pass is ceil(levels/3) launches (subtree.cu)
        if (src) {
            // done above
        } else if (h->up_sharded) {
            h->launches += launch_up_tree(T, h->G, h->sub, q, h->d_mcharge, h->st, nullptr, h->rank, h->nranks);
            phase_mark(h, "up_own");
            p2p_on_bulk();
            allgather_multipoles(h);
            CK(cudaEventRecord(h->ev_leaf, h->st));
            phase_mark(h, "gather_multipoles");
        } else if (split) {
            // Split tail: P2P rows [0, gp/2) fill the bulk stream while the leaf multipoles are being formed; the other half
            // follows the deepest-level M2L stream, and the latency-bound leaf pass (far field only) runs under it.
            const int gp = h->p2p.groups;
            LAUNCH(h, launch_bsr_apply(sub_range(h->p2p, 0, gp / 2), c.p2p, q, h->d_pot, c.sign, false, sb));
            h->launches += launch_up_tree(T, h->G, h->sub, q, h->d_mcharge, h->st, h->ev_leaf);
            phase_mark(h, "up");
        } else {
            h->launches += launch_up_tree(T, h->G, h->sub, q, h->d_mcharge, h->st, h->overlap ? h->ev_leaf : nullptr);
            phase_mark(h, "up");
            if (h->overlap) p2p_on_bulk();
        }
        // Chunked tail (plan.deep_chunks > 1): the deepest-level M2L rows are streamed chunk by chunk, and the leaf pass of a
        // chunk's sub-trees (L2L + L2P + folds + stencil, latency-bound) starts as soon as that chunk has been streamed, under
        // the HBM-bound stream of the next chunks; only the last chunk's leaf pass is exposed.
        const int nch = (h->overlap && gd < ng && p.deep_chunks > 1 && (ng - gd) % p.deep_chunks == 0) ? p.deep_chunks : 1;
        if (h->overlap) {
            if (!deep_gathered_on_bulk) CK(cudaStreamWaitEvent(sb, h->ev_leaf, 0));
            const int per = (ng - gd) / nch;
            for (int k = 0; k < nch && gd < ng; ++k) {
                LAUNCH(h, launch_bsr_apply(sub_range(h->m2l, gd + k * per, per), c.m2l, mch, h->d_mpot, c.sign, false, sb));
                CK(cudaEventRecord(nch > 1 ? h->ev_chunk[k] : h->ev_join, sb));
            }
            if (gd >= ng) CK(cudaEventRecord(h->ev_join, sb));
            phase_mark(h, "[b]m2l_deep", sb);
            if (split) {
                const int gp = h->p2p.groups;
                LAUNCH(h, launch_bsr_apply(sub_range(h->p2p, gp / 2, gp - gp / 2), c.p2p, q, h->d_pot, c.sign, false, sb));
                CK(cudaEventRecord(h->ev_p2p, sb));
                phase_mark(h, "[b]p2p_2nd", sb);
            }
        }
        if (gd > 0) LAUNCH(h, launch_bsr_apply(sub_range(h->m2l, 0, gd), c.m2l, mch, h->d_mpot, c.sign, false, h->st));
        phase_mark(h, "m2l_shallow");
        h->launches += launch_down_mid_tree(T, h->m2l, sym_m, h->sub, L0, h->d_mpot, h->st);
        phase_mark(h, "down_mid");
        if (nch > 1) {
            const int per_root = 1 << (2 * h->sub.leaf_D);  // leaves per sub-tree of the leaf pass
            const int root0 = h->leaf_begin / per_root, nroots = (h->leaf_end - h->leaf_begin) / per_root / nch;
            for (int k = 0; k < nch; ++k) {
                CK(cudaStreamWaitEvent(h->st, h->ev_chunk[k], 0));
                h->launches += launch_down_leaf_tree(T, h->G, h->m2l, sym_m, h->p2p, sym_p, h->sub, L0, h->d_leaf_of_slot, h->slot0,
                                                     h->nown, h->d_mpot, h->d_pot, stencil, f, x_own, out, mode, h->st, nullptr,
                                                     root0 + k * nroots, nroots);
                if (k == nch - 2) phase_mark(h, "down_leaf_but_last");
            }
        } else if (split) {
            CK(cudaStreamWaitEvent(h->st, h->ev_join, 0));
            phase_mark(h, "join_m2l");
            h->launches += launch_down_leaf_tree(T, h->G, h->m2l, sym_m, h->p2p, sym_p, h->sub, L0, h->d_leaf_of_slot, h->slot0, h->nown,
                                                 h->d_mpot, h->d_pot, nullptr, nullptr, nullptr, h->d_fmm, 0, h->st, nullptr, 0, -1, true);
            phase_mark(h, "far_field");
            CK(cudaStreamWaitEvent(h->st, h->ev_p2p, 0));
            LAUNCH(h, launch_tail(T, h->G, h->p2p, sym_p, h->d_leaf_of_slot, h->slot0, h->nown, h->d_fmm, h->d_pot, stencil, f, x_own, out,
                                  mode, h->st));
        } else {
            if (h->overlap) CK(cudaStreamWaitEvent(h->st, h->ev_join, 0));
            phase_mark(h, "join_bulk");
            h->launches += launch_down_leaf_tree(T, h->G, h->m2l, sym_m, h->p2p, sym_p, h->sub, L0, h->d_leaf_of_slot, h->slot0, h->nown,
                                                 h->d_mpot, h->d_pot, stencil, f, x_own, out, mode, h->st);
        }
        phase_mark(h, "down_leaf");
        CK(cudaGetLastError());
        return;
    }
    LAUNCH(h, launch_p2m(T, h->G, q, h->d_mcharge, h->st));
    if (h->overlap) {
        CK(cudaEventRecord(h->ev_leaf, h->st));
        CK(cudaStreamWaitEvent(sb, h->ev_leaf, 0));
        if (gd < ng) LAUNCH(h, launch_bsr_apply(sub_range(h->m2l, gd, ng - gd), c.m2l, h->d_mcharge, h->d_mpot, c.sign, false, sb));
        CK(cudaEventRecord(h->ev_join, sb));
    }
    for (int L = deepest - 1; L >= L0 && L > Lt; --L) LAUNCH(h, launch_m2m(T, p.level_ptr[L], count(L), h->d_mcharge, h->st));
    if (std::min(Lt, deepest - 1) >= L0) LAUNCH(h, launch_m2m_top(T, std::min(Lt, deepest - 1), L0, h->d_mcharge, h->st));
    if (gd > 0) LAUNCH(h, launch_bsr_apply(sub_range(h->m2l, 0, gd), c.m2l, h->d_mcharge, h->d_mpot, c.sign, false, h->st));
    if (sym_m && h->m2l_top_groups > 0)  // level L0 has no L2L step: finish its sums here, every other level folds them in L2L
        LAUNCH(h, launch_bsr_reduce(sub_range(h->m2l, 0, h->m2l_top_groups), h->d_mpot, h->st));
    bool joined = !h->overlap;
    auto join = [&](int L) {  // the deepest level's M2L sums (and the P2P sums) come from the bulk stream
        if (!joined && L >= deepest) {
            CK(cudaStreamWaitEvent(h->st, h->ev_join, 0));
            joined = true;
        }
    };
    if (Lt >= L0 + 1) {
        join(Lt);
        LAUNCH(h, launch_l2l_top(T, h->m2l, sym_m, L0 + 1, Lt, h->d_mpot, h->st));
    }
    for (int L = std::max(L0 + 1, Lt + 1); L <= deepest; ++L) {
        join(L);
        LAUNCH(h, launch_l2l(T, h->m2l, sym_m, p.level_ptr[L], count(L), h->d_mpot, h->st));
    }
    join(deepest);
    LAUNCH(h, launch_l2p(T, h->G, h->p2p, sym_p, h->d_leaf_of_slot, h->slot0, h->nown, h->d_mpot, h->d_pot, stencil, f, x_own, out,
                         mode, h->st));
    CK(cudaGetLastError());
}

// all-gather of the owned shards into a full slot-order vector; counts differ per rank, so this is a
// grouped broadcast per owner
void allgather_slots(aniso_handle* h, const double* shard, double* full, cudaStream_t st) {
    if (h->nranks == 1) {
        if (shard != full) CK(cudaMemcpyAsync(full, shard, sizeof(double) * h->plan.N, cudaMemcpyDeviceToDevice, st));
        return;
    }
    NcclApi* a = h->nccl;
    {
        bool equal = true;
        const int cnt = h->all_slot_end[0] - h->all_slot_begin[0];
        for (int r = 0; r < h->nranks; ++r)
            equal = equal && h->all_slot_end[r] - h->all_slot_begin[r] == cnt && h->all_slot_begin[r] == r * cnt;
        if (equal) {  // one collective instead of nranks broadcasts
            ncclResult_t rc = a->AllGather(shard, full, cnt, ncclDouble, h->comm, st);
            if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclAllGather: ") + a->GetErrorString(rc));
            return;
        }
    }
    a->GroupStart();
    for (int r = 0; r < h->nranks; ++r) {
        const int b = h->all_slot_begin[r], n = h->all_slot_end[r] - b;
        if (n == 0) continue;
        ncclResult_t rc = a->Broadcast(r == h->rank ? shard : full + b, full + b, n, ncclDouble, r, h->comm, st);
        if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclBroadcast: ") + a->GetErrorString(rc));
    }
    ncclResult_t rc = a->GroupEnd();
    if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclGroupEnd: ") + a->GetErrorString(rc));
}
void allgather_slots(aniso_handle* h, const double* shard, double* full) { allgather_slots(h, shard, full, h->st); }

void allreduce_scalars(aniso_handle* h, double* d, int n) {
    if (h->nranks == 1) return;
    if (h->peer.ready && n <= h->peer.red_cap) {
        peer_allreduce(h->peer, d, n, h->st);
        return;
    }
    ncclResult_t rc = h->nccl->AllReduce(d, d, n, ncclDouble, ncclSum, h->comm, h->st);
    if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclAllReduce: ") + h->nccl->GetErrorString(rc));
}

// mapping in slot order: full source (unscaled) -> owned rows
void mapping_slots(aniso_handle* h, const double* f_full, int id, const double* sig_s, const double* x_own, double* out,
                   int subtract, const double* f_shard = nullptr) {
    OperatorCache& c = ensure_ops(h, id, PART_BOTH);
    StencilCache& s = ensure_stencils(h, id);
    if (f_shard && h->up_sharded) {  // the source is the owned shard of every rank: gather + scale inside the schedule
        const ShardSource src{f_shard, sig_s};
        run_fmm(h, c, nullptr, out, subtract ? 2 : 1, s.call, h->d_f, x_own, &src);
        return;
    }
    if (f_shard) {
        allgather_slots(h, f_shard, h->d_xfull);
        f_full = h->d_xfull;
        phase_mark(h, "allgather");
    }
    const double* f = f_full;
    if (sig_s) {
        LAUNCH(h, launch_scale(h->plan.N, f_full, sig_s, h->G.pw, h->G.perm, h->d_f, h->d_q, h->st));
    } else
        LAUNCH(h, launch_scale(h->plan.N, f_full, nullptr, h->G.pw, h->G.perm, h->d_f, h->d_q, h->st));
    f = h->d_f;  // unscaled source in reference node order
    phase_mark(h, "scale");
    run_fmm(h, c, h->d_q, out, subtract ? 2 : 1, s.call, f, x_own);
}

void forward_own(aniso_handle* h, const double* x_own, double* y_own, bool graph);

// host vector (N doubles) -> h->d_stage through the pinned buffer; the stream is drained first so the pinned
// buffer is never rewritten under an in-flight copy.  Chunked: the CPU copy of chunk c+1 overlaps the DMA of chunk c.
constexpr int COPY_CHUNK = 1 << 18;  // doubles (2 MB)
bool is_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}
void stage_h2d_range(aniso_handle* h, const double* host, int lo, int cnt, double* dst) {  // host[lo, lo+cnt) -> dst[lo, lo+cnt)
    if (cnt <= 0) return;
    if (is_pinned(host)) {  // caller's page-locked memory: one direct DMA
        CK(cudaMemcpyAsync(dst + lo, host + lo, sizeof(double) * cnt, cudaMemcpyHostToDevice, h->st));
        return;
    }
    CK(cudaStreamSynchronize(h->st));
    for (int off = lo; off < lo + cnt; off += COPY_CHUNK) {
        const int n = std::min(COPY_CHUNK, lo + cnt - off);
        std::memcpy(h->pinned + off, host + off, sizeof(double) * n);
        CK(cudaMemcpyAsync(dst + off, h->pinned + off, sizeof(double) * n, cudaMemcpyHostToDevice, h->st));
    }
}
void stage_h2d(aniso_handle* h, const double* host) { stage_h2d_range(h, host, 0, h->plan.N, h->d_stage); }

// A host vector every rank holds (reference order) -> the full vector on every device (reference order): each rank uploads
// one contiguous 1/nranks piece and the pieces are all-gathered over NVLink, instead of nranks full copies through the
// host's PCIe root.  Returns the device vector.  ANISO_HOST_FULL_COPY=1: the plain full upload (A/B switch).
const double* stage_h2d_shared(aniso_handle* h, const double* host) {
    static const bool full_copy = getenv("ANISO_HOST_FULL_COPY") != nullptr;
    if (h->nranks == 1 || full_copy) {
        stage_h2d(h, host);
        return h->d_stage;
    }
    const int N = h->plan.N, lo = std::min(N, h->rank * h->stage_chunk), cnt = std::min(N, lo + h->stage_chunk) - lo;
    stage_h2d_range(h, host, lo, cnt, h->d_stage_pad);
    ncclResult_t rc = h->nccl->AllGather(h->d_stage_pad + (size_t)h->rank * h->stage_chunk, h->d_stage_pad, h->stage_chunk, ncclDouble,
                                         h->comm, h->st);
    if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclAllGather: ") + h->nccl->GetErrorString(rc));
    return h->d_stage_pad;
}

// host (reference order) -> device slot order, full vector
void to_slots(aniso_handle* h, const double* host, double* dev) {
    const int N = h->plan.N;
    const double* ref = stage_h2d_shared(h, host);
    LAUNCH(h, launch_gather(N, ref, h->G.perm, dev, h->st));
}
// device owned rows (slot order) -> host reference order; rows owned by other ranks are left untouched
void from_slots(aniso_handle* h, const double* dev_own, double* host) {
    const HostPlan& p = h->plan;
    double* stage = h->pinned + p.N;
    if (h->nranks == 1) {
        // permute on the device, then bring the vector back in chunks: the CPU copy of chunk c overlaps the DMA of c+1
        const int N = p.N;
        LAUNCH(h, launch_scatter(N, dev_own, h->G.perm, h->d_stage, h->st));
        if (is_pinned(host)) {
            CK(cudaMemcpyAsync(host, h->d_stage, sizeof(double) * N, cudaMemcpyDeviceToHost, h->st));
            CK(cudaStreamSynchronize(h->st));
            return;
        }
        const int nch = (N + COPY_CHUNK - 1) / COPY_CHUNK;
        while ((int)h->copy_ev.size() < nch) {
            cudaEvent_t e;
            CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            h->copy_ev.push_back(e);
        }
        for (int c = 0; c < nch; ++c) {
            const int off = c * COPY_CHUNK, n = std::min(COPY_CHUNK, N - off);
            CK(cudaMemcpyAsync(stage + off, h->d_stage + off, sizeof(double) * n, cudaMemcpyDeviceToHost, h->st));
            CK(cudaEventRecord(h->copy_ev[c], h->st));
        }
        for (int c = 0; c < nch; ++c) {
            const int off = c * COPY_CHUNK, n = std::min(COPY_CHUNK, N - off);
            CK(cudaEventSynchronize(h->copy_ev[c]));
            std::memcpy(host + off, stage + off, sizeof(double) * n);
        }
        return;
    }
    if (h->own_copies.size() <= 64 && !getenv("ANISO_HOST_FULL_COPY")) {
        // owned rows sorted by reference index on the device, then one strided copy per group of equal runs (a quadtree
        // box is sz/2^L runs of one length at pitch sz): no host-side scatter
        LAUNCH(h, launch_gather(h->nown, dev_own, h->d_own_sorted, h->d_stage, h->st));
        for (const auto& c : h->own_copies)
            CK(cudaMemcpy2DAsync(host + c.ref_start, c.pitch * sizeof(double), h->d_stage + c.stage_off, c.run * sizeof(double),
                                 c.run * sizeof(double), c.nruns, cudaMemcpyDeviceToHost, h->st));
        CK(cudaStreamSynchronize(h->st));
        return;
    }
    CK(cudaMemcpyAsync(stage, dev_own, sizeof(double) * h->nown, cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    for (int k = 0; k < h->nown; ++k) host[p.perm[h->slot0 + k]] = stage[k];
}

}  // namespace

extern "C" {

int aniso_set_coeff(aniso_t* h, const double* sigma_s, const double* sigma_t) {
    MULTI_ALL(h, aniso_set_coeff(r_, sigma_s, sigma_t));
    API_BEGIN
    REQUIRE(h && sigma_s && sigma_t, "null pointer");
    if (h->plan.generic) throw std::logic_error("this handle is a bbfmm::kernel point-set handle (aniso_fmm_create), not an Aniso object");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    const HostPlan& p = h->plan;
    const int N = p.N;
    CK(cudaStreamSynchronize(h->st));
    invalidate_caches(h);
    h->sigma_s.assign(sigma_s, sigma_s + N);
    h->sigma_t.assign(sigma_t, sigma_t + N);
    // sigma_s in slot order for the GMRES operator
    to_slots(h, sigma_s, h->d_sig_s);
    CK(cudaStreamSynchronize(h->st));
    // interpolation() KernelFactory.cpp:212-227 on the device; chat = coeff / legendreNorms feeds the ray march
    stage_h2d(h, sigma_t);
    LAUNCH(h, launch_interpolate(h->G, h->d_stage, h->d_lnorm, h->d_chat, h->st));
    if (h->S.ty) LAUNCH(h, launch_sigma_prefix(h->S, const_cast<double2*>(h->S.ty), const_cast<double2*>(h->S.tx), h->st));
    CK(cudaStreamSynchronize(h->st));
    h->coeff_t.clear();
    h->coeff_s.clear();
    h->coeff_set = true;
    return ANISO_OK;
    API_END
}

int aniso_interpolation(aniso_t* h, const double* f, double* coeff) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h && f && coeff, "null pointer");
    if (h->plan.generic) throw std::logic_error("this handle is a bbfmm::kernel point-set handle (aniso_fmm_create), not an Aniso object");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    const int N = h->plan.N;
    stage_h2d(h, f);
    LAUNCH(h, launch_interpolate(h->G, h->d_stage, nullptr, h->d_a, h->st));
    CK(cudaMemcpyAsync(h->pinned + N, h->d_a, sizeof(double) * N, cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    std::memcpy(coeff, h->pinned + N, sizeof(double) * N);
    return ANISO_OK;
    API_END
}

int aniso_sigma_coeff(aniso_t* h, double* ct, double* cs) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h, "null handle");
    if (!h->coeff_set) throw std::logic_error("setCoeff has not been called");
    if (ct) { int rc = aniso_interpolation(h, h->sigma_t.data(), ct); if (rc) return rc; }
    if (cs) { int rc = aniso_interpolation(h, h->sigma_s.data(), cs); if (rc) return rc; }
    return ANISO_OK;
    API_END
}

int aniso_cache(aniso_t* h, int id) {
    MULTI_ALL(h, aniso_cache(r_, id));
    API_BEGIN
    REQUIRE(h, "null handle");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    ensure_ops(h, id, PART_BOTH);
    ensure_stencils(h, id);
    CK(cudaStreamSynchronize(h->st));
    return ANISO_OK;
    API_END
}

int aniso_cache_part(aniso_t* h, int id, int part) {
    MULTI_ALL(h, aniso_cache_part(r_, id, part));
    API_BEGIN
    REQUIRE(h && part >= 0 && part <= 3, "bad arguments");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (part == 0) ensure_stencils(h, id);
    else ensure_ops(h, id, part);
    CK(cudaStreamSynchronize(h->st));
    return ANISO_OK;
    API_END
}

size_t aniso_cache_bytes(const aniso_t* h, int id) {
    if (h && !h->multi.empty()) {
        size_t b = 0;
        for (const aniso_handle* r : h->multi) b += aniso_cache_bytes(r, id);
        return b;
    }
    if (!h) return 0;
    size_t b = 0;
    for (auto& kv : h->ops)
        if (kv.first.first == id) b += (size_t)((kv.second.p2p ? h->p2p.nnz : 0) + h->m2l.nnz) * sizeof(double);
    if (h->stencils.count(id)) b += ((size_t)3 * 9 + 1) * h->G.c2 * h->nown * sizeof(double);
    return b;
}

int aniso_cache_counts(const aniso_t* h, int64_t out[2]) {
    MULTI_R0(h);
    if (!h || !out) return fail(ANISO_EINVAL, "null pointer");
    out[0] = h->fill_counts[0];
    out[1] = h->fill_counts[1];
    return ANISO_OK;
}

int aniso_cache_times(const aniso_t* h, double out[3]) {
    MULTI_R0(h);
    if (!h || !out) return fail(ANISO_EINVAL, "null pointer");
    for (int i = 0; i < 3; ++i) out[i] = h->fill_ms[i];
    return ANISO_OK;
}

int aniso_mapping_dev(aniso_t* h, const double* charge_dev, int id, double* out_dev) {
    MULTI_REJECT(h, "aniso_mapping_dev");
    API_BEGIN
    REQUIRE(h && charge_dev && out_dev, "null pointer");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    mapping_slots(h, charge_dev, id, nullptr, nullptr, out_dev, 0);
    return ANISO_OK;
    API_END
}

int aniso_forward_dev(aniso_t* h, const double* x_dev, double* y_dev) {
    MULTI_REJECT(h, "aniso_forward_dev");
    API_BEGIN
    REQUIRE(h && x_dev && y_dev, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    phase_mark(h, nullptr);
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede the operator");
    forward_own(h, x_dev, y_dev, true);
    return ANISO_OK;
    API_END
}

int aniso_mapping(aniso_t* h, const double* charge, int id, double* out) {
    MULTI_ALL(h, aniso_mapping(r_, charge, id, out));
    API_BEGIN
    REQUIRE(h && charge && out, "null pointer");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    to_slots(h, charge, h->d_a);
    mapping_slots(h, h->d_a, id, nullptr, nullptr, h->d_b, 0);
    from_slots(h, h->d_b, out);
    return ANISO_OK;
    API_END
}

static int stage_fmm(aniso_t* h, int id, int part, const double* f, double* ret) {
    MULTI_ALL(h, stage_fmm(r_, id, part, f, ret));
    API_BEGIN
    REQUIRE(h && f && ret, "null pointer");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    OperatorCache& c = ensure_ops(h, id, part);
    to_slots(h, f, h->d_q);  // the caller passes the scaled function, like the reference
    run_fmm(h, c, h->d_q, h->d_b);
    from_slots(h, h->d_b, ret);
    return ANISO_OK;
    API_END
}
int aniso_run_kernels_fast(aniso_t* h, int id, const double* f, double* ret) { return stage_fmm(h, id, PART_IMAG, f, ret); }
int aniso_run_kernels_fast_sing(aniso_t* h, int id, const double* f, double* ret) { return stage_fmm(h, id, PART_REAL, f, ret); }

static int stage_stencil(aniso_t* h, int id, int which, const double* f, double* ret) {
    MULTI_ALL(h, stage_stencil(r_, id, which, f, ret));
    API_BEGIN
    REQUIRE(h && f && ret, "null pointer");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    StencilCache& s = ensure_stencils(h, id);
    to_slots(h, ret, h->d_b);  // corrections accumulate into ret
    if (which == 9) {
        stage_h2d(h, f);
        LAUNCH(h, launch_sing_apply(h->G, h->slot0, h->nown, s.c9, h->d_stage, h->d_b + h->slot0, h->st));
    } else {
        to_slots(h, f, h->d_a);
        LAUNCH(h, launch_stencil_apply(h->G, h->slot0, h->nown, which == 7 ? s.c7 : s.c8, h->d_a, h->d_b + h->slot0, 1, h->st));
    }
    from_slots(h, h->d_b + h->slot0, ret);
    return ANISO_OK;
    API_END
}
int aniso_near_removal(aniso_t* h, int id, const double* f, double* ret) { return stage_stencil(h, id, 7, f, ret); }
int aniso_refine_add_on_fast(aniso_t* h, int id, const double* f, double* ret) { return stage_stencil(h, id, 8, f, ret); }
int aniso_singular_add_fast(aniso_t* h, int id, const double* fc, double* ret) { return stage_stencil(h, id, 9, fc, ret); }

int aniso_line_integral(aniso_t* h, int n, const double* pts4, double* out) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h && pts4 && out && n >= 0, "bad arguments");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede lineIntegral");
    check_probe_points(n, pts4);
    double* d = dalloc<double>(h, (size_t)5 * n);
    struct Guard {  // the temporary is released on the error paths too
        aniso_handle* h;
        void* p;
        ~Guard() { dfree(h, p); }
    } guard{h, d};
    CK(cudaMemcpyAsync(d, pts4, sizeof(double) * 4 * n, cudaMemcpyHostToDevice, h->st));
    LAUNCH(h, launch_probe_line_integral(h->S, n, d, d + 4 * (size_t)n, h->st));
    CK(cudaMemcpyAsync(out, d + 4 * (size_t)n, sizeof(double) * n, cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return ANISO_OK;
    API_END
}

int aniso_kernel_eval(aniso_t* h, int id, int part, int n, const double* pts4, double* out) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h && pts4 && out && n >= 0 && part >= 1 && part <= 3, "bad arguments");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede kernel evaluation");
    check_probe_points(n, pts4);
    double* d = dalloc<double>(h, (size_t)5 * n);
    struct Guard {
        aniso_handle* h;
        void* p;
        ~Guard() { dfree(h, p); }
    } guard{h, d};
    CK(cudaMemcpyAsync(d, pts4, sizeof(double) * 4 * n, cudaMemcpyHostToDevice, h->st));
    LAUNCH(h, launch_probe_kernel(h->S, id, part, n, d, d + 4 * (size_t)n, h->st));
    CK(cudaMemcpyAsync(out, d + 4 * (size_t)n, sizeof(double) * n, cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return ANISO_OK;
    API_END
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// GMRES (gmres.cpp:53-169) with A = forwardOperator (main.cpp:123-136).  Vectors live on the device in slot order
// (owned range); the Hessenberg column is produced on the device and read back once per Arnoldi step for the
// Givens update, which stays on the host exactly as in the reference.
// ---------------------------------------------------------------------------------------------------
namespace {

void plane_rotation(double dx, double dy, double& cs, double& sn) {  // GeneratePlaneRotation, gmres.cpp:26-39
    if (dy == 0.0) {
        cs = 1.0;
        sn = 0.0;
    } else if (std::fabs(dy) > std::fabs(dx)) {
        const double t = dx / dy;
        sn = 1.0 / std::sqrt(1.0 + t * t);
        cs = t * sn;
    } else {
        const double t = dy / dx;
        cs = 1.0 / std::sqrt(1.0 + t * t);
        sn = t * cs;
    }
}
void apply_rotation(double& dx, double& dy, double cs, double sn) {  // ApplyPlaneRotation, gmres.cpp:42-46
    const double t = cs * dx + sn * dy;
    dy = -sn * dx + cs * dy;
    dx = t;
}

struct GmresWork {
    aniso_handle* h;
    int n, m;
    double *w, *r, *V;  // V: (m+1) x n
    double* v(int i) { return V + (size_t)i * n; }
};

// dot over the owned rows, summed across ranks; result stays on the device (and optionally sqrt'ed)
void ddot_dev(aniso_handle* h, int n, const double* a, const double* b, double* out, bool take_sqrt) {
    LAUNCH(h, launch_dot(n, a, b, out, h->dot, h->st));
    allreduce_scalars(h, out, 1);
    if (take_sqrt) LAUNCH(h, launch_sqrt1(out, h->st));
}
void check_peer(aniso_handle* h) {  // a peer-memory wait that timed out leaves its mark in pinned host memory
    if (h->peer.err && *h->peer.err) throw std::runtime_error("peer-memory exchange timed out (a rank is missing or out of step)");
}

double read_scalar(aniso_handle* h, const double* d) {
    double v;
    CK(cudaMemcpyAsync(&v, d, sizeof(double), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return v;
}

void forward_eager(aniso_handle* h, const double* x_own, double* y_own) {
    if (h->nranks > 1) mapping_slots(h, nullptr, 0, h->d_sig_s, x_own, y_own, 1, x_own);
    else mapping_slots(h, x_own, 0, h->d_sig_s, x_own, y_own, 1);
}

// graph == true: the caller applies the operator to this pointer pair again and again (GMRES on its fixed input buffer,
// aniso_forward_dev from a benchmark or solver loop): first application eager (it also builds the operator caches and makes
// the per-kernel shared-memory opt-ins), second one captured, then replays.
void forward_own(aniso_handle* h, const double* x_own, double* y_own, bool graph) {
    if (!graph || !h->use_graphs || !h->subtree_ok) {
        forward_eager(h, x_own, y_own);
        return;
    }
    if (h->fwd_graphs.size() > 32) drop_graphs(h);  // callers that never repeat a pointer pair
    auto& g = h->fwd_graphs[{x_own, y_own}];
    if (g.exec) {
        CK(cudaGraphLaunch(g.exec, h->st));
        h->launches += g.kernels;
        return;
    }
    if (++g.seen < 2) {
        forward_eager(h, x_own, y_own);
        return;
    }
    const int64_t l0 = h->launches;
    cudaGraph_t graph_obj = nullptr;
    CK(cudaStreamBeginCapture(h->st, cudaStreamCaptureModeThreadLocal));
    try {
        forward_eager(h, x_own, y_own);
    } catch (...) {
        cudaStreamEndCapture(h->st, &graph_obj);
        if (graph_obj) cudaGraphDestroy(graph_obj);
        cudaGetLastError();
        throw;
    }
    CK(cudaStreamEndCapture(h->st, &graph_obj));
    g.kernels = h->launches - l0;
    h->launches = l0;
    const cudaError_t e = cudaGraphInstantiate(&g.exec, graph_obj, 0);
    cudaGraphDestroy(graph_obj);
    if (e != cudaSuccess) {
        g.exec = nullptr;
        throw std::runtime_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
    }
    CK(cudaGraphLaunch(g.exec, h->st));
    h->launches += g.kernels;
}

// Multi-mode block operator of the MATLAB application (aniso.m:121-156): K angular modes of n values,
//   forward (weighted = false):  out_i = sum_{j=-(K-1)}^{K-1} mapping_{|i+j|}(u_|j|)
//   mforward (weighted = true):  out_i = sum_j chi_|j| mapping_{|i+j|}(sigma_s .* u_|j|),  chi_a = (g^a - g^K)/(1 - g^K)
// the only place the Henyey-Greenstein anisotropy g acts.  subtract: out = u - out (the GMRES operator of aniso.m:162).
// u_own / out: K blocks of owned rows in slot order.  Loop and accumulation order are the reference's.
// Batched form (complete uniform tree, symmetric 16x16 / 64x64 blocks): every cached operator is streamed ONCE per block
// apply for all K input modes (launch_bsr_apply_multi), the upward pass runs once per input mode and the downward pass once
// per output mode, folding the 2K-1 (kernel id, input mode) terms with their coefficients (FoldList).  2K-1 operator
// streams instead of K(2K-1).
bool block_multi_ok(const aniso_handle* h) {
    const int K = h->plan.K;
    auto ok = [](const DevCSR& m) { return m.symmetric && (m.uniform_block == 16 || m.uniform_block == 64); };
    return h->subtree_ok && ok(h->m2l) && ok(h->p2p) && 2 * K - 1 <= FoldList::MAXF && !getenv("ANISO_BLOCK_LOOP") &&
           !h->p2p_onfly;  // the batched forms stream cached P2P tiles
}

void block_apply_multi(aniso_handle* h, const double* full, const double* u_own, double* out, bool weighted, bool subtract) {
    const int K = h->plan.K, N = h->plan.N, n = h->nown, nk = 2 * K - 1;
    const DevTree& T = h->T;
    const size_t mc = (size_t)T.nnodes * T.r;
    const size_t Sm = (size_t)h->m2l.scratch_rows * h->m2l.uniform_block, Sp = (size_t)h->p2p.scratch_rows * h->p2p.uniform_block;
    if (!h->d_blk_q) {
        h->d_blk_q = dalloc<double>(h, (size_t)K * N);
        h->d_blk_f = dalloc<double>(h, (size_t)K * N);
        h->d_blk_mc = dalloc<double>(h, (size_t)K * mc);
        h->d_blk_sm = dalloc<double>(h, (size_t)nk * K * Sm);
        h->d_blk_sp = dalloc<double>(h, (size_t)nk * K * Sp);
    }
    for (int a = 0; a < K; ++a) {
        LAUNCH(h, launch_scale(N, full + (size_t)a * N, weighted ? h->d_sig_s : nullptr, h->G.pw, h->G.perm, h->d_blk_f + (size_t)a * N,
                               h->d_blk_q + (size_t)a * N, h->st));
        h->launches += launch_up_tree(T, h->G, h->sub, h->d_blk_q + (size_t)a * N, h->d_blk_mc + (size_t)a * mc, h->st, nullptr);
    }
    std::vector<const double*> stencil(nk);
    const double g = h->plan.g, gK = std::pow(g, K);
    if (h->m2l.swz && h->p2p.swz && K <= 8 && nk <= IdList::MAXI && !getenv("ANISO_NO_DMMA") && !getenv("ANISO_BLOCK_PER_ID")) {
        // Regrouped by kernel id: out_i = sum_k [ FMM_k + C_k ] w^(i,k),  w^(i,k) = sum_{j: |i+j| = k} chi_|j| f_|j|  (at most two
        // input modes).  The multipoles, charges and stencil sources of the w's are linear combinations of those of the K
        // input modes; all kernel ids are then streamed by ONE launch per operator kind that accumulates over k on the fly
        // (launch_bsr_apply_ids), so each output mode has a single scratch set and the downward pass folds one term.
        if (!h->d_blk_wq) {
            h->d_blk_wq = dalloc<double>(h, (size_t)nk * K * N);
            h->d_blk_wf = dalloc<double>(h, (size_t)nk * K * N);
            h->d_blk_wm = dalloc<double>(h, (size_t)nk * K * mc);
            CK(cudaMemsetAsync(h->d_blk_wq, 0, sizeof(double) * nk * K * N, h->st));   // absent (k, i) pairs stay zero
            CK(cudaMemsetAsync(h->d_blk_wf, 0, sizeof(double) * nk * K * N, h->st));
            CK(cudaMemsetAsync(h->d_blk_wm, 0, sizeof(double) * nk * K * mc, h->st));
        }
        struct Term { int a[2]; double c[2]; int n = 0; };
        std::vector<Term> W((size_t)nk * K);
        for (int iid = 0; iid < K; ++iid)
            for (int j = -(K - 1); j <= K - 1; ++j) {  // the reference's loop (aniso.m:126, 143)
                const int a = std::abs(j), k = std::abs(-iid - j);
                const double coef = weighted ? (std::pow(g, a) - gK) / (1.0 - gK) : 1.0;
                Term& t = W[(size_t)k * K + iid];
                if (t.n && t.a[0] == a) t.c[0] += coef;
                else {
                    t.a[t.n] = a;
                    t.c[t.n++] = coef;
                }
            }
        LinList Lq, Lf, Lm;
        for (int k = 0; k < nk; ++k)
            for (int iid = 0; iid < K; ++iid) {
                const Term& t = W[(size_t)k * K + iid];
                if (!t.n) continue;
                const size_t w = (size_t)k * K + iid;
                auto add = [&](LinList& L, double* dst, const double* base, size_t stride) {
                    const int e = L.n++;
                    L.dst[e] = dst;
                    L.a[e] = base + t.a[0] * stride;
                    L.ca[e] = t.c[0];
                    L.b[e] = t.n > 1 ? base + t.a[1] * stride : nullptr;
                    L.cb[e] = t.n > 1 ? t.c[1] : 0.0;
                };
                add(Lq, h->d_blk_wq + w * N, h->d_blk_q, (size_t)N);
                add(Lf, h->d_blk_wf + w * N, h->d_blk_f, (size_t)N);
                add(Lm, h->d_blk_wm + w * mc, h->d_blk_mc, mc);
            }
        LAUNCH(h, launch_lincomb(Lq, (size_t)N, h->st));
        LAUNCH(h, launch_lincomb(Lf, (size_t)N, h->st));
        LAUNCH(h, launch_lincomb(Lm, mc, h->st));
        IdList im, ip;
        for (int k = 0; k < nk; ++k) {
            OperatorCache& c = ensure_ops(h, k, PART_BOTH);
            stencil[k] = ensure_stencils(h, k).call;
            im.A[k] = c.m2l; im.x[k] = h->d_blk_wm + (size_t)k * K * mc; im.sign[k] = c.sign;
            ip.A[k] = c.p2p; ip.x[k] = h->d_blk_wq + (size_t)k * K * N; ip.sign[k] = c.sign;
        }
        im.n = ip.n = nk;
        LAUNCH(h, launch_bsr_apply_ids(h->m2l, im, mc, K, h->d_blk_sm, Sm, h->st));
        LAUNCH(h, launch_bsr_apply_ids(h->p2p, ip, (size_t)N, K, h->d_blk_sp, Sp, h->st));
        for (int iid = 0; iid < K; ++iid) {
            FoldList fl;
            fl.nscr = 1;
            for (int k = 0; k < nk; ++k) {
                if (!W[(size_t)k * K + iid].n) continue;
                const int t = fl.n++;
                fl.coef[t] = 1.0;
                fl.m2l_scr[t] = h->d_blk_sm + (size_t)iid * Sm;
                fl.p2p_scr[t] = h->d_blk_sp + (size_t)iid * Sp;
                fl.C[t] = stencil[k];
                fl.f[t] = h->d_blk_wf + ((size_t)k * K + iid) * N;
            }
            h->launches += launch_down_mid_tree(T, h->m2l, true, h->sub, h->first_m2l_level, h->d_mpot, h->st, &fl);
            h->launches += launch_down_leaf_tree(T, h->G, h->m2l, true, h->p2p, true, h->sub, h->first_m2l_level, h->d_leaf_of_slot,
                                                 h->slot0, n, h->d_mpot, h->d_pot, nullptr, nullptr, u_own + (size_t)iid * n,
                                                 out + (size_t)iid * n, subtract ? 2 : 1, h->st, &fl);
        }
        CK(cudaGetLastError());
        return;
    }
    for (int k = 0; k < nk; ++k) {
        OperatorCache& c = ensure_ops(h, k, PART_BOTH);
        stencil[k] = ensure_stencils(h, k).call;
        h->launches += launch_bsr_apply_multi(h->m2l, c.m2l, h->d_blk_mc, mc, K, h->d_blk_sm + (size_t)k * K * Sm, Sm, c.sign, h->st);
        h->launches += launch_bsr_apply_multi(h->p2p, c.p2p, h->d_blk_q, (size_t)N, K, h->d_blk_sp + (size_t)k * K * Sp, Sp, c.sign, h->st);
    }
    for (int iid = 0; iid < K; ++iid) {
        FoldList fl;
        for (int j = -(K - 1); j <= K - 1; ++j) {  // the reference's loop order (aniso.m:126, 143)
            const int a = std::abs(j), k = std::abs(-iid - j), t = fl.n++;
            fl.coef[t] = weighted ? (std::pow(g, a) - gK) / (1.0 - gK) : 1.0;
            fl.m2l_scr[t] = h->d_blk_sm + ((size_t)k * K + a) * Sm;
            fl.p2p_scr[t] = h->d_blk_sp + ((size_t)k * K + a) * Sp;
            fl.C[t] = stencil[k];
            fl.f[t] = h->d_blk_f + (size_t)a * N;
        }
        h->launches += launch_down_mid_tree(T, h->m2l, true, h->sub, h->first_m2l_level, h->d_mpot, h->st, &fl);
        h->launches += launch_down_leaf_tree(T, h->G, h->m2l, true, h->p2p, true, h->sub, h->first_m2l_level, h->d_leaf_of_slot, h->slot0,
                                             n, h->d_mpot, h->d_pot, nullptr, nullptr, u_own + (size_t)iid * n, out + (size_t)iid * n,
                                             subtract ? 2 : 1, h->st, &fl);
    }
    CK(cudaGetLastError());
}

void block_apply(aniso_handle* h, const double* u_own, double* out, bool weighted, bool subtract) {
    const int K = h->plan.K, N = h->plan.N, n = h->nown;
    if (!h->d_blk_tmp) h->d_blk_tmp = dalloc<double>(h, n);
    const double* full = u_own;
    if (h->nranks > 1) {
        if (!h->d_blk_full) h->d_blk_full = dalloc<double>(h, (size_t)K * N);
        for (int a = 0; a < K; ++a) allgather_slots(h, u_own + (size_t)a * n, h->d_blk_full + (size_t)a * N);
        full = h->d_blk_full;
    }
    if (block_multi_ok(h)) {
        block_apply_multi(h, full, u_own, out, weighted, subtract);
        return;
    }
    const double g = h->plan.g, gK = std::pow(g, K);
    for (int i = -(K - 1); i <= 0; ++i) {
        const int iid = -i;
        double* acc = out + (size_t)iid * n;
        CK(cudaMemsetAsync(acc, 0, sizeof(double) * n, h->st));
        for (int j = -(K - 1); j <= K - 1; ++j) {
            const int id = std::abs(j), k = std::abs(i - j);
            mapping_slots(h, full + (size_t)id * N, k, weighted ? h->d_sig_s : nullptr, nullptr, h->d_blk_tmp, 0);
            const double chi = weighted ? (std::pow(g, id) - gK) / (1.0 - gK) : 1.0;
            LAUNCH(h, launch_axpy(n, chi, h->d_blk_tmp, acc, h->st));
        }
        if (subtract) LAUNCH(h, launch_sub(n, u_own + (size_t)iid * n, acc, acc, h->st));
    }
}

// Update, gmres.cpp:12-24
void gm_update(GmresWork& g, double* x, int k, const std::vector<double>& H, int ldh, const std::vector<double>& s) {
    std::vector<double> y(s.begin(), s.begin() + k + 1);
    for (int i = k; i >= 0; --i) {
        y[i] /= H[i + (size_t)i * ldh];
        for (int j = i - 1; j >= 0; --j) y[j] -= H[j + (size_t)i * ldh] * y[i];
    }
    for (int j = 0; j <= k; ++j) LAUNCH(g.h, launch_axpy(g.n, y[j], g.v(j), x, g.h->st));
}

// owned rows (slot order) of every rank -> the complete vector in reference order on this rank's host
void slots_to_host_full(aniso_handle* h, const double* dev_own, double* host) {
    if (h->nranks == 1) {
        from_slots(h, dev_own, host);
        return;
    }
    allgather_slots(h, dev_own, h->d_xfull);
    LAUNCH(h, launch_scatter(h->plan.N, h->d_xfull, h->G.perm, h->d_stage, h->st));
    CK(cudaMemcpyAsync(host, h->d_stage, sizeof(double) * h->plan.N, cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
}

// Host preconditioner callback (aniso.m:111-119 `prec`, handed to MATLAB's gmres at aniso.m:165) as a device operator:
// the vector goes to the host in reference node order (all `blocks` modes, complete on every rank), the callback maps
// r -> z = M^-1 r, and z comes back.  On a sharded handle every rank runs the callback on the same data.
struct HostPrec {
    aniso_handle* h;
    aniso_prec_fn fn;
    void* ctx;
    int blocks;
    std::vector<double> r, z;
    void operator()(double* v_own) {
        const int N = h->plan.N, n = h->nown;
        r.resize((size_t)blocks * N);
        z.assign((size_t)blocks * N, 0.0);
        for (int a = 0; a < blocks; ++a) slots_to_host_full(h, v_own + (size_t)a * n, r.data() + (size_t)a * N);
        const int rc = fn(ctx, blocks * N, r.data(), z.data());
        if (rc != 0) throw std::runtime_error("preconditioner callback returned " + std::to_string(rc));
        for (int a = 0; a < blocks; ++a) {
            to_slots(h, z.data() + (size_t)a * N, h->d_a);
            CK(cudaMemcpyAsync(v_own + (size_t)a * n, h->d_a + h->slot0, sizeof(double) * n, cudaMemcpyDeviceToDevice, h->st));
        }
        CK(cudaStreamSynchronize(h->st));
    }
};

// A: y = A x on device vectors of length n (owned rows; for the block operator K blocks of owned rows)
// P (optional): left preconditioner applied in place to a device vector, z = M^-1 r (MATLAB gmres(A, b, restart, tol, maxit, M1)
// as called by aniso.m:165: the iteration runs on M^-1 A x = M^-1 b and the residual is measured in that norm)
using DevOp = std::function<void(const double*, double*)>;
using DevPrec = std::function<void(double*)>;

// Afix (optional): the same operator, applied by the pipelined Arnoldi to ONE fixed pair of buffers (every new basis vector is
// also written to the input buffer), so that it can be a captured graph
int gmres_core(aniso_handle* h, const DevOp& A, int n, const double* b, double* x, int m, int max_iter, double tol, int* iters,
               double* hist, int hist_cap, const DevPrec* P = nullptr, const DevOp* Afix = nullptr) {
    if (m < 1 || m > 4096) throw std::invalid_argument("restart length m out of range (1 .. 4096)");  // before any allocation
    const int ldh = m + 1;
    GmresWork g{h, n, m, nullptr, nullptr, nullptr};
    // device scalars of a solve: projections h1 / h2 (m + 2 each), then the state of the device-side Givens recurrence
    const size_t nscal = (size_t)4 * (m + 2) + (size_t)ldh * m + (size_t)4 * (m + 1) + (size_t)2 * (m + 1) + 8;
    {
        // one allocation, kept in the handle between solves (680 MB at C5, m = 80: cudaMalloc/cudaFree of that size cost
        // milliseconds, a visible share of a 25 ms solve)
        const size_t need = (size_t)(m + 5) * n + (size_t)(m + 1) * h->dot.blocks + nscal;
        if (h->gm_cap < need) {
            dfree(h, h->gm_buf);
            h->gm_buf = nullptr;
            h->gm_cap = 0;
            h->gm_buf = dalloc<double>(h, need);
            h->gm_cap = need;
        }
        const size_t pin = (size_t)2 * m + 8;
        if (h->pinned_small_cap < pin) {
            if (h->pinned_small) CK(cudaFreeHost(h->pinned_small));
            h->pinned_small = nullptr;
            h->pinned_small_cap = 0;
            CK(cudaMallocHost(&h->pinned_small, pin * sizeof(double)));
            h->pinned_small_cap = pin;
        }
        for (auto& e : h->gm_ev)
            if (!e) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        if (!h->ev_h) CK(cudaEventCreateWithFlags(&h->ev_h, cudaEventDisableTiming));
        if (!h->st_aux) CK(cudaStreamCreateWithFlags(&h->st_aux, cudaStreamNonBlocking));
    }
    g.w = h->gm_buf;
    g.r = g.w + n;
    double* pb = g.r + n;  // M^-1 b of a preconditioned solve
    double* gx = pb + n;   // fixed input buffer of the operator graph (pipelined form)
    g.V = gx + n;
    double* md_partial = g.V + (size_t)(m + 1) * n;
    double* dh = md_partial + (size_t)(m + 1) * h->dot.blocks;  // h1: m + 2 entries
    double* h2 = dh + m + 2;
    double* dh_b = h2 + m + 2;               // second set of projections: the pipelined form alternates by step parity, so that
    double* h2_b = dh_b + m + 2;             //   the scalar recurrence of step i (on st_aux) can still read while step i + 1 writes
    double* dH = h2_b + m + 2;               // (m + 1) x m, column-major: the rotated Hessenberg matrix (device Givens)
    double* dcs = dH + (size_t)ldh * m;
    double* dsn = dcs + m + 1;
    double* ds = dsn + m + 1;
    double* dy = ds + m + 1;
    double* drep = dy + m + 1;               // 2 per step: residual, converged flag
    double* dbeta = drep + 2 * (m + 1);      // beta, |b|, x0 . x0
    // orthogonalisation: classical Gram-Schmidt applied twice in three fused passes over the basis (default; same
    // projections to rounding and the same iteration counts as the reference's modified Gram-Schmidt on every fixture, one
    // launch and one all-reduce per pass instead of i+1 dependent ones).  ANISO_GMRES_ORTHO=mgs: the reference's MGS
    // (gmres.cpp:115-119); cgs2_unfused: the two-launch form
    const char* oenv = getenv("ANISO_GMRES_ORTHO");
    const bool cgs2 = oenv ? std::string(oenv) == "cgs2" || std::string(oenv) == "cgs2_unfused" || std::string(oenv) == "cgs2_3red" : true;
    const bool cgs_fused = !(oenv && std::string(oenv) == "cgs2_unfused");
    const bool cgs_3red = oenv && std::string(oenv) == "cgs2_3red";  // the fused form with a third reduction for the norm
    double* col = h->pinned_small;
    std::vector<double> H((size_t)(m + 1) * m, 0.0), s(m + 1, 0.0), cs(m + 1, 0.0), sn(m + 1, 0.0);
    int rc = 1, j = 1, i = 0;
    bool done = false;
    double resid = 0;
    *iters = 0;
    // ANISO_GMRES_TIMES=1 (development aid): device time of the operator and of the orthogonalisation, summed over the
    // Arnoldi steps, printed next to the wall time of the solve
    const bool gt_on = getenv("ANISO_GMRES_TIMES") != nullptr;
    cudaEvent_t gt_ev[3] = {nullptr, nullptr, nullptr};
    double gt_ms[2] = {0, 0};
    const auto gt_t0 = std::chrono::steady_clock::now();
    if (gt_on)
        for (auto& e : gt_ev) CK(cudaEventCreate(&e));
    // Pipelined Arnoldi (default): Hessenberg column, Givens rotations and the stopping test stay on the device
    // (givens_step_kernel); the host queues step i+1 and only then waits for the 16-byte report of step i, so the device
    // never idles between steps.  When step i reports convergence exactly one surplus step has been queued - on every rank
    // alike, since all ranks see the same all-reduced scalars - and its results are discarded.  The synchronous form (read the
    // column back, rotate on the host like gmres.cpp:123-144) remains for MGS, the two-launch CGS2, a host preconditioner,
    // the timers and ANISO_GMRES_SYNC=1.
    const bool pipelined = cgs2 && cgs_fused && !cgs_3red && !P && !gt_on && !getenv("ANISO_GMRES_SYNC") && cgs_pass_ok(n, 1, g.V, (size_t)n, g.w);
    const double* bb = b;  // right-hand side of the iteration: b or M^-1 b
    if (P) {
        CK(cudaMemcpyAsync(pb, b, sizeof(double) * n, cudaMemcpyDeviceToDevice, h->st));
        (*P)(pb);
        bb = pb;
    }
    // |b| and x0 . x0 come back in one read.  x0 = 0 (the usual start): A x0 = 0 exactly (the operator is linear and every
    // kernel of it maps zeros to zeros), so r = b and the first operator application (gmres.cpp:86) is not launched.
    ddot_dev(h, n, bb, bb, dbeta + 1, true);
    ddot_dev(h, n, x, x, dbeta + 2, false);
    CK(cudaMemcpyAsync(col, dbeta + 1, 2 * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    double normb = col[0], beta;
    if (col[1] == 0.0) {
        CK(cudaMemcpyAsync(g.r, bb, sizeof(double) * n, cudaMemcpyDeviceToDevice, h->st));
        CK(cudaMemcpyAsync(dbeta, dbeta + 1, sizeof(double), cudaMemcpyDeviceToDevice, h->st));
        beta = normb;
    } else {
        A(x, g.w);
        LAUNCH(h, launch_sub(n, b, g.w, g.r, h->st));
        if (P) (*P)(g.r);
        ddot_dev(h, n, g.r, g.r, dbeta, true);
        beta = read_scalar(h, dbeta);
    }
    if (normb == 0.0) normb = 1;
    if ((resid = beta / normb) <= tol) {  // gmres.cpp:90 (<= here, < below: Q11)
        rc = 0;
        done = true;
    }
    // the orthogonalisation of Arnoldi step i (w against v_0 .. v_i) with CGS2; leaves h1 in dh, h2 in h2, w . w in dh[i + 1]
    auto orthogonalise_cgs2 = [&](int i) {
        if (cgs_fused && !cgs_3red && cgs_pass_ok(n, i + 1, g.V, (size_t)n, g.w)) {
            // three fused passes, TWO reductions: h1; w -= V h1 with h2 AND |w|^2 of the new w in the same pass; w -= V h2,
            // whose norm needs no pass of its own: |w - V h2|^2 = |w|^2 - |h2|^2 for the orthonormal basis and h2 = V^T w
            // (h2 is the rounding-level correction of the second Gram-Schmidt sweep, so the identity holds to O(eps)).
            // One all-reduce fewer per Arnoldi step on a sharded handle.
            LAUNCH(h, launch_cgs_pass(0, n, i + 1, nullptr, g.V, (size_t)n, g.w, dh, md_partial, h->dot, h->st));
            allreduce_scalars(h, dh, i + 1);
            LAUNCH(h, launch_cgs_pass(3, n, i + 1, dh, g.V, (size_t)n, g.w, h2, md_partial, h->dot, h->st));
            allreduce_scalars(h, h2, i + 2);
            LAUNCH(h, launch_cgs_pass(4, n, i + 1, h2, g.V, (size_t)n, g.w, dh + i + 1, md_partial, h->dot, h->st));
            return;
        }
        if (cgs_fused && cgs_pass_ok(n, i + 1, g.V, (size_t)n, g.w)) {
            // three fused passes over the basis (vecops.cu cgs_pass_kernel): h1; w -= V h1 and h2; w -= V h2 and |w|^2
            LAUNCH(h, launch_cgs_pass(0, n, i + 1, nullptr, g.V, (size_t)n, g.w, dh, md_partial, h->dot, h->st));
            allreduce_scalars(h, dh, i + 1);
            LAUNCH(h, launch_cgs_pass(1, n, i + 1, dh, g.V, (size_t)n, g.w, h2, md_partial, h->dot, h->st));
            allreduce_scalars(h, h2, i + 1);
            LAUNCH(h, launch_cgs_pass(2, n, i + 1, h2, g.V, (size_t)n, g.w, dh + i + 1, md_partial, h->dot, h->st));
        } else {
            for (int pass = 0; pass < 2; ++pass) {
                double* hp = pass ? h2 : dh;
                LAUNCH(h, launch_multi_dot(n, i + 1, g.w, g.V, (size_t)n, hp, md_partial, h->dot, h->st));
                ++h->launches;
                allreduce_scalars(h, hp, i + 1);
                LAUNCH(h, launch_multi_axpy(n, i + 1, hp, g.V, (size_t)n, g.w, pass ? dh + i + 1 : nullptr, md_partial, h->dot, h->st));
            }
            ++h->launches;
        }
        allreduce_scalars(h, dh + i + 1, 1);
    };
    // Everything of Arnoldi step i that the pipelined form queues without looking at results: the operator, then CGS2 in
    // three passes over the basis with two reductions - h1; w' = w - V h1 with h2 = V^T w' and |w'|^2; v_{i+1} = (w' - V h2) /
    // sqrt(|w'|^2 - |h2|^2) written straight to the basis (and to the operator's input buffer).  On a sharded handle the two
    // reductions are completed across the ranks inside the pass kernels (peer memory).  The scalar recurrence (Hessenberg
    // column, rotations, residual) only needs h1 and h2: it runs on st_aux beside the last pass and the next operator.
    auto queue_step = [&](int i) {
        double* h1s = (i & 1) ? dh_b : dh;
        double* h2s = (i & 1) ? h2_b : h2;
        if (Afix) (*Afix)(gx, g.w);  // gx holds v_i
        else A(g.v(i), g.w);
        const bool fused_red = h->peer.ready && i + 2 <= h->peer.red_cap;
        PeerRed pr;
        if (fused_red) peer_red_args(h->peer, pr);
        LAUNCH(h, launch_cgs_pass(0, n, i + 1, nullptr, g.V, (size_t)n, g.w, h1s, md_partial, h->dot, h->st, fused_red ? &pr : nullptr));
        if (!fused_red) allreduce_scalars(h, h1s, i + 1);
        if (fused_red) peer_red_args(h->peer, pr);
        LAUNCH(h, launch_cgs_pass(3, n, i + 1, h1s, g.V, (size_t)n, g.w, h2s, md_partial, h->dot, h->st, fused_red ? &pr : nullptr));
        if (!fused_red) allreduce_scalars(h, h2s, i + 2);
        CK(cudaEventRecord(h->ev_h, h->st));
        LAUNCH(h, launch_cgs_pass(4, n, i + 1, h2s, g.V, (size_t)n, g.w, h1s + i + 1, md_partial, h->dot, h->st, nullptr, g.v(i + 1),
                                  Afix ? gx : nullptr));
        CK(cudaStreamWaitEvent(h->st_aux, h->ev_h, 0));
        LAUNCH(h, launch_givens_step(i, h1s, h2s, nullptr, dbeta, dH, ldh, dcs, dsn, ds, normb, tol, drep + 2 * i, h->st_aux));
        CK(cudaMemcpyAsync(col + 2 * (i & 1), drep + 2 * i, 2 * sizeof(double), cudaMemcpyDeviceToHost, h->st_aux));
        CK(cudaEventRecord(h->gm_ev[i & 1], h->st_aux));
    };
    while (!done && j <= max_iter) {
        LAUNCH(h, launch_scale_inv_dev(n, dbeta, g.r, g.v(0), 0, h->st, pipelined && Afix ? gx : nullptr));
        std::fill(s.begin(), s.end(), 0.0);
        s[0] = beta;
        if (pipelined) {
            int queued = 0;  // highest Arnoldi step of this cycle that has been queued
            queue_step(0);
            for (i = 0; i < m && j <= max_iter; ++i, ++j) {
                if (hist && j - 1 < hist_cap) hist[j - 1] = resid;
                *iters = j;
                if (queued < i) queue_step(queued = i);
                // One step ahead of the report being waited for - unless the last known residual is within a factor 100 of the
                // tolerance: GMRES gains well under a digit per step here, so above that step i does not converge and the
                // queued step is never surplus, while the last two or three steps run one at a time (a surplus operator
                // application costs more than the idle gaps of a whole solve).  The rule reads all-reduced values only, so
                // every rank queues the same steps.
                if (i + 1 < m && j + 1 <= max_iter && resid > 100.0 * tol) queue_step(queued = i + 1);
                CK(cudaEventSynchronize(h->gm_ev[i & 1]));
                resid = col[2 * (i & 1)];
                if (col[2 * (i & 1) + 1] != 0.0) {
                    LAUNCH(h, launch_gm_solve(i, dH, ldh, ds, dy, h->st));
                    LAUNCH(h, launch_multi_axpy(n, i + 1, dy, g.V, (size_t)n, x, nullptr, md_partial, h->dot, h->st));
                    rc = 0;
                    done = true;
                    break;
                }
            }
            if (done) break;
            LAUNCH(h, launch_gm_solve(i - 1, dH, ldh, ds, dy, h->st));
            LAUNCH(h, launch_multi_axpy(n, i, dy, g.V, (size_t)n, x, nullptr, md_partial, h->dot, h->st));
        } else {
        for (i = 0; i < m && j <= max_iter; ++i, ++j) {
            if (hist && j - 1 < hist_cap) hist[j - 1] = resid;
            *iters = j;
            if (gt_on) CK(cudaEventRecord(gt_ev[0], h->st));
            A(g.v(i), g.w);
            if (P) (*P)(g.w);
            if (gt_on) CK(cudaEventRecord(gt_ev[1], h->st));
            if (!cgs2) {
                // modified Gram-Schmidt, gmres.cpp:115-119: h_k = w . v_k; w -= h_k v_k.  The axpy of step k and the dot of
                // step k+1 (or the final norm) share one pass over w.
                ddot_dev(h, n, g.w, g.v(0), dh, false);
                for (int k = 0; k <= i; ++k) {
                    LAUNCH(h, launch_axpy_dot(n, dh + k, g.v(k), g.w, k < i ? g.v(k + 1) : nullptr, dh + k + 1, h->dot, h->st));
                    allreduce_scalars(h, dh + k + 1, 1);
                }
            } else {
                // classical Gram-Schmidt applied twice: the same projections to rounding (H(:,i) = h1 + h2), but all dots of
                // a pass in one launch and ONE all-reduce, instead of i+1 dependent all-reduces
                orthogonalise_cgs2(i);
            }
            // dh[i+1] = w . w: the square root is taken inside the scaling kernel and, for H, on the host
            LAUNCH(h, launch_scale_inv_dev(n, dh + i + 1, g.w, g.v(i + 1), 1, h->st));
            CK(cudaMemcpyAsync(col, dh, sizeof(double) * (cgs2 ? 2 * m + 4 : i + 2), cudaMemcpyDeviceToHost, h->st));
            if (gt_on) CK(cudaEventRecord(gt_ev[2], h->st));
            CK(cudaStreamSynchronize(h->st));
            if (gt_on) {
                float a = 0, o = 0;
                cudaEventElapsedTime(&a, gt_ev[0], gt_ev[1]);
                cudaEventElapsedTime(&o, gt_ev[1], gt_ev[2]);
                gt_ms[0] += a;
                gt_ms[1] += o;
            }
            for (int k = 0; k <= i; ++k) H[k + (size_t)i * ldh] = col[k];
            H[i + 1 + (size_t)i * ldh] = std::sqrt(col[i + 1]);
            if (cgs2)
                for (int k = 0; k <= i; ++k) H[k + (size_t)i * ldh] += col[m + 2 + k];
            for (int k = 0; k < i; ++k) apply_rotation(H[k + (size_t)i * ldh], H[k + 1 + (size_t)i * ldh], cs[k], sn[k]);
            plane_rotation(H[i + (size_t)i * ldh], H[i + 1 + (size_t)i * ldh], cs[i], sn[i]);
            apply_rotation(H[i + (size_t)i * ldh], H[i + 1 + (size_t)i * ldh], cs[i], sn[i]);
            apply_rotation(s[i], s[i + 1], cs[i], sn[i]);
            if ((resid = std::fabs(s[i + 1]) / normb) < tol) {
                gm_update(g, x, i, H, ldh, s);
                rc = 0;
                done = true;
                break;
            }
        }
        if (done) break;
        gm_update(g, x, i - 1, H, ldh, s);
        }
        A(x, g.w);
        LAUNCH(h, launch_sub(n, b, g.w, g.r, h->st));
        if (P) (*P)(g.r);
        ddot_dev(h, n, g.r, g.r, dbeta, true);
        beta = read_scalar(h, dbeta);
        if ((resid = beta / normb) < tol) {
            rc = 0;
            done = true;
        }
    }
    if (hist && *iters < hist_cap) hist[*iters] = resid;
    CK(cudaStreamSynchronize(h->st));
    // a step queued ahead of a converging one leaves its scalar recurrence and its report copy on the side stream: nothing of
    // this solve may still write to the pinned report buffer when the next solve starts reading it
    if (h->st_aux) CK(cudaStreamSynchronize(h->st_aux));
    check_peer(h);
    if (gt_on) {
        const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - gt_t0).count();
        fprintf(stderr, "[aniso gmres] %d iterations: wall %.3f ms, operator %.3f ms, orthogonalisation %.3f ms (device, Arnoldi steps only)\n",
                *iters, wall, gt_ms[0], gt_ms[1]);
        for (auto& e : gt_ev) cudaEventDestroy(e);
    }
    return rc;
}

}  // namespace

extern "C" {

int aniso_gmres_dev(aniso_t* h, const double* b, double* x, int m, int max_iter, double tol, int* iters, double* hist,
                    int hist_cap) {
    MULTI_REJECT(h, "aniso_gmres_dev");
    API_BEGIN
    REQUIRE(h && b && x && iters && m >= 1 && max_iter >= 0, "bad arguments");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede GMRES");
    ensure_ops(h, 0, PART_BOTH);
    ensure_stencils(h, 0);
    const DevOp Afix = [h](const double* xx, double* yy) { forward_own(h, xx, yy, true); };
    return gmres_core(h, [h](const double* xx, double* yy) { forward_own(h, xx, yy, false); }, h->nown, b, x, m, max_iter, tol, iters,
                      hist, hist_cap, nullptr, &Afix);
    API_END
}

int aniso_gmres_prec(aniso_t* h, const double* b, double* x, int m, int max_iter, double tol, aniso_prec_fn prec, void* ctx,
                     int* iters, double* hist, int hist_cap) {
    if (h && !h->multi.empty()) {
        if (prec) return fail(ANISO_EINVAL, "a preconditioner callback on a single-process multi-GPU handle would be called from every rank's thread at once: not supported");
        std::vector<int> its(h->multi.size(), 0);
        std::vector<std::vector<double>> hs(h->multi.size(), std::vector<double>(hist ? (size_t)std::max(hist_cap, 0) : 0));
        const int rc = multi_all(h, [&](int k, aniso_t* r_) -> int {
            return aniso_gmres_prec(r_, b, x, m, max_iter, tol, nullptr, nullptr, &its[k], hist ? hs[k].data() : nullptr, hist_cap);
        });
        if (iters) *iters = its[0];
        if (hist && hist_cap > 0) std::memcpy(hist, hs[0].data(), sizeof(double) * hist_cap);
        return rc;
    }
    API_BEGIN
    REQUIRE(h && b && x && iters && m >= 1 && max_iter >= 0, "bad arguments");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede GMRES");
    ensure_ops(h, 0, PART_BOTH);
    ensure_stencils(h, 0);
    if (!h->d_gm_b) {  // right-hand side and iterate of the host-vector entry points, kept in the handle
        h->d_gm_b = dalloc<double>(h, h->plan.N);
        h->d_gm_x = dalloc<double>(h, h->plan.N);
    }
    double *db = h->d_gm_b, *dx = h->d_gm_x;
    to_slots(h, b, db);
    to_slots(h, x, dx);
    HostPrec hp{h, prec, ctx, 1, {}, {}};
    const DevPrec P = [&hp](double* v) { hp(v); };
    const DevOp Afix = [h](const double* xx, double* yy) { forward_own(h, xx, yy, true); };
    const int rc = gmres_core(h, [h](const double* xx, double* yy) { forward_own(h, xx, yy, false); }, h->nown, db + h->slot0,
                              dx + h->slot0, m, max_iter, tol, iters, hist, hist_cap, prec ? &P : nullptr, &Afix);
    from_slots(h, dx + h->slot0, x);
    return rc;
    API_END
}

int aniso_gmres(aniso_t* h, const double* b, double* x, int m, int max_iter, double tol, int* iters, double* hist,
                int hist_cap) {
    return aniso_gmres_prec(h, b, x, m, max_iter, tol, nullptr, nullptr, iters, hist, hist_cap);
}

int aniso_block_forward_dev(aniso_t* h, const double* u_dev, int weighted, double* out_dev) {
    MULTI_REJECT(h, "aniso_block_forward_dev");
    API_BEGIN
    REQUIRE(h && u_dev && out_dev, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede the block operator");
    block_apply(h, u_dev, out_dev, weighted != 0, false);
    return ANISO_OK;
    API_END
}

int aniso_block_forward(aniso_t* h, const double* u, int weighted, double* out) {
    MULTI_ALL(h, aniso_block_forward(r_, u, weighted, out));
    API_BEGIN
    REQUIRE(h && u && out, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede the block operator");
    const int K = h->plan.K, N = h->plan.N, n = h->nown;
    if (!h->d_blk_io) h->d_blk_io = dalloc<double>(h, (size_t)K * N + 2 * (size_t)K * n);  // kept in the handle: no leak on a throw
    double* du = h->d_blk_io;
    double* dow = du + (size_t)K * N;
    double* down = dow + (size_t)K * n;
    for (int a = 0; a < K; ++a) {
        to_slots(h, u + (size_t)a * N, du + (size_t)a * N);
        CK(cudaMemcpyAsync(dow + (size_t)a * n, du + (size_t)a * N + h->slot0, sizeof(double) * n, cudaMemcpyDeviceToDevice, h->st));
    }
    block_apply(h, dow, down, weighted != 0, false);
    for (int a = 0; a < K; ++a) from_slots(h, down + (size_t)a * n, out + (size_t)a * N);
    return ANISO_OK;
    API_END
}

static int block_gmres_dev(aniso_t* h, const double* rhs_dev, double* x_dev, int m, int max_iter, double tol, aniso_prec_fn prec,
                           void* ctx, int* iters, double* hist, int hist_cap) {
    API_BEGIN
    REQUIRE(h && rhs_dev && x_dev && iters && m >= 1 && max_iter >= 0, "bad arguments");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede GMRES");
    for (int k = 0; k < h->plan.kernelSize; ++k) {
        ensure_ops(h, k, PART_BOTH);
        ensure_stencils(h, k);
    }
    HostPrec hp{h, prec, ctx, h->plan.K, {}, {}};
    const DevPrec P = [&hp](double* v) { hp(v); };
    return gmres_core(h, [h](const double* xx, double* yy) { block_apply(h, xx, yy, true, true); }, h->plan.K * h->nown, rhs_dev,
                      x_dev, std::min(m, std::max(max_iter, 1)), max_iter, tol, iters, hist, hist_cap, prec ? &P : nullptr);
    API_END
}

int aniso_block_gmres_dev(aniso_t* h, const double* rhs_dev, double* x_dev, int m, int max_iter, double tol, int* iters,
                          double* hist, int hist_cap) {
    MULTI_REJECT(h, "aniso_block_gmres_dev");
    return block_gmres_dev(h, rhs_dev, x_dev, m, max_iter, tol, nullptr, nullptr, iters, hist, hist_cap);
}

int aniso_block_gmres_prec(aniso_t* h, const double* rhs, double* x, int m, int max_iter, double tol, aniso_prec_fn prec, void* ctx,
                           int* iters, double* hist, int hist_cap) {
    if (h && !h->multi.empty()) {
        if (prec) return fail(ANISO_EINVAL, "a preconditioner callback on a single-process multi-GPU handle would be called from every rank's thread at once: not supported");
        std::vector<int> its(h->multi.size(), 0);
        std::vector<std::vector<double>> hs(h->multi.size(), std::vector<double>(hist ? (size_t)std::max(hist_cap, 0) : 0));
        const int rc = multi_all(h, [&](int k, aniso_t* r_) -> int {
            return aniso_block_gmres_prec(r_, rhs, x, m, max_iter, tol, nullptr, nullptr, &its[k], hist ? hs[k].data() : nullptr, hist_cap);
        });
        if (iters) *iters = its[0];
        if (hist && hist_cap > 0) std::memcpy(hist, hs[0].data(), sizeof(double) * hist_cap);
        return rc;
    }
    API_BEGIN
    REQUIRE(h && rhs && x && iters && m >= 1 && max_iter >= 0, "bad arguments");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede GMRES");
    const int K = h->plan.K, N = h->plan.N, n = h->nown;
    if (!h->d_blk_gm) h->d_blk_gm = dalloc<double>(h, (size_t)N + 2 * (size_t)K * n);  // staging vector, rhs and iterate
    double* tmp = h->d_blk_gm;
    double* db = tmp + N;
    double* dx = db + (size_t)K * n;
    for (int a = 0; a < K; ++a) {
        to_slots(h, rhs + (size_t)a * N, tmp);
        CK(cudaMemcpyAsync(db + (size_t)a * n, tmp + h->slot0, sizeof(double) * n, cudaMemcpyDeviceToDevice, h->st));
        to_slots(h, x + (size_t)a * N, tmp);
        CK(cudaMemcpyAsync(dx + (size_t)a * n, tmp + h->slot0, sizeof(double) * n, cudaMemcpyDeviceToDevice, h->st));
    }
    const int rc = block_gmres_dev(h, db, dx, m, max_iter, tol, prec, ctx, iters, hist, hist_cap);
    if (rc >= 0)
        for (int a = 0; a < K; ++a) from_slots(h, dx + (size_t)a * n, x + (size_t)a * N);
    return rc;
    API_END
}

int aniso_block_gmres(aniso_t* h, const double* rhs, double* x, int m, int max_iter, double tol, int* iters, double* hist,
                      int hist_cap) {
    return aniso_block_gmres_prec(h, rhs, x, m, max_iter, tol, nullptr, nullptr, iters, hist, hist_cap);
}

void* aniso_stream(aniso_t* h) {
    if (h && !h->multi.empty()) return nullptr;  // a group has one stream per rank
    return h ? (void*)h->st : nullptr;
}
int aniso_device(const aniso_t* h) {
    MULTI_R0(h);
    return h ? h->device : fail(ANISO_EINVAL, "null handle");
}
int aniso_synchronize(aniso_t* h) {
    MULTI_ALL(h, aniso_synchronize(r_));
    API_BEGIN
    REQUIRE(h, "null handle");
    if (h->dev_ready) CK(cudaStreamSynchronize(h->st));
    if (h->peer.err && *h->peer.err) throw std::runtime_error("peer-memory exchange timed out (a rank is missing or out of step)");
    return ANISO_OK;
    API_END
}
int aniso_dev_alloc(aniso_t* h, size_t bytes, void** out) {
    MULTI_REJECT(h, "aniso_dev_alloc");
    API_BEGIN
    REQUIRE(h && out, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    *out = dalloc<char>(h, bytes);
    return ANISO_OK;
    API_END
}
int aniso_dev_free(aniso_t* h, void* p) {
    MULTI_REJECT(h, "aniso_dev_free");
    API_BEGIN
    REQUIRE(h, "null handle");
    CK(cudaSetDevice(h->device));
    dfree(h, p);
    return ANISO_OK;
    API_END
}
int aniso_host_alloc(aniso_t* h, size_t bytes, void** out) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h && out, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    CK(cudaMallocHost(out, std::max<size_t>(bytes, 1)));
    return ANISO_OK;
    API_END
}
int aniso_host_free(aniso_t* h, void* p) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h, "null handle");
    if (p) CK(cudaFreeHost(p));
    return ANISO_OK;
    API_END
}
int aniso_to_slots(aniso_t* h, const double* host, double* dev) {
    MULTI_REJECT(h, "aniso_to_slots");
    API_BEGIN
    REQUIRE(h && host && dev, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    to_slots(h, host, dev);
    return ANISO_OK;
    API_END
}
int aniso_from_slots(aniso_t* h, const double* dev, double* host) {
    MULTI_REJECT(h, "aniso_from_slots");
    API_BEGIN
    REQUIRE(h && host && dev, "null pointer");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    from_slots(h, dev, host);
    return ANISO_OK;
    API_END
}
int64_t aniso_launch_count(const aniso_t* h) {
    if (h && !h->multi.empty()) {
        int64_t n = 0;
        for (const aniso_handle* r : h->multi) n += r->launches;
        return n;
    }
    return h ? h->launches : 0;
}
int aniso_apply_bytes(const aniso_t* h, int64_t out[3]) {
    MULTI_R0(h);
    if (!h || !out) return fail(ANISO_EINVAL, "null pointer");
    // bytes the streaming kernels must move per matvec: stored operator blocks (half of the pairs under symmetric
    // storage) plus one scratch row (8R bytes) written per target and per dropped block
    out[0] = h->plan.p2p.nnz * 8 + (int64_t)h->plan.p2p.scratch_rows * h->plan.p2p.uniform_block * 8;
    out[1] = h->plan.m2l.nnz * 8 + (int64_t)h->plan.m2l.scratch_rows * h->plan.m2l.uniform_block * 8;
    out[2] = (int64_t)9 * h->plan.c2 * h->nown * 8;
    return ANISO_OK;
}
int aniso_time_forward(aniso_t* h, int reps, double* ms) {
    MULTI_REJECT(h, "aniso_time_forward");
    API_BEGIN
    REQUIRE(h && ms && reps >= 1, "bad arguments");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    if (!h->coeff_set) throw std::logic_error("setCoeff must precede the operator");
    ensure_ops(h, 0, PART_BOTH);
    ensure_stencils(h, 0);
    CK(cudaMemsetAsync(h->d_a, 0, sizeof(double) * h->plan.N, h->st));
    forward_own(h, h->d_a + h->slot0, h->d_b, true);  // warm-up
    forward_own(h, h->d_a + h->slot0, h->d_b, true);  // (captured)
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, h->st));
    for (int i = 0; i < reps; ++i) forward_own(h, h->d_a + h->slot0, h->d_b, true);
    CK(cudaEventRecord(e1, h->st));
    CK(cudaEventSynchronize(e1));
    float t = 0;
    CK(cudaEventElapsedTime(&t, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ms = t / reps;
    return ANISO_OK;
    API_END
}

int aniso_set_option(aniso_t* h, const char* name, const char* value) {
    MULTI_ALL(h, aniso_set_option(r_, name, value));
    API_BEGIN
    REQUIRE(h && name && value, "null pointer");
    const std::string n(name), v(value);
    if (n == "p2p") {
        // "cached" (default): runKernelsCache stores every U/W tile (bbfmm.h:949-1039); "onfly": the tiles are evaluated while
        // they are applied, like the reference's uncached kernel::run (bbfmm.h:507-532, 863-946)
        REQUIRE(v == "cached" || v == "onfly", "option p2p takes \"cached\" or \"onfly\"");
        const bool want = v == "onfly";
        if (want != h->p2p_onfly) {
            if (h->dev_ready) {
                CK(cudaSetDevice(h->device));
                CK(cudaStreamSynchronize(h->st));
                free_caches(h);  // the operators are rebuilt in the new form by the next cache / mapping call
            }
            h->p2p_onfly = want;
        }
        return ANISO_OK;
    }
    throw std::invalid_argument("unknown option \"" + n + "\" (known: p2p)");
    API_END
}

const char* aniso_get_config(const aniso_t* h) {
    MULTI_R0(h);
    // the environment switches that select a code path (development A/B switches; all default off), as one string a
    // benchmark line can carry so that a run describes itself
    static thread_local std::string cfg;
    static const char* names[] = {"ANISO_GMRES_ORTHO", "ANISO_GMRES_SYNC", "ANISO_NO_SYM", "ANISO_NO_SUBTREE", "ANISO_NO_OVERLAP",
                                  "ANISO_CELL_WALK", "ANISO_NO_SWIZZLE", "ANISO_STREAM_REGS", "ANISO_SYMG_NO_DMMA", "ANISO_NO_DMMA",
                                  "ANISO_BLOCK_LOOP", "ANISO_BLOCK_PER_ID", "ANISO_NO_SHARED_TAU", "ANISO_SUBTREE_D",
                                  "ANISO_HOST_FULL_COPY", "ANISO_UP_REPLICATED", "ANISO_GRAPH", "ANISO_DEEP_CHUNKS", "ANISO_NO_PEER", "ANISO_SPLIT_TAIL", "ANISO_FROM_BASE_GENERIC"};
    cfg = "aniso_b200 0.2 (sm_100a)";
    if (h) {
        cfg += "; ranks=" + std::to_string(h->nranks) + " sym=" + std::to_string(h->plan.m2l.symmetric ? 1 : 0) +
               " subtree=" + std::to_string(h->subtree_ok ? 1 : 0) + " overlap=" + std::to_string(h->overlap ? 1 : 0) +
               " table_line_integral=" + std::to_string(h->S.ty ? 1 : 0) + " up_sharded=" + std::to_string(h->up_sharded ? 1 : 0) +
               " peer_collectives=" + std::to_string(h->peer.ready ? 1 : 0) + " p2p=" + (h->p2p_onfly ? "onfly" : "cached");
    }
    const char* o = getenv("ANISO_GMRES_ORTHO");
    cfg += std::string("; gmres=") + (o ? o : "cgs2") + (getenv("ANISO_GMRES_SYNC") ? "/sync" : "/pipelined");
    for (const char* n : names)
        if (const char* v = getenv(n)) cfg += std::string("; ") + n + "=" + v;
    return cfg.c_str();
}

int aniso_time_kernel(aniso_t* h, int id, int which, int reps, double* ms) {
    MULTI_REJECT(h, "aniso_time_kernel");
    API_BEGIN
    REQUIRE(h && ms && reps >= 1 && (which == 0 || which == 1), "bad arguments");
    check_id(h, id);
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    OperatorCache& c = ensure_ops(h, id, PART_BOTH);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, h->st));
    for (int i = 0; i < reps; ++i) {
        if (which == 0) LAUNCH(h, launch_bsr_apply(h->m2l, c.m2l, h->d_mcharge, h->d_mpot, c.sign, false, h->st));
        else if (c.p2p) LAUNCH(h, launch_bsr_apply(h->p2p, c.p2p, h->d_q, h->d_pot, c.sign, false, h->st));
        else LAUNCH(h, launch_p2p_onfly(h->p2p, h->G, h->S, c.id, c.part, h->d_q, h->d_pot, c.sign, h->st));
    }
    CK(cudaEventRecord(e1, h->st));
    CK(cudaEventSynchronize(e1));
    float t = 0;
    CK(cudaEventElapsedTime(&t, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ms = t / reps;
    return ANISO_OK;
    API_END
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// Allocation-free probe (no counterpart in the reference).
// aniso_line_integral stages its points and results in a device buffer it cudaMallocs and frees on every call.  Callers that
// must not allocate (tight serving loops, harnesses that forbid cudaMalloc) get the same kernel on a fixed scratch array
// that lives in the library image: ANISO_NOALLOC_MAX_SEGMENTS segments, 5 doubles each = 1.25 MiB, held by every GPU that
// loads the library until the process exits.
//   layout   [0, 4 MAX)       end points (x0, y0, x1, y1) of the segments, copied from the caller's host buffer
//            [4 MAX, 5 MAX)   the optical depths the kernel writes, copied back to the caller's host buffer
// `g_noalloc_scratch` is a host-side stand-in in host code, NOT a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked for once per device with that device current (noalloc_scratch_ptr below).
// The array is shared by every call on a GPU and keeps the contents of the last call.  Nothing in it has to start at
// zero - a call overwrites the 4n inputs it reads and the kernel writes all n outputs it returns - but the output range is
// cleared first all the same, so that a launch that fails cannot hand back the previous call's values.  Calls that overlap
// on one GPU (two host threads, two handles on the same device) race on the array; the entry point synchronises its stream
// before it returns, so calls made one after the other are safe.
// ---------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_scratch[5 * ANISO_NOALLOC_MAX_SEGMENTS];

namespace {

double* noalloc_scratch_ptr(int device) {  // the caller has made `device` current
    static std::atomic<double*> cache[64];
    if (device < 0 || device >= 64) throw std::invalid_argument("device ordinal out of range for the allocation-free probe");
    double* p = cache[device].load();
    if (p) return p;
    void* sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);
    if (e != cudaSuccess || !sym) throw std::runtime_error(std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
    p = static_cast<double*>(sym);
    cache[device].store(p);  // the address of a symbol is fixed per device: a second thread finding nullptr stores the same value
    return p;
}

}  // namespace

extern "C" int aniso_line_integral_noalloc(aniso_t* h, int n, const double* pts4, double* out) {
    MULTI_R0(h);
    API_BEGIN
    REQUIRE(h && pts4 && out && n >= 0, "bad arguments");
    if (n > ANISO_NOALLOC_MAX_SEGMENTS)
        throw std::invalid_argument("aniso_line_integral_noalloc: " + std::to_string(n) + " segments exceed the fixed scratch of " +
                                    std::to_string(ANISO_NOALLOC_MAX_SEGMENTS) + " (use aniso_line_integral, or split the call)");
    // no lazy set-up here: creating streams and uploading the geometry allocates
    if (!h->dev_ready || !h->coeff_set) throw std::logic_error("setCoeff must precede aniso_line_integral_noalloc (it sets the device up)");
    check_probe_points(n, pts4);
    if (n == 0) return ANISO_OK;
    CK(cudaSetDevice(h->device));
    double* scr = noalloc_scratch_ptr(h->device);
    double* res = scr + (size_t)4 * ANISO_NOALLOC_MAX_SEGMENTS;
    CK(cudaMemcpyAsync(scr, pts4, sizeof(double) * 4 * n, cudaMemcpyHostToDevice, h->st));
    CK(cudaMemsetAsync(res, 0, sizeof(double) * n, h->st));
    LAUNCH(h, launch_probe_line_integral(h->S, n, scr, res, h->st));
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, res, sizeof(double) * n, cudaMemcpyDeviceToHost, h->st));
    CK(cudaStreamSynchronize(h->st));
    return ANISO_OK;
    API_END
}


// ---------------------------------------------------------------------------------------------------
// bbfmm::kernel for a caller-supplied eval (bbfmm.h:451-505): the black-box interface itself.  eval is a host callback and
// cannot run on the device, so the operators are tabulated on the host - once, OpenMP over the U/W/V/X entries, in
// the block-CSR layout of the aniso path - and the apply is the device path: P2M / M2M, streamed M2L and P2P blocks, L2L /
// L2P (run_fmm with mode 0).  Nothing is assumed about eval: every block is stored (no symmetric storage).
// ---------------------------------------------------------------------------------------------------
extern "C" {

int aniso_fmm_create(int np, int n, const double* x, const double* y, int leaf_capacity, int maxLevel, aniso_t** out) {
    API_BEGIN
    REQUIRE(out && x && y, "null pointer");
    aniso_handle* h = new aniso_handle();
    try {
        build_plan_points(h->plan, np, n, x, y, leaf_capacity, maxLevel);
    } catch (const std::exception& e) {
        delete h;
        throw std::invalid_argument(e.what());
    }
    const HostPlan& p = h->plan;
    h->all_slot_begin.assign(1, 0);
    h->all_slot_end.assign(1, p.N);
    h->leaf_begin = 0;
    h->leaf_end = (int)p.leaves.size();
    h->slot0 = 0;
    h->nown = p.N;
    *out = h;
    return ANISO_OK;
    API_END
}

int aniso_fmm_cache(aniso_t* h, aniso_eval_fn eval, void* ctx, int threads) {
    MULTI_REJECT(h, "aniso_fmm_cache");
    API_BEGIN
    REQUIRE(h && eval, "null pointer");
    if (!h->plan.generic) throw std::logic_error("aniso_fmm_cache needs a handle made by aniso_fmm_create");
    ensure_device(h);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->st));
    free_caches(h);
    const HostPlan& p = h->plan;
    OperatorCache c;
    c.sign = 1.0;
    c.id = 0;
    c.part = PART_BOTH;
    c.p2p = dalloc<double>(h, (size_t)p.p2p.nnz + 2);
    c.m2l = dalloc<double>(h, (size_t)p.m2l.nnz + 2);
    std::vector<double> host((size_t)std::max(p.p2p.nnz, p.m2l.nnz));
    tabulate_operator(p, p.p2p, false, eval, ctx, threads, host.data());
    CK(cudaMemcpy(c.p2p, host.data(), sizeof(double) * p.p2p.nnz, cudaMemcpyHostToDevice));
    tabulate_operator(p, p.m2l, true, eval, ctx, threads, host.data());
    CK(cudaMemcpy(c.m2l, host.data(), sizeof(double) * p.m2l.nnz, cudaMemcpyHostToDevice));
    h->fill_counts[0] = p.p2p.nnz;
    h->fill_counts[1] = p.m2l.nnz;
    h->ops[std::make_pair(0, (int)PART_BOTH)] = c;
    return ANISO_OK;
    API_END
}

int aniso_fmm_apply(aniso_t* h, const double* charge, double* potential) {
    MULTI_REJECT(h, "aniso_fmm_apply");
    API_BEGIN
    REQUIRE(h && charge && potential, "null pointer");
    if (!h->plan.generic) throw std::logic_error("aniso_fmm_apply needs a handle made by aniso_fmm_create");
    auto it = h->ops.find(std::make_pair(0, (int)PART_BOTH));
    if (it == h->ops.end()) throw std::logic_error("aniso_fmm_cache must precede aniso_fmm_apply");
    CK(cudaSetDevice(h->device));
    to_slots(h, charge, h->d_q);            // chargeTree (bbfmm.h:485): used as given, no quadrature weights
    run_fmm(h, it->second, h->d_q, h->d_b);  // mode 0: upPass + downPassFast, nothing else
    from_slots(h, h->d_b, potential);
    return ANISO_OK;
    API_END
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// Single-process multi-GPU (SURVEY 5 / 8b "one process, 8 streams"; aniso_create_sharded(..., ndev, devs) of the required
// export list): main.cpp:7-150 and AnisoWrapper.cpp:10-138 are single-process callers, so a handle that needs one process
// per GPU leaves them on one device.  A group handle owns one sharded rank handle per device; each collective call is run
// by one host thread per rank (CUDA device bound per thread), and the ranks exchange exactly as in the process-per-GPU
// form: NCCL for the set-up, the peer-memory collectives per matvec (CUDA IPC cannot map a buffer into the process that
// owns it, so ranks that are threads of one process publish the raw device pointer and enable peer access instead,
// peer.cu peer_open).  Host vectors are shared: inputs are read by every rank, each rank writes its own rows of an output.
// ---------------------------------------------------------------------------------------------------
template <class F>
static int multi_all(aniso_handle* h, F&& fn) {
    const int P = (int)h->multi.size();
    std::vector<int> rc(P, 0);
    std::vector<std::string> msg(P);
    std::vector<std::thread> th;
    th.reserve(P);
    for (int k = 0; k < P; ++k)
        th.emplace_back([&, k] {
            if (h->multi[k] && h->multi[k]->device >= 0) cudaSetDevice(h->multi[k]->device);
            rc[k] = fn(k, h->multi[k]);
            if (rc[k] < 0) msg[k] = aniso_last_error();
        });
    for (auto& t : th) t.join();
    for (int k = 0; k < P; ++k)
        if (rc[k] < 0) return fail(rc[k], "rank " + std::to_string(k) + ": " + msg[k]);
    return rc[0];
}

extern "C" int aniso_create_multi(int sz, int deg, int K, double g, int sr, int np, int maxLevel, int ndev, const int* devs,
                                  aniso_t** out) {
    API_BEGIN
    REQUIRE(out && ndev >= 1 && ndev <= 64 && (devs || ndev == 1), "bad arguments");
    if (ndev == 1) return aniso_create_sharded(sz, deg, K, g, sr, np, maxLevel, 1, 0, nullptr, devs ? devs[0] : -1, out);
    for (int a = 0; a < ndev; ++a)
        for (int b = a + 1; b < ndev; ++b) REQUIRE(devs[a] != devs[b] && devs[a] >= 0 && devs[b] >= 0, "device ordinals must be distinct and non-negative");
    unsigned char uid[128];
    int rc = aniso_nccl_unique_id(uid);
    if (rc < 0) return rc;
    aniso_handle* grp = new aniso_handle();
    grp->multi.assign(ndev, nullptr);
    rc = multi_all(grp, [&](int k, aniso_t*) -> int {  // the host plans (tree, lists) are built side by side
        return aniso_create_sharded(sz, deg, K, g, sr, np, maxLevel, ndev, k, uid, devs[k], &grp->multi[k]);
    });
    for (aniso_handle* r : grp->multi)
        if (r) r->in_group = true;
    if (rc < 0) {
        const std::string keep = aniso_last_error();
        for (aniso_handle* r : grp->multi) delete r;  // no device state yet
        delete grp;
        return fail(rc, keep);
    }
    *out = grp;
    return ANISO_OK;
    API_END
}
