/* aniso_b200 - C-ABI of the B200-native aniso hot path (integral-operator matvec `mapping` + GMRES).
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  Each entry point names the
 * reference interface it replaces (file:line into lowrank/aniso).  INTEGRATION.md shows the MEX / C++ / ctypes
 * bindings a maintainer would put on top.
 *
 * Conventions
 *   - every function returns 0 on success or a negative ANISO_E* code; aniso_last_error() gives the text
 *     (thread-local).  No exception crosses this boundary.
 *   - vectors named `..._host` are caller-owned host buffers of length N = aniso_num_nodes() in the REFERENCE node
 *     order (Geometry.cpp:28-60); `..._dev` are device pointers on the handle's device in the library's SLOT order
 *     (depth-first leaf order of the quadtree, see aniso_slot_permutation).
 *   - all device memory is owned by the handle; calls on one handle must be serialised by the caller (as in the
 *     reference, which is not re-entrant either).
 *   - there is NO CPU fallback: every compute entry point fails with ANISO_ENODEVICE when no CUDA device is usable.
 *     The host-only entry points (geometry, tree, partition introspection) work without a GPU.
 */
#ifndef ANISO_B200_H
#define ANISO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct aniso_handle aniso_t;

enum {
    ANISO_OK = 0,
    ANISO_EINVAL = -1,    /* bad argument (null pointer, id out of range, bad size) */
    ANISO_ENODEVICE = -2, /* no usable CUDA device / driver */
    ANISO_ECUDA = -3,     /* a CUDA call failed */
    ANISO_ESTATE = -4,    /* call order violated (setCoeff before cache, cache before mapping) */
    ANISO_ENOMEM = -5,
    ANISO_ENCCL = -6
};

/* Sharded handles (aniso_create_sharded with nranks > 1) are SPMD objects: every rank makes the same calls in the same
 * order, as in any NCCL program.  The entry points that take HOST vectors move 1/nranks of each vector per rank over PCIe
 * and all-gather the pieces over NVLink, so they are collective too. */

/* kernel parts of KernelFactory.h:22-23: realParts = singular cos(k theta)/r, imagParts = attenuated remainder */
enum { ANISO_PART_REAL = 1, ANISO_PART_IMAG = 2, ANISO_PART_BOTH = 3 };

const char* aniso_last_error(void);
const char* aniso_version(void);

/* ---- life cycle ------------------------------------------------------------------------------------------ */
/* Aniso::Aniso(geometry_size, geometry_degree, kernel_size, kernel_g, kernel_degree, np, maxLevel)  Aniso.h:14-15,
 * Aniso.cpp:7-13; MEX op 'new' AnisoWrapper.cpp:10-25.  Host set-up only (geometry, quadtree, lists). */
int aniso_create(int geometry_size, int geometry_degree, int kernel_size, double kernel_g, int kernel_degree, int np,
                 int maxLevel, aniso_t** out);
/* Same, as rank `rank` of `nranks` devices on one node: target leaves are partitioned by quadtree ownership
 * (SURVEY 8e).  nccl_unique_id = the 128 bytes of aniso_nccl_unique_id() from rank 0, identical on all ranks.
 * device = CUDA device ordinal to use (-1: current device). */
int aniso_create_sharded(int geometry_size, int geometry_degree, int kernel_size, double kernel_g, int kernel_degree,
                         int np, int maxLevel, int nranks, int rank, const void* nccl_unique_id, int device,
                         aniso_t** out);
int aniso_nccl_unique_id(void* out128);
/* Same partition, ONE process: a group handle that drives `ndev` GPUs (device ordinals `devs`) of this node from a
 * single-process caller such as main.cpp:7-150 or the MEX gateway AnisoWrapper.cpp:10-138 (SURVEY 8b:
 * aniso_create_sharded(..., ndev, devs)).  Every entry point that takes HOST vectors works on the group as on a one-GPU
 * handle - inputs are read by all ranks, outputs come back complete - and is run by one host thread per GPU inside the
 * call.  Entry points that take device pointers or a stream (`*_dev`, aniso_stream, aniso_to_slots, ...) return
 * ANISO_EINVAL on a group, and a preconditioner callback is not accepted (it would run on every rank's thread at once).
 * ndev == 1 gives an ordinary handle on devs[0]. */
int aniso_create_multi(int geometry_size, int geometry_degree, int kernel_size, double kernel_g, int kernel_degree, int np,
                       int maxLevel, int ndev, const int* devs, aniso_t** out);
/* MEX op 'delete' AnisoWrapper.cpp:27-31 / Aniso::~Aniso */
int aniso_destroy(aniso_t* h);

/* ---- geometry (host only) -------------------------------------------------------------------------------- */
int aniso_num_nodes(const aniso_t* h);   /* Geometry::numberOfNodes, Geometry.h:60 */
int aniso_kernel_size(const aniso_t* h); /* KernelFactory::kernelSize = 2*kernel_size-1, KernelFactory.cpp:10-13 */
/* MEX op 'getNodes' AnisoWrapper.cpp:33-44 (+ Geometry::weights); any pointer may be NULL */
int aniso_get_nodes(const aniso_t* h, double* x, double* y, double* w);
/* get_legendre_data, Quadrature.cpp:5418-22191 (table order) */
int aniso_gl_table(int degree, double* x, double* w);
/* Geometry tables: interpolate (c2*c2 row-major), nearMapping (16c2*c2), legendreNorms (c2), refine_quad_x/y,
 * refine_weight (16c2 each); Geometry.h:49-58.  NULL pointers are skipped. */
int aniso_geometry_tables(const aniso_t* h, double* interpolate, double* nearMapping, double* legendreNorms,
                          double* refine_x, double* refine_y, double* refine_w);
/* singX/singY/singW, KernelFactory.h:36 (c2 * 8*sr*sr each), filled by singPrecompute KernelFactory.cpp:929-986 */
int aniso_sing_tables(const aniso_t* h, double* sx, double* sy, double* sw);
/* bbfmm::tree statistics (bbfmm.h:176-447): out[9] = nodes, leaves, maxLevel, sum|U|, sum|V|, sum|W|, sum|X|,
 * P2P pairs, max points per leaf */
int aniso_tree_stats(const aniso_t* h, int64_t out[9]);
/* per tree node (library numbering, level-major): geom[4n] = cx,cy,rx,ry; info[4n] = level, leaf, nSource, parent.
 * Returns the node count (also when cap is too small). */
int aniso_tree_nodes(const aniso_t* h, int cap, double* geom, int* info);
/* interaction lists of one node: kinds 0=V 1=X 2=U 3=W; returns the entry count */
int aniso_node_lists(const aniso_t* h, int node, int cap, int* ids, int* kinds);
/* slot order used by the *_dev entry points: perm[slot] = reference node index */
int aniso_slot_permutation(const aniso_t* h, int* perm);
/* ownership of rank `rank` out of `nranks`: slots [*slot_begin, *slot_end) and leaf count; host only */
int aniso_partition(const aniso_t* h, int nranks, int rank, int* slot_begin, int* slot_end, int* nleaves);
int aniso_owned_range(const aniso_t* h, int* slot_begin, int* slot_end);

/* ---- coefficients and caches ----------------------------------------------------------------------------- */
/* MEX op 'setCoeff' AnisoWrapper.cpp:46-69: stores sigma_s/sigma_t, runs interpolation() KernelFactory.cpp:212-227
 * (singPrecompute / makeKernels are part of aniso_create here).  Invalidates every cache. */
int aniso_set_coeff(aniso_t* h, const double* sigma_s_host, const double* sigma_t_host);
/* sigma_t_Coeff / sigma_s_Coeff (KernelFactory.h:77-78), N doubles each, cell-major; NULL skipped */
int aniso_sigma_coeff(aniso_t* h, double* coeff_t_host, double* coeff_s_host);
/* MEX op 'cache' AnisoWrapper.cpp:72-90 = runKernelsCache + runKernelsCacheSing + refineAddOnCache +
 * singularAddCache (KernelFactory.cpp:323-334, 279-288, 550-609, 752-788): evaluates every P2P/M2L operator block
 * of kernel `id` on the device (fused real+imag), and the correction stencils. */
int aniso_cache(aniso_t* h, int id);
/* The four cache steps of the reference one at a time (main.cpp:60-76): part = ANISO_PART_IMAG: runKernelsCache
 * (KernelFactory.cpp:323-334), ANISO_PART_REAL: runKernelsCacheSing (:279-288), ANISO_PART_BOTH: the fused operators
 * aniso_cache builds, 0: refineAddOnCache + singularAddCache (+ the nearRemoval table) as correction stencils. */
int aniso_cache_part(aniso_t* h, int id, int part);
/* Options of a handle.  "p2p" = "cached" (default; runKernelsCache keeps every U/W tile, bbfmm.h:949-1039) or "onfly": the
 * near-field tiles are evaluated while they are applied, as in the reference's uncached kernel::run / downPass
 * (bbfmm.h:507-532, 863-946), which saves 8 bytes per stored P2P pair and kernel id for memory-constrained runs (SURVEY 8f
 * N2).  Same results to rounding.  Changing it drops the cached operators; leaves of at most 64 points. */
int aniso_set_option(aniso_t* h, const char* name, const char* value);
/* bytes held for kernel id (Aniso::displayKernelCacheSize, Aniso.cpp:19-47) */
size_t aniso_cache_bytes(const aniso_t* h, int id);
/* number of kernel evaluations the last aniso_cache(id) performed: out[0] = P2P pairs, out[1] = M2L pairs */
int aniso_cache_counts(const aniso_t* h, int64_t out[2]);
/* device time (CUDA events, ms) of the last operator fill: out[0] = P2P evaluator, out[1] = M2L evaluator,
 * out[2] = correction-stencil build */
int aniso_cache_times(const aniso_t* h, double out[3]);

/* ---- the matvec ------------------------------------------------------------------------------------------- */
/* MEX op 'mapping' AnisoWrapper.cpp:92-136 == apply_mapping main.cpp:78-119.  Host in, host out (H2D + D2H inside). */
int aniso_mapping(aniso_t* h, const double* charge_host, int id, double* out_host);
/* Device-resident form, slot order.  charge_dev: full vector (N); out_dev: the owned range (N on one GPU).
 * Runs on the handle's stream; returns without synchronising. */
int aniso_mapping_dev(aniso_t* h, const double* charge_dev, int id, double* out_dev);
/* y = x - mapping_0(sigma_s .* x): forwardOperator of main.cpp:123-136, device-resident. */
int aniso_forward_dev(aniso_t* h, const double* x_dev, double* y_dev);

/* stage entry points for stage-wise parity (host vectors, reference order).  The three corrections ACCUMULATE
 * into ret like the reference; the two FMM calls overwrite it (bbfmm.h:579). */
int aniso_interpolation(aniso_t* h, const double* f_host, double* coeff_host);            /* KernelFactory.cpp:988-1005 */
int aniso_run_kernels_fast(aniso_t* h, int id, const double* f_host, double* ret_host);      /* :354-363 (imagParts) */
int aniso_run_kernels_fast_sing(aniso_t* h, int id, const double* f_host, double* ret_host); /* :379-387 (realParts) */
int aniso_near_removal(aniso_t* h, int id, const double* f_host, double* ret_host);          /* :445-478 */
int aniso_refine_add_on_fast(aniso_t* h, int id, const double* f_host, double* ret_host);    /* :662-709 */
int aniso_singular_add_fast(aniso_t* h, int id, const double* fcoeff_host, double* ret_host);/* :828-860 */

/* ---- GMRES ------------------------------------------------------------------------------------------------ */
/* GMRES(A, x, b, m, max_iter, tol) gmres.cpp:53-169 with A = forwardOperator (main.cpp:123-136), device-resident
 * Arnoldi.  x_host: initial guess in, solution out.  Returns 0 converged / 1 not converged (like the reference) or a
 * negative error.  resid_hist[j-1] = relative residual the reference prints at iteration j, resid_hist[*iters] = the
 * final one (hist_cap entries available; may be NULL).  The Krylov basis is orthogonalised by classical Gram-Schmidt
 * applied twice (same Hessenberg entries to rounding and the same iteration counts as the reference's modified
 * Gram-Schmidt, gmres.cpp:115-119, which ANISO_GMRES_ORTHO=mgs selects); a zero x_host skips the first operator
 * application (r = b). */
int aniso_gmres(aniso_t* h, const double* b_host, double* x_host, int m, int max_iter, double tol, int* iters,
                double* resid_hist, int hist_cap);
/* Same with device vectors in slot order (owned range), no host copies of vectors. */
int aniso_gmres_dev(aniso_t* h, const double* b_dev, double* x_dev, int m, int max_iter, double tol, int* iters,
                    double* resid_hist, int hist_cap);

/* Left-preconditioned GMRES: the DSA preconditioner hook of the MATLAB application, `prec` of aniso.m:111-119 handed to
 * gmres(A, rhs, 400, 1e-11, 400, @obj.prec) at aniso.m:165 (gmres.cpp:7 lists it as "todo: preconditioner").  The callback
 * maps r -> z = M^-1 r on HOST vectors of n doubles in reference node order (n = N, or K*N mode-major for the block solve) and
 * returns 0; it is applied to b, to the initial residual and after every operator application, so the iteration runs on
 * M^-1 A x = M^-1 b and the reported residuals are |M^-1 (b - A x)| / |M^-1 b| (MATLAB's convention).  prec == NULL is the
 * unpreconditioned solve.  A non-zero return aborts with ANISO_ECUDA-class error text.  The FEM solve behind aniso.m's
 * callback (femm) is outside this library. */
typedef int (*aniso_prec_fn)(void* ctx, int n, const double* r_host, double* z_host);
int aniso_gmres_prec(aniso_t* h, const double* b_host, double* x_host, int m, int max_iter, double tol, aniso_prec_fn prec,
                     void* ctx, int* iters, double* resid_hist, int hist_cap);

/* ---- multi-mode block operator (MATLAB application layer: aniso.m:121-156 forward / mforward, aniso.m:158-172 solve) ----
 * Block vectors hold K = kernel_size angular modes of N values, mode-major: mode a at [a*N, (a+1)*N), reference node order.
 *   weighted = 0  `forward`:  out_i = sum_{j=-(K-1)}^{K-1} mapping_{|i+j|}(u_|j|)                         (aniso.m:121-136)
 *   weighted = 1  `mforward`: out_i = sum_j chi_|j| mapping_{|i+j|}(sigma_s .* u_|j|), chi_a = (g^a - g^K)/(1 - g^K)
 *                                                                                                       (aniso.m:138-156)
 * aniso_block_gmres solves (I - mforward) x = rhs with the reference's own GMRES (gmres.cpp:53-169; MATLAB's gmres is not
 * available to the oracle either): aniso.m:165-169 uses restart 400, tol 1e-11, maxit 400.  The _dev forms take K blocks of
 * OWNED rows in slot order (block a at [a*nown, (a+1)*nown)). */
int aniso_block_forward(aniso_t* h, const double* u_host, int weighted, double* out_host);
int aniso_block_forward_dev(aniso_t* h, const double* u_dev, int weighted, double* out_dev);
int aniso_block_gmres(aniso_t* h, const double* rhs_host, double* x_host, int m, int max_iter, double tol, int* iters,
                      double* resid_hist, int hist_cap);
int aniso_block_gmres_dev(aniso_t* h, const double* rhs_dev, double* x_dev, int m, int max_iter, double tol, int* iters,
                          double* resid_hist, int hist_cap);
/* aniso.m:165 with pr == 1: gmres(A, rhs, 400, 1e-11, 400, @obj.prec); callback on K*N host doubles, mode-major */
int aniso_block_gmres_prec(aniso_t* h, const double* rhs_host, double* x_host, int m, int max_iter, double tol,
                           aniso_prec_fn prec, void* ctx, int* iters, double* resid_hist, int hist_cap);

/* ---- the black-box engine for a caller-supplied kernel: bbfmm::kernel (bbfmm.h:451-505) --------------------------------
 * kernel::eval : std::function<double(point& source, point& target)> (bbfmm.h:457) is a host callback and cannot run on the
 * device, so the operators are tabulated on the host (once, OpenMP over the U/W/V/X entries: kernelEvalIndex /
 * kernelEvalChebyshev, bbfmm.h:770-804) and applied by the device path.  Nothing is assumed about eval (every block is
 * stored).  Sources and targets are one point set, as in every call the reference makes (KernelFactory.cpp:284-286).
 *   aniso_fmm_create  kernel::initialize(np, points, points, ., n, n, rank, maxLevel) (bbfmm.h:476-505): tree, lists, transfer
 *                     operators; leaf_capacity = the `_rank` argument (leaf iff nSource <= rank, bbfmm.h:264); host only
 *   aniso_fmm_cache   the evaluation half of runCache (bbfmm.h:535-562 -> downPassCache 949-1039); eval is called from
 *                     `threads` OpenMP threads at once (0 = all), as under the reference's OpenMP tasks
 *   aniso_fmm_apply   runFast (bbfmm.h:565-592): potential = FMM(charge); host vectors in the caller's point order
 * The handle is an aniso_t: destroy it with aniso_destroy; the Aniso entry points reject it with ANISO_ESTATE. */
typedef double (*aniso_eval_fn)(void* ctx, double source_x, double source_y, double target_x, double target_y);
int aniso_fmm_create(int np, int n, const double* x, const double* y, int leaf_capacity, int maxLevel, aniso_t** out);
int aniso_fmm_cache(aniso_t* h, aniso_eval_fn eval, void* ctx, int threads);
int aniso_fmm_apply(aniso_t* h, const double* charge_host, double* potential_host);

/* ---- device probes (parity tests of K1/K2 pieces) ---------------------------------------------------------- */
/* lineIntegral(x0,y0,x1,y1) KernelFactory.cpp:67-166 for n segments, pts4 = n x 4 host doubles */
int aniso_line_integral(aniso_t* h, int n, const double* pts4_host, double* out_host);
/* realParts[id].eval / imagParts[id].eval (makeKernels KernelFactory.cpp:240-267): pts4 = (ax,ay,bx,by) */
int aniso_kernel_eval(aniso_t* h, int id, int part, int n, const double* pts4_host, double* out_host);

/* aniso_line_integral without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that must not call cudaMalloc (tight serving loops, harnesses that forbid it).  Same kernel, same results; the
 * device staging buffer is a fixed array in the library image (5 doubles per segment: 1.25 MiB on every GPU that loads the
 * library, until the process exits) instead of a cudaMalloc / cudaFree pair per call.  n > ANISO_NOALLOC_MAX_SEGMENTS fails
 * with ANISO_EINVAL; a handle whose aniso_set_coeff has not run fails with ANISO_ESTATE (no lazy set-up: that would allocate).
 * The array is shared by all calls on a GPU and keeps the last call's contents: calls that overlap on one GPU (two host
 * threads, two handles on one device) race; calls made one after the other are safe (the call synchronises before returning). */
enum { ANISO_NOALLOC_MAX_SEGMENTS = 32768 };
int aniso_line_integral_noalloc(aniso_t* h, int n, const double* pts4_host, double* out_host);

/* ---- plumbing ---------------------------------------------------------------------------------------------- */
void* aniso_stream(aniso_t* h);          /* cudaStream_t all work of the handle is launched on */
int aniso_device(const aniso_t* h);
int aniso_synchronize(aniso_t* h);
/* device buffer helpers for callers without a CUDA runtime of their own */
int aniso_dev_alloc(aniso_t* h, size_t bytes, void** out);
int aniso_dev_free(aniso_t* h, void* p);
/* page-locked host buffers: vectors passed to aniso_mapping / aniso_gmres / aniso_to_slots / aniso_from_slots from such a
 * buffer (or from any memory the caller registered with cudaHostRegister) are DMA'd directly, others are staged through
 * the handle's own pinned buffer in 2 MB chunks */
int aniso_host_alloc(aniso_t* h, size_t bytes, void** out);
int aniso_host_free(aniso_t* h, void* p);
int aniso_to_slots(aniso_t* h, const double* ref_order_host, double* slot_order_dev);   /* H2D + permute */
int aniso_from_slots(aniso_t* h, const double* slot_order_dev, double* ref_order_host); /* permute + D2H (owned rows only
                                                                                            when sharded: others untouched) */
/* kernels launched by this handle since creation (bench.py's gpu_launches) */
int64_t aniso_launch_count(const aniso_t* h);
/* bytes the cached-apply kernels must move per mapping: out[0] = P2P, out[1] = M2L (stored blocks, i.e. one of each
 * symmetric pair when the blocks are uniform, plus the per-target scratch rows they write), out[2] = stencil */
int aniso_apply_bytes(const aniso_t* h, int64_t out[3]);
/* time the streaming apply kernels alone: `reps` back-to-back launches of (0) M2L stream, (1) P2P stream on the cache of `id`,
 * CUDA events on the handle's stream; returns average milliseconds per launch */
int aniso_time_kernel(aniso_t* h, int id, int which, int reps, double* ms_per_launch);
/* `reps` applications of the GMRES operator y = x - mapping_0(sigma_s x) (main.cpp:123-136) on device vectors, CUDA events on
 * the handle's stream: average milliseconds per application (collective on a sharded handle) */
int aniso_time_forward(aniso_t* h, int reps, double* ms_per_apply);
/* one line naming the code paths in effect (library version, handle layout, GMRES form, every ANISO_* development switch
 * set in the environment): lets a benchmark record describe itself.  Thread-local storage. */
const char* aniso_get_config(const aniso_t* h);

#ifdef __cplusplus
}
#endif
#endif /* ANISO_B200_H */
