"""Host-side mirror of the reference's plugin interface for the hot path.

  Aniso            class/@Aniso/Aniso.m:1-34 (new / delete / setCoeff / getNodes / cache / mapping) plus the C++
                   stage methods of KernelFactory.h:65-75 and GMRES (gmres.h:23), all forwarding to the C-ABI.
  AnisoWrapper     the string-dispatched MEX gateway AnisoWrapper.cpp:10-138 with its integer-handle session table.

Vectors are numpy float64 arrays in the reference node order.  Nothing here computes: every call lands in
libaniso_b200.so, which fails with ANISO_ENODEVICE when no CUDA device is present.
"""
import ctypes as C

import numpy as np

from . import _lib


def _vec(a, n=None, name="vector"):
    v = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))
    if n is not None and v.size != n:
        raise ValueError("%s has %d entries, expected %d" % (name, v.size, n))
    return v


class Aniso:
    """Aniso(ds, qr, ks, as, sr, fn, fm) == Aniso(geometry_size, geometry_degree, kernel_size, kernel_g,
    kernel_degree, np, maxLevel) (Aniso.h:14-15)."""

    def __init__(self, ds, qr, ks, g, sr, fn, fm, nranks=1, rank=0, nccl_uid=None, device=-1, devices=None):
        """devices=[0, 1, ...]: one process drives these GPUs (aniso_create_multi); nranks / rank / nccl_uid: this process is
        one rank of a process-per-GPU job (aniso_create_sharded)."""
        self.L = _lib.load()
        h = C.c_void_p()
        # MEX passes doubles and truncates with int(*p) (AnisoWrapper.cpp:22-24)
        args = (int(ds), int(qr), int(ks), float(g), int(sr), int(fn), int(fm))
        if devices is not None:
            devs = np.ascontiguousarray(devices, np.int32)
            _lib.check(self.L.aniso_create_multi(*args, devs.size, devs, C.byref(h)))
        elif nranks == 1:
            _lib.check(self.L.aniso_create(*args, C.byref(h)))
        else:
            buf = C.create_string_buffer(bytes(nccl_uid), 128)
            _lib.check(self.L.aniso_create_sharded(*args, int(nranks), int(rank), buf, int(device), C.byref(h)))
        self.h = h
        self.nranks, self.rank = nranks, rank
        self.numberOfNodes = self.L.aniso_num_nodes(h)
        self.kernelSize = self.L.aniso_kernel_size(h)
        self.sz, self.deg, self.sr, self.np, self.maxLevel = int(ds), int(qr), int(sr), int(fn), int(fm)
        self.sigma_s = self.sigma_t = None
        self._pinned = []

    # -- life cycle --------------------------------------------------------------------------------------
    def delete(self):
        if getattr(self, "h", None):
            for p in self._pinned:
                self.L.aniso_host_free(self.h, p)
            self._pinned = []
            self.L.aniso_destroy(self.h)
            self.h = None

    close = delete

    def __del__(self):
        try:
            self.delete()
        except Exception:
            pass

    # -- MEX ops -------------------------------------------------------------------------------------------
    def getNodes(self):
        """N x 2 array (x column, y column), AnisoWrapper.cpp:33-44."""
        x, y, _ = self.nodes()
        return np.stack([x, y], axis=1)

    def nodes(self):
        n = self.numberOfNodes
        x, y, w = np.zeros(n), np.zeros(n), np.zeros(n)
        _lib.check(self.L.aniso_get_nodes(self.h, x, y, w))
        return x, y, w

    @property
    def weights(self):
        return self.nodes()[2]

    def setCoeff(self, sigma_s, sigma_t):
        n = self.numberOfNodes
        self.sigma_s, self.sigma_t = _vec(sigma_s, n, "sigma_s"), _vec(sigma_t, n, "sigma_t")
        _lib.check(self.L.aniso_set_coeff(self.h, self.sigma_s, self.sigma_t))

    def set_option(self, name, value):
        """aniso_set_option: e.g. ("p2p", "onfly") evaluates the near-field tiles during the apply instead of caching them"""
        _lib.check(self.L.aniso_set_option(self.h, str(name).encode(), str(value).encode()))

    def cache(self, kid):
        _lib.check(self.L.aniso_cache(self.h, int(kid)))

    def mapping(self, charge, kid):
        out = np.zeros(self.numberOfNodes)
        _lib.check(self.L.aniso_mapping(self.h, _vec(charge, self.numberOfNodes, "charge"), int(kid), out))
        return out

    # -- C++ stage methods (KernelFactory.h:65-75) ---------------------------------------------------------
    def interpolation(self, f):
        """interpolation(h, coeff): per-cell Legendre coefficients, returned as (cells, deg^2)."""
        out = np.zeros(self.numberOfNodes)
        _lib.check(self.L.aniso_interpolation(self.h, _vec(f, self.numberOfNodes), out))
        return out.reshape(-1, self.deg * self.deg)

    def runKernelsFast(self, kid, f):
        ret = np.zeros(self.numberOfNodes)
        _lib.check(self.L.aniso_run_kernels_fast(self.h, int(kid), _vec(f, self.numberOfNodes), ret))
        return ret

    def runKernelsFastSing(self, kid, f):
        ret = np.zeros(self.numberOfNodes)
        _lib.check(self.L.aniso_run_kernels_fast_sing(self.h, int(kid), _vec(f, self.numberOfNodes), ret))
        return ret

    def nearRemoval(self, kid, f, ret):
        _lib.check(self.L.aniso_near_removal(self.h, int(kid), _vec(f, self.numberOfNodes), ret))
        return ret

    def refineAddOnFast(self, kid, f, ret):
        _lib.check(self.L.aniso_refine_add_on_fast(self.h, int(kid), _vec(f, self.numberOfNodes), ret))
        return ret

    def singularAddFast(self, kid, fcoeff, ret):
        _lib.check(self.L.aniso_singular_add_fast(self.h, int(kid), _vec(fcoeff, self.numberOfNodes), ret))
        return ret

    def mapping_stages(self, charge, kid):
        """The body of MEX 'mapping' (AnisoWrapper.cpp:111-130) stage by stage, for stage-wise parity tests."""
        charge = _vec(charge, self.numberOfNodes)
        scaled = charge * self.weights
        coeff = self.interpolation(charge)
        o = {"imag_fmm": self.runKernelsFast(kid, scaled)}
        t = self.runKernelsFastSing(kid, scaled)
        o["real_fmm"] = t.copy()
        o["k7"] = self.nearRemoval(kid, scaled, t).copy()
        o["k8"] = self.refineAddOnFast(kid, scaled, t).copy()
        o["k9"] = self.singularAddFast(kid, coeff, t).copy()
        o["out"] = (t + o["imag_fmm"]) * (1.0 / np.pi / 2.0)
        return o

    def displayKernelCacheSize(self):
        """Aniso.cpp:19-47: megabytes held per kernel id."""
        return {k: self.L.aniso_cache_bytes(self.h, k) / 1e6 for k in range(self.kernelSize)}

    # -- GMRES (gmres.cpp:53-169 with the operator of main.cpp:123-136) ---------------------------------------
    @staticmethod
    def _prec_callback(prec):
        """Wraps prec(r) -> z (numpy vectors, reference node order) as the C callback of aniso_gmres_prec."""
        def cb(ctx, n, r, z):
            try:
                out = np.asarray(prec(np.ctypeslib.as_array(r, shape=(n,)).copy()), dtype=np.float64).reshape(-1)
                if out.size != n:
                    return 2
                np.ctypeslib.as_array(z, shape=(n,))[:] = out
                return 0
            except Exception:  # no exception may cross the C boundary
                return 1
        return _lib.PREC_FN(cb)

    def gmres(self, b, x0=None, m=80, max_iter=400, tol=1e-12, prec=None):
        """GMRES(A, x, b, m, max_iter, tol) of gmres.cpp:53-169 on the operator of main.cpp:123-136; prec: optional left
        preconditioner r -> M^-1 r on host vectors (the hook aniso.m:165 hands to MATLAB's gmres)."""
        n = self.numberOfNodes
        x = np.zeros(n) if x0 is None else _vec(x0, n).copy()
        it = C.c_int(0)
        hist = np.zeros(max_iter + 2)
        if prec is None:
            rc = _lib.check(self.L.aniso_gmres(self.h, _vec(b, n), x, int(m), int(max_iter), float(tol), C.byref(it), hist,
                                               hist.size))
        else:
            cb = self._prec_callback(prec)
            rc = _lib.check(self.L.aniso_gmres_prec(self.h, _vec(b, n), x, int(m), int(max_iter), float(tol), cb, None,
                                                    C.byref(it), hist, hist.size))
        return x, rc, it.value, hist[: it.value + 1]

    # -- MATLAB application layer (aniso.m:121-172): multi-mode block operator ------------------------------------
    def forward(self, u):
        """aniso.m:121-136: rhs_i = sum_j mapping(u_|j|, |i-j|); u, rhs: K*N, mode-major."""
        n = self.kernelSize // 2 + 1
        out = np.zeros(n * self.numberOfNodes)
        _lib.check(self.L.aniso_block_forward(self.h, _vec(u, n * self.numberOfNodes, "u"), 0, out))
        return out

    def mforward(self, u):
        """aniso.m:138-156: rhs_i = sum_j chi_|j| mapping(sigma_s .* u_|j|, |i-j|), chi_a = (g^a - g^K)/(1 - g^K)."""
        n = self.kernelSize // 2 + 1
        out = np.zeros(n * self.numberOfNodes)
        _lib.check(self.L.aniso_block_forward(self.h, _vec(u, n * self.numberOfNodes, "u"), 1, out))
        return out

    def solve(self, charge, m=400, max_iter=400, tol=1e-11, prec=None):
        """aniso.m:158-172: rhs = forward(charge); (I - mforward) y = rhs by GMRES(400, 1e-11, 400) (the reference's
        gmres.cpp; MATLAB's gmres is the caller there); prec = the `pr == 1` branch (aniso.m:165), a callable r -> M^-1 r on
        K*N host values.  max_iter counts Arnoldi steps in total (gmres.cpp semantics); MATLAB's gmres(restart 400, maxit 400)
        would allow 400 outer cycles of 400 - a difference that only matters for solves that do not converge in 400 steps.
        Returns (y, rc, iterations, residual history)."""
        K = self.kernelSize // 2 + 1
        rhs = self.forward(charge)
        x = np.zeros(K * self.numberOfNodes)
        it = C.c_int(0)
        hist = np.zeros(max_iter + 2)
        if prec is None:
            rc = _lib.check(self.L.aniso_block_gmres(self.h, rhs, x, int(m), int(max_iter), float(tol), C.byref(it), hist, hist.size))
        else:
            cb = self._prec_callback(prec)
            rc = _lib.check(self.L.aniso_block_gmres_prec(self.h, rhs, x, int(m), int(max_iter), float(tol), cb, None,
                                                          C.byref(it), hist, hist.size))
        return x, rc, it.value, hist[: it.value + 1]

    # -- probes / introspection ---------------------------------------------------------------------------------
    def line_integral(self, pts4):
        p = np.ascontiguousarray(np.asarray(pts4, np.float64).reshape(-1, 4))
        out = np.zeros(len(p))
        _lib.check(self.L.aniso_line_integral(self.h, len(p), p.reshape(-1), out))
        return out

    def line_integral_noalloc(self, pts4):
        """line_integral on the library's fixed scratch array: no cudaMalloc / cudaFree (at most _lib.NOALLOC_MAX_SEGMENTS segments)"""
        p = np.ascontiguousarray(np.asarray(pts4, np.float64).reshape(-1, 4))
        out = np.zeros(len(p))
        _lib.check(self.L.aniso_line_integral_noalloc(self.h, len(p), p.reshape(-1), out))
        return out

    def kernel_eval(self, kid, part, pts4):
        p = np.ascontiguousarray(np.asarray(pts4, np.float64).reshape(-1, 4))
        out = np.zeros(len(p))
        _lib.check(self.L.aniso_kernel_eval(self.h, int(kid), int(part), len(p), p.reshape(-1), out))
        return out

    def sigma_coeff(self):
        c = self.deg * self.deg
        ct, cs = np.zeros(self.numberOfNodes), np.zeros(self.numberOfNodes)
        _lib.check(self.L.aniso_sigma_coeff(self.h, ct, cs))
        return ct.reshape(-1, c), cs.reshape(-1, c)

    def tree_stats(self):
        o = np.zeros(9, np.int64)
        _lib.check(self.L.aniso_tree_stats(self.h, o))
        return dict(zip(("nodes", "leaves", "max_level", "sumU", "sumV", "sumW", "sumX", "p2p_pairs", "max_leaf"),
                        (int(v) for v in o)))

    def tree_nodes(self):
        n = self.L.aniso_tree_nodes(self.h, 0, np.zeros(4), np.zeros(4, np.int32))
        geom, info = np.zeros(4 * n), np.zeros(4 * n, np.int32)
        self.L.aniso_tree_nodes(self.h, n, geom, info)
        return geom.reshape(n, 4), info.reshape(n, 4)

    def node_lists(self, node):
        ids, kinds = np.zeros(512, np.int32), np.zeros(512, np.int32)
        c = _lib.check(self.L.aniso_node_lists(self.h, int(node), 512, ids, kinds))
        return ids[:c].copy(), kinds[:c].copy()

    def geometry_tables(self):
        c = self.deg * self.deg
        r = 16 * c
        o = dict(interpolate=np.zeros(c * c), nearMapping=np.zeros(r * c), legendreNorms=np.zeros(c), rqx=np.zeros(r),
                 rqy=np.zeros(r), rw=np.zeros(r))
        _lib.check(self.L.aniso_geometry_tables(self.h, o["interpolate"], o["nearMapping"], o["legendreNorms"],
                                                o["rqx"], o["rqy"], o["rw"]))
        o["interpolate"] = o["interpolate"].reshape(c, c)
        o["nearMapping"] = o["nearMapping"].reshape(r, c)
        return o

    def sing_tables(self):
        c, m = self.deg * self.deg, 8 * self.sr * self.sr
        sx, sy, sw = np.zeros(c * m), np.zeros(c * m), np.zeros(c * m)
        _lib.check(self.L.aniso_sing_tables(self.h, sx, sy, sw))
        return sx.reshape(c, m), sy.reshape(c, m), sw.reshape(c, m)

    def slot_permutation(self):
        p = np.zeros(self.numberOfNodes, np.int32)
        _lib.check(self.L.aniso_slot_permutation(self.h, p))
        return p

    def partition(self, nranks, rank):
        b, e, nl = C.c_int(), C.c_int(), C.c_int()
        _lib.check(self.L.aniso_partition(self.h, nranks, rank, C.byref(b), C.byref(e), C.byref(nl)))
        return b.value, e.value, nl.value

    def owned_range(self):
        b, e = C.c_int(), C.c_int()
        _lib.check(self.L.aniso_owned_range(self.h, C.byref(b), C.byref(e)))
        return b.value, e.value

    def launch_count(self):
        return int(self.L.aniso_launch_count(self.h))

    def apply_bytes(self):
        o = np.zeros(3, np.int64)
        _lib.check(self.L.aniso_apply_bytes(self.h, o))
        return dict(p2p=int(o[0]), m2l=int(o[1]), stencil=int(o[2]))

    def cache_counts(self):
        o = np.zeros(2, np.int64)
        _lib.check(self.L.aniso_cache_counts(self.h, o))
        return dict(p2p=int(o[0]), m2l=int(o[1]))

    def cache_times(self):
        o = np.zeros(3)
        _lib.check(self.L.aniso_cache_times(self.h, o))
        return dict(p2p_fill_ms=float(o[0]), m2l_fill_ms=float(o[1]), stencil_ms=float(o[2]))

    def time_forward(self, reps):
        """average device milliseconds of one application of the GMRES operator (CUDA events on the handle's stream)"""
        ms = C.c_double()
        _lib.check(self.L.aniso_time_forward(self.h, int(reps), C.byref(ms)))
        return ms.value

    def config(self):
        return self.L.aniso_get_config(self.h).decode()

    def time_kernel(self, kid, which, reps):
        ms = C.c_double()
        _lib.check(self.L.aniso_time_kernel(self.h, int(kid), int(which), int(reps), C.byref(ms)))
        return ms.value

    def pinned_empty(self, n):
        """float64 array of n entries in page-locked host memory (aniso_host_alloc): vectors passed to mapping()/gmres()
        from it are DMA'd directly instead of being staged.  Freed with the handle."""
        import ctypes as C
        p = C.c_void_p()
        _lib.check(self.L.aniso_host_alloc(self.h, 8 * n, C.byref(p)))
        self._pinned.append(p)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(n,))

    def synchronize(self):
        _lib.check(self.L.aniso_synchronize(self.h))


class BBFMMKernel:
    """bbfmm::kernel (bbfmm.h:451-592) for a caller-supplied eval(sx, sy, tx, ty) -> float: the black-box interface itself.
        k = BBFMMKernel(); k.eval = f; k.initialize(np, x, y, charge, rank, maxLevel); pot = k.runCache() / k.runFast() / k.run()
    eval runs on the host (the reference's std::function does too): the operators are tabulated once on the host and applied
    on the device.  Sources and targets are one point set.  A Python eval is called from one thread (the GIL); a C callback
    passed as `eval_c` (an _lib.EVAL_FN) is called from all host threads."""

    def __init__(self):
        self.L = _lib.load()
        self.h = None
        self.eval = None
        self.eval_c = None
        self.nChebyshev = self.rank = 0
        self.chargeTree = None
        self._cached = False
        self._key = None

    def initialize(self, nChebyshev, x, y, charge, rank, maxLevel):
        x, y = _vec(x), _vec(y, None)
        if x.size != y.size:
            raise ValueError("x and y differ in length")
        key = (int(nChebyshev), int(rank), int(maxLevel), x.tobytes(), y.tobytes())
        if key != self._key:  # the reference re-initialises before every run (KernelFactory.cpp:356, 380): same tree, kept
            self.delete()
            h = C.c_void_p()
            _lib.check(self.L.aniso_fmm_create(int(nChebyshev), x.size, x, y, int(rank), int(maxLevel), C.byref(h)))
            self.h, self._key, self._cached = h, key, False
        self.nChebyshev, self.rank, self.n = int(nChebyshev), int(nChebyshev) ** 2, x.size
        self.chargeTree = _vec(charge, x.size, "charge").copy()

    def _tabulate(self):
        if self.eval_c is not None:
            cb, threads = self.eval_c, 0
        else:
            f = self.eval
            cb, threads = _lib.EVAL_FN(lambda ctx, sx, sy, tx, ty: float(f(sx, sy, tx, ty))), 1
        _lib.check(self.L.aniso_fmm_cache(self.h, cb, None, threads))
        self._cached = True

    def _apply(self):
        out = np.zeros(self.n)
        _lib.check(self.L.aniso_fmm_apply(self.h, self.chargeTree, out))
        return out

    def run(self):
        """kernel::run (bbfmm.h:507-532): evaluates the kernel during the pass; nothing is kept"""
        self._tabulate()
        out = self._apply()
        self._cached = False
        return out

    def runCache(self):
        """kernel::runCache (bbfmm.h:535-562): evaluates every entry, keeps it, applies"""
        self._tabulate()
        return self._apply()

    def runFast(self):
        """kernel::runFast (bbfmm.h:565-592): applies the cached operators"""
        if not self._cached:
            raise _lib.AnisoError(-4, "runCache must precede runFast")
        return self._apply()

    def tree_stats(self):
        o = np.zeros(9, np.int64)
        _lib.check(self.L.aniso_tree_stats(self.h, o))
        return o

    def delete(self):
        if self.h:
            self.L.aniso_destroy(self.h)
            self.h = None
            self._key = None

    def __del__(self):
        try:
            self.delete()
        except Exception:
            pass


def gl_table(deg):
    x, w = np.zeros(max(deg, 0)), np.zeros(max(deg, 0))
    _lib.check(_lib.load().aniso_gl_table(int(deg), x, w))
    return x, w


class _Session:
    """Session<Aniso> of mexplus (dispatch.h:183-275): integer handles into a static table."""
    table = {}
    next_id = 1


def AnisoWrapper(op, *args):
    """String-dispatched gateway with the six ops of AnisoWrapper.cpp:10-136."""
    if op == "new":
        if len(args) != 7:
            raise TypeError("AnisoWrapper('new', ds, qr, ks, as, sr, fn, fm)")
        hid = _Session.next_id
        _Session.next_id += 1
        _Session.table[hid] = Aniso(*args)
        return hid
    if op not in ("delete", "getNodes", "setCoeff", "cache", "mapping"):
        raise KeyError("Invalid operation: %s" % op)  # mexplus dispatch.h:302-317
    if not args or args[0] not in _Session.table:
        raise KeyError("Invalid id %r" % (args[0] if args else None))  # dispatch.h:207-215
    a = _Session.table[args[0]]
    if op == "delete":
        a.delete()
        del _Session.table[args[0]]
        return None
    if op == "getNodes":
        return a.getNodes()
    if op == "setCoeff":
        return a.setCoeff(args[1], args[2])
    if op == "cache":
        return a.cache(int(args[1]))
    return a.mapping(args[1], int(args[2])).reshape(-1, 1)
