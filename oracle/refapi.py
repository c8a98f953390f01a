"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libaniso_ref*.so (the UNMODIFIED
reference compiled by oracle/Makefile, wrapped by oracle/ref_capi.cpp).

Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
Never imported by the product package (aniso_b200/).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_lp = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def lib_path(fast=False):
    return os.path.join(_HERE, "_ref", "libaniso_ref_fast.so" if fast else "libaniso_ref.so")


def available(fast=False):
    return os.path.exists(lib_path(fast))


_libs = {}


def _load(fast):
    if fast in _libs:
        return _libs[fast]
    L = C.CDLL(lib_path(fast))
    L.ref_create.restype = C.c_void_p
    L.ref_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]
    L.ref_destroy.argtypes = [C.c_void_p]
    L.ref_num_nodes.argtypes = [C.c_void_p]
    L.ref_kernel_size.argtypes = [C.c_void_p]
    L.ref_get_nodes.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.ref_set_coeff.argtypes = [C.c_void_p, _dp, _dp]
    L.ref_cache.argtypes = [C.c_void_p, C.c_int, _dp]
    L.ref_mapping.argtypes = [C.c_void_p, _dp, C.c_int, _dp]
    L.ref_mapping_stages.argtypes = [C.c_void_p, _dp, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
    L.ref_gmres.argtypes = [C.c_void_p, _dp, _dp, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), _dp, C.c_int]
    L.ref_applies.restype = C.c_long
    L.ref_applies.argtypes = [C.c_void_p]
    L.ref_line_integral.restype = C.c_double
    L.ref_line_integral.argtypes = [C.c_void_p] + [C.c_double] * 4
    L.ref_evaluate.restype = C.c_double
    L.ref_evaluate.argtypes = [C.c_void_p, C.c_double, C.c_double]
    L.ref_kernel_eval.restype = C.c_double
    L.ref_kernel_eval.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_double] * 4
    L.ref_tree_stats.argtypes = [C.c_void_p, C.c_int, _lp]
    L.ref_tree_nodes.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _ip]
    L.ref_cache_block.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int),
                                  C.POINTER(C.c_int), _dp, C.c_int]
    L.ref_node_lists.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _ip, _ip]
    L.ref_node_points.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _ip]
    L.ref_gl_table.argtypes = [C.c_int, _dp, _dp]
    L.ref_geometry_tables.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp, _dp]
    L.ref_sing_tables.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.ref_sigma_coeff.argtypes = [C.c_void_p, _dp, _dp]
    L.ref_bbfmm_run.argtypes = [C.c_int, C.c_int, _dp, _dp, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp, _lp]
    _libs[fast] = L
    return L


def gl_table(deg, fast=False):
    x = np.zeros(deg)
    w = np.zeros(deg)
    _load(fast).ref_gl_table(deg, x, w)
    return x, w


def bbfmm_eval(kind, sx, sy, tx, ty):
    """the test kernels of ref_bbfmm_run (oracle/ref_capi.cpp), restated for the callbacks handed to the product"""
    dx, dy = sx - tx, sy - ty
    r2 = dx * dx + dy * dy
    if kind == 0:
        return 0.0 if r2 == 0.0 else 0.5 * np.log(r2)
    if kind == 1:
        return (dx + 0.3) / (r2 + 0.01)
    r = np.sqrt(r2)
    return np.exp(-3.0 * r) / (0.05 + r)


def bbfmm_run(np_, x, y, rank, max_level, kind, mode, charge, fast=False):
    """the reference's bbfmm::kernel (initialize + run / runCache / runFast) with a non-family eval on arbitrary points"""
    x, y, q = (np.ascontiguousarray(v, np.float64) for v in (x, y, charge))
    pot = np.zeros(x.size)
    stats = np.zeros(9, np.int64)
    _load(fast).ref_bbfmm_run(np_, x.size, x, y, rank, max_level, kind, mode, q, pot, stats)
    return pot, stats


class RefAniso:
    """The reference `Aniso` object driven through the op sequence of AnisoWrapper.cpp."""

    def __init__(self, sz, deg, K, g, sr, np_, max_level, fast=False):
        self.L = _load(fast)
        self.h = self.L.ref_create(sz, deg, K, g, sr, np_, max_level)
        self.sz, self.deg, self.K, self.sr, self.np = sz, deg, K, sr, np_
        self.n = self.L.ref_num_nodes(self.h)
        self.kernel_size = self.L.ref_kernel_size(self.h)

    def close(self):
        if self.h:
            self.L.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def num_threads(self):
        return self.L.ref_num_threads()

    def nodes(self):
        x, y, w = np.zeros(self.n), np.zeros(self.n), np.zeros(self.n)
        self.L.ref_get_nodes(self.h, x, y, w)
        return x, y, w

    def set_coeff(self, sigma_s, sigma_t):
        self.L.ref_set_coeff(self.h, np.ascontiguousarray(sigma_s, np.float64),
                             np.ascontiguousarray(sigma_t, np.float64))

    def cache(self, kid):
        t = np.zeros(4)
        self.L.ref_cache(self.h, kid, t)
        return t

    def mapping(self, charge, kid):
        out = np.zeros(self.n)
        self.L.ref_mapping(self.h, np.ascontiguousarray(charge, np.float64), kid, out)
        return out

    def mapping_stages(self, charge, kid):
        o = {k: np.zeros(self.n) for k in ("imag_fmm", "real_fmm", "k7", "k8", "k9", "out")}
        self.L.ref_mapping_stages(self.h, np.ascontiguousarray(charge, np.float64), kid, o["imag_fmm"],
                                  o["real_fmm"], o["k7"], o["k8"], o["k9"], o["out"])
        return o

    def gmres(self, b, x0=None, m=80, max_iter=400, tol=1e-12):
        x = np.zeros(self.n) if x0 is None else np.array(x0, np.float64)
        it = C.c_int(0)
        hist = np.zeros(max_iter + 2)
        rc = self.L.ref_gmres(self.h, np.ascontiguousarray(b, np.float64), x, m, max_iter, tol, C.byref(it), hist,
                              hist.size)
        return x, rc, it.value, hist[: it.value]

    def line_integral(self, x0, y0, x1, y1):
        return self.L.ref_line_integral(self.h, x0, y0, x1, y1)

    def evaluate(self, x, y):
        return self.L.ref_evaluate(self.h, x, y)

    def kernel_eval(self, kid, part, ax, ay, bx, by):
        return self.L.ref_kernel_eval(self.h, kid, part, ax, ay, bx, by)

    def tree_stats(self, kid=0):
        o = np.zeros(9, np.int64)
        self.L.ref_tree_stats(self.h, kid, o)
        return dict(zip(("nodes", "leaves", "max_level", "sumU", "sumV", "sumW", "sumX", "p2p_pairs",
                         "max_leaf"), (int(v) for v in o)))

    def tree_nodes(self, kid=0):
        cap = 1 << 22
        geom = np.zeros(4 * cap)
        info = np.zeros(4 * cap, np.int32)
        n = self.L.ref_tree_nodes(self.h, kid, cap, geom, info)
        return geom[: 4 * n].reshape(n, 4), info[: 4 * n].reshape(n, 4)

    def node_lists(self, node, kid=0):
        ids = np.zeros(256, np.int32)
        kinds = np.zeros(256, np.int32)
        c = self.L.ref_node_lists(self.h, kid, node, 256, ids, kinds)
        return ids[:c].copy(), kinds[:c].copy()

    def node_points(self, node, kid=0):
        cap = 1 << 16
        idx = np.zeros(cap, np.int32)
        c = self.L.ref_node_points(self.h, kid, node, cap, idx)
        return idx[:c].copy()

    def cache_block(self, kid, part, node, slot):
        r, c = C.c_int(0), C.c_int(0)
        buf = np.zeros(1 << 14)
        n = self.L.ref_cache_block(self.h, kid, part, node, slot, C.byref(r), C.byref(c), buf, buf.size)
        if n == 0:
            return None
        return buf[:n].reshape(c.value, r.value).T.copy()  # column-major -> [row, col]

    def geometry_tables(self):
        c = self.deg * self.deg
        r = 16 * c
        o = dict(interpolate=np.zeros(c * c), nearMapping=np.zeros(r * c), legendreNorms=np.zeros(c),
                 rqx=np.zeros(r), rqy=np.zeros(r), rw=np.zeros(r))
        self.L.ref_geometry_tables(self.h, o["interpolate"], o["nearMapping"], o["legendreNorms"], o["rqx"],
                                   o["rqy"], o["rw"])
        o["interpolate"] = o["interpolate"].reshape(c, c).T.copy()
        o["nearMapping"] = o["nearMapping"].reshape(c, r).T.copy()
        return o

    def sing_tables(self):
        c = self.deg * self.deg
        m = 8 * self.sr * self.sr
        sx, sy, sw = np.zeros(c * m), np.zeros(c * m), np.zeros(c * m)
        self.L.ref_sing_tables(self.h, sx, sy, sw)
        return sx.reshape(c, m), sy.reshape(c, m), sw.reshape(c, m)

    def sigma_coeff(self):
        c = self.deg * self.deg
        ct = np.zeros(self.sz * self.sz * c)
        cs = np.zeros(self.sz * self.sz * c)
        self.L.ref_sigma_coeff(self.h, ct, cs)
        return ct.reshape(-1, c), cs.reshape(-1, c)
