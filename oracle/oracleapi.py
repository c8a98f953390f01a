"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/liboracle.so (the plain-C restatement,
oracle/aniso_oracle.c).  Same Python surface as oracle/refapi.RefAniso so tests can drive either.
Never imported by the product package (aniso_b200/)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_lp = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_lib = None


def lib_path():
    return os.path.join(_HERE, "liboracle.so")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])


def _load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(lib_path()):
        build()
    L = C.CDLL(lib_path())
    L.oracle_create.restype = C.c_void_p
    L.oracle_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]
    L.oracle_destroy.argtypes = [C.c_void_p]
    L.oracle_num_nodes.argtypes = [C.c_void_p]
    L.oracle_kernel_size.argtypes = [C.c_void_p]
    L.oracle_get_nodes.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.oracle_set_coeff.argtypes = [C.c_void_p, _dp, _dp]
    L.oracle_cache.argtypes = [C.c_void_p, C.c_int]
    L.oracle_mapping.argtypes = [C.c_void_p, _dp, C.c_int, _dp]
    L.oracle_mapping_stages.argtypes = [C.c_void_p, _dp, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
    L.oracle_gmres.argtypes = [C.c_void_p, _dp, _dp, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), _dp, C.c_int]
    L.oracle_line_integral.restype = C.c_double
    L.oracle_line_integral.argtypes = [C.c_void_p] + [C.c_double] * 4
    L.oracle_evaluate.restype = C.c_double
    L.oracle_evaluate.argtypes = [C.c_void_p, C.c_double, C.c_double]
    L.oracle_kernel_eval.restype = C.c_double
    L.oracle_kernel_eval.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_double] * 4
    L.oracle_tree_stats.argtypes = [C.c_void_p, _lp]
    L.oracle_tree_nodes.argtypes = [C.c_void_p, C.c_int, _dp, _ip]
    L.oracle_node_lists.argtypes = [C.c_void_p, C.c_int, C.c_int, _ip, _ip]
    L.oracle_gl_table.argtypes = [C.c_int, _dp, _dp]
    L.oracle_geometry_tables.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp, _dp]
    L.oracle_sing_tables.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.oracle_sigma_coeff.argtypes = [C.c_void_p, _dp, _dp]
    L.oracle_fmm.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
    _lib = L
    return L


def gl_table(deg):
    x, w = np.zeros(deg), np.zeros(deg)
    _load().oracle_gl_table(deg, x, w)
    return x, w


class OracleAniso:
    def __init__(self, sz, deg, K, g, sr, np_, max_level):
        self.L = _load()
        self.h = self.L.oracle_create(sz, deg, K, g, sr, np_, max_level)
        self.sz, self.deg, self.K, self.sr, self.np = sz, deg, K, sr, np_
        self.n = self.L.oracle_num_nodes(self.h)
        self.kernel_size = self.L.oracle_kernel_size(self.h)

    def close(self):
        if self.h:
            self.L.oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def num_threads(self):
        return 1

    def nodes(self):
        x, y, w = np.zeros(self.n), np.zeros(self.n), np.zeros(self.n)
        self.L.oracle_get_nodes(self.h, x, y, w)
        return x, y, w

    def set_coeff(self, sigma_s, sigma_t):
        self.L.oracle_set_coeff(self.h, np.ascontiguousarray(sigma_s, np.float64),
                                np.ascontiguousarray(sigma_t, np.float64))

    def cache(self, kid):
        self.L.oracle_cache(self.h, kid)

    def mapping(self, charge, kid):
        out = np.zeros(self.n)
        self.L.oracle_mapping(self.h, np.ascontiguousarray(charge, np.float64), kid, out)
        return out

    def mapping_stages(self, charge, kid):
        o = {k: np.zeros(self.n) for k in ("imag_fmm", "real_fmm", "k7", "k8", "k9", "out")}
        self.L.oracle_mapping_stages(self.h, np.ascontiguousarray(charge, np.float64), kid, o["imag_fmm"],
                                     o["real_fmm"], o["k7"], o["k8"], o["k9"], o["out"])
        return o

    def fmm(self, kid, part, charge):
        out = np.zeros(self.n)
        self.L.oracle_fmm(self.h, kid, part, np.ascontiguousarray(charge, np.float64), out)
        return out

    def gmres(self, b, x0=None, m=80, max_iter=400, tol=1e-12):
        x = np.zeros(self.n) if x0 is None else np.array(x0, np.float64)
        it = C.c_int(0)
        hist = np.zeros(max_iter + 2)
        rc = self.L.oracle_gmres(self.h, np.ascontiguousarray(b, np.float64), x, m, max_iter, tol, C.byref(it), hist,
                                 hist.size)
        return x, rc, it.value, hist[: it.value + 1]

    def line_integral(self, x0, y0, x1, y1):
        return self.L.oracle_line_integral(self.h, x0, y0, x1, y1)

    def evaluate(self, x, y):
        return self.L.oracle_evaluate(self.h, x, y)

    def kernel_eval(self, kid, part, ax, ay, bx, by):
        return self.L.oracle_kernel_eval(self.h, kid, part, ax, ay, bx, by)

    def tree_stats(self, kid=0):
        o = np.zeros(9, np.int64)
        self.L.oracle_tree_stats(self.h, o)
        return dict(zip(("nodes", "leaves", "max_level", "sumU", "sumV", "sumW", "sumX", "p2p_pairs",
                         "max_leaf"), (int(v) for v in o)))

    def tree_nodes(self, kid=0):
        cap = 1 << 20
        geom = np.zeros(4 * cap)
        info = np.zeros(4 * cap, np.int32)
        n = self.L.oracle_tree_nodes(self.h, cap, geom, info)
        return geom[: 4 * n].reshape(n, 4), info[: 4 * n].reshape(n, 4)

    def node_lists(self, node, kid=0):
        ids = np.zeros(256, np.int32)
        kinds = np.zeros(256, np.int32)
        c = self.L.oracle_node_lists(self.h, node, 256, ids, kinds)
        return ids[:c].copy(), kinds[:c].copy()

    def geometry_tables(self):
        c = self.deg * self.deg
        r = 16 * c
        o = dict(interpolate=np.zeros(c * c), nearMapping=np.zeros(r * c), legendreNorms=np.zeros(c),
                 rqx=np.zeros(r), rqy=np.zeros(r), rw=np.zeros(r))
        self.L.oracle_geometry_tables(self.h, o["interpolate"], o["nearMapping"], o["legendreNorms"], o["rqx"],
                                      o["rqy"], o["rw"])
        o["interpolate"] = o["interpolate"].reshape(c, c)
        o["nearMapping"] = o["nearMapping"].reshape(r, c)
        return o

    def sing_tables(self):
        c = self.deg * self.deg
        m = 8 * self.sr * self.sr
        sx, sy, sw = np.zeros(c * m), np.zeros(c * m), np.zeros(c * m)
        self.L.oracle_sing_tables(self.h, sx, sy, sw)
        return sx.reshape(c, m), sy.reshape(c, m), sw.reshape(c, m)

    def sigma_coeff(self):
        c = self.deg * self.deg
        ct, cs = np.zeros(self.n), np.zeros(self.n)
        self.L.oracle_sigma_coeff(self.h, ct, cs)
        return ct.reshape(-1, c), cs.reshape(-1, c)
