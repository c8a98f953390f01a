"""Timing of the multi-mode block operator (run under gpurun): batched form vs the loop of K(2K-1) mappings.
  python scripts/block_time.py sz K [np]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso, _lib
import ctypes as C

sz, K = int(sys.argv[1]), int(sys.argv[2])
np_ = int(sys.argv[3]) if len(sys.argv) > 3 else 4
a = Aniso(sz, 1, K, 0.9, 10, np_, 20)
x, y, w = a.nodes(); n = x.size
ss, st = cases.sigma_fields("layers", x, y, sz, 1)
a.setCoeff(ss, st)
t0 = time.time()
for k in range(2 * K - 1):
    a.cache(k)
t_cache = time.time() - t0
L = a.L
du, dout = C.c_void_p(), C.c_void_p()
_lib.check(L.aniso_dev_alloc(a.h, 8 * K * n, C.byref(du))); _lib.check(L.aniso_dev_alloc(a.h, 8 * K * n, C.byref(dout)))
tmp = C.c_void_p(); _lib.check(L.aniso_dev_alloc(a.h, 8 * n, C.byref(tmp)))
u = np.tile(cases.gaussian_source(x, y), K)
for m in range(K):
    _lib.check(L.aniso_to_slots(a.h, u[m * n:(m + 1) * n], C.c_void_p(du.value + 8 * m * n)))
for env in ("", "1"):
    if env: os.environ["ANISO_BLOCK_LOOP"] = env
    else: os.environ.pop("ANISO_BLOCK_LOOP", None)
    for _ in range(3): _lib.check(L.aniso_block_forward_dev(a.h, du, 1, dout))
    a.synchronize(); l0 = a.launch_count(); t0 = time.perf_counter(); reps = 10
    for _ in range(reps): _lib.check(L.aniso_block_forward_dev(a.h, du, 1, dout))
    a.synchronize(); dt = (time.perf_counter() - t0) / reps
    print("sz", sz, "K", K, "np", np_, "loop" if env else "batched", "mforward %.3f ms" % (dt * 1e3), "launches", (a.launch_count() - l0) // reps, "cache all ids %.3f s" % t_cache)
a.delete()
