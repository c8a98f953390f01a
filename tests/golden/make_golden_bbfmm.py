"""Generates tests/golden/bbfmm_points.npz by RUNNING THE COMPILED, UNMODIFIED REFERENCE ENGINE bbfmm::kernel
(oracle/_ref/libaniso_ref.so, driver ref_bbfmm_run in oracle/ref_capi.cpp) with kernels that are not the two families of
makeKernels, on an arbitrary (clustered, non-grid) point set.  Build container only:
    python tests/golden/make_golden_bbfmm.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cases  # noqa: E402
from oracle import refapi  # noqa: E402

# name -> (n, np, leaf capacity `rank`, maxLevel)
SETS = {"p1500n4": (1500, 4, 16, 20), "p900n6": (900, 6, 20, 20), "p700cap": (700, 3, 9, 2)}


def points(n):
    u = cases.splitmix64(4242, 2 * n)
    return u[:n] * 0.9 + 0.05 * u[n:] ** 2, u[n:] ** 1.5  # clustered towards y = 0: a non-uniform tree with W / X lists


out = {}
for name, (n, np_, rank, ml) in SETS.items():
    x, y = points(n)
    q = 2 * cases.splitmix64(9, n) - 1
    out[name + "_cfg"] = np.array([n, np_, rank, ml], np.int64)
    out[name + "_x"], out[name + "_y"], out[name + "_q"] = x, y, q
    for kind in (0, 1, 2):
        pot, stats = refapi.bbfmm_run(np_, x, y, rank, ml, kind, 2, q)  # runCache, then initialize + runFast
        pot_run, _ = refapi.bbfmm_run(np_, x, y, rank, ml, kind, 0, q)  # run (uncached)
        assert np.array_equal(pot, pot_run)
        out["%s_pot%d" % (name, kind)] = pot
    out[name + "_stats"] = stats
    print(name, "tree stats", stats, flush=True)
np.savez_compressed(os.path.join(HERE, "bbfmm_points.npz"), **out)

# The drop-in proof: tests/data/bbfmm_blackbox.cpp compiled against the reference's own bbfmm.h (make -C oracle bbfmm ->
# oracle/_ref/bbfmm_ref) and run here; the same source compiled against include/compat runs on the B200 (tests/test_dropin_gpu.py).
import subprocess  # noqa: E402
import tempfile  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(HERE))
n, np_, rank, ml = 1200, 4, 16, 20
u = cases.splitmix64(777, 2 * n)
x, y, q = u[:n] * 0.8 + 0.1, u[n:] ** 1.3, 2 * cases.splitmix64(5, n) - 1
with tempfile.NamedTemporaryFile("w", suffix=".txt", delete=False) as f:
    f.write("%d %d %d %d\n" % (n, np_, rank, ml) + "".join("%.17g %.17g %.17g\n" % t for t in zip(x, y, q)))
res = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "bbfmm_ref"), f.name], capture_output=True, text=True, check=True)
pot = np.array([float(t) for t in res.stdout.split()])
assert pot.size == 2 * n
np.savez_compressed(os.path.join(HERE, "bbfmm_blackbox.npz"), cfg=np.array([n, np_, rank, ml], np.int64), x=x, y=y, q=q, pot=pot)
print("bbfmm_blackbox: reference binary, %d potentials" % pot.size)
