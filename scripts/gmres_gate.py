"""GMRES gate evidence (run under gpurun): iteration counts of every orthogonalisation form of the device GMRES next to the
compiled reference's own count (tests/golden), on the single-mode fixtures (main.cpp:141: GMRES(80, 400, 1e-12)) and the
multi-mode block fixtures (aniso.m:165: restart 400, tol 1e-11).
    python scripts/gmres_gate.py > profiles/r02_gmres_gate.txt
Forms (ANISO_GMRES_ORTHO / ANISO_GMRES_SYNC): default = CGS2, two reductions, pipelined; cgs2_3red; cgs2_unfused; mgs = the
reference's modified Gram-Schmidt (gmres.cpp:115-119)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso

FORMS = [("default", {}), ("default/sync", {"ANISO_GMRES_SYNC": "1"}), ("cgs2_3red", {"ANISO_GMRES_ORTHO": "cgs2_3red"}),
         ("cgs2_unfused", {"ANISO_GMRES_ORTHO": "cgs2_unfused"}), ("mgs", {"ANISO_GMRES_ORTHO": "mgs"})]


def golden(name):
    return dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))


def with_env(env, fn):
    for k in ("ANISO_GMRES_SYNC", "ANISO_GMRES_ORTHO"):
        os.environ.pop(k, None)
    os.environ.update(env)
    try:
        return fn()
    finally:
        for k in env:
            os.environ.pop(k, None)


print("case | reference iterations (last printed residual) | " + " | ".join(f for f, _ in FORMS) + "   [iterations, residual at the reference's last step, rel. L2 of the solution vs the reference]")
for name in ["a16", "b8d3", "nu6d5", "ml12d2", "h32", "c2", "c1", "c3", "c4n4", "c4n8", "c5"]:
    g = golden(name)
    if "gm_iters" not in g:
        continue
    sz, deg, K, gg, sr, np_, ml, kind = cases.CASES[name]
    a = Aniso(sz, deg, K, gg, sr, np_, ml)
    x, y, w = a.nodes()
    a.setCoeff(*cases.sigma_fields(kind, x, y, sz, deg))
    a.cache(0)
    stride = int(g["stride"])
    b = g["gm_b"] if stride == 1 else a.mapping(cases.gaussian_source(x, y), 0)
    ref_it = int(g["gm_iters"])
    cells = []
    for form, env in FORMS:
        xs, rc, it, hist = with_env(env, lambda: a.gmres(b, m=80, max_iter=400, tol=1e-12))
        err = np.linalg.norm(xs[::stride] - g["gm_x"]) / np.linalg.norm(g["gm_x"])
        cells.append("%d, %.2e, %.1e" % (it, hist[min(it, ref_it) - 1] if min(it, ref_it) >= 1 else hist[0], err))
    print("%s | %d (%.3e) | %s" % (name, ref_it, float(g["gm_hist"][-1]), " | ".join(cells)), flush=True)
    a.delete()
for name in ["a16", "k3h16", "b8d3"]:
    g = golden(name + "_block")
    sz, deg, K, gg, sr, np_, ml, kind = cases.BLOCK_CASES[name]
    a = Aniso(sz, deg, K, gg, sr, np_, ml)
    x, y, w = a.nodes()
    a.setCoeff(*cases.sigma_fields(kind, x, y, sz, deg))
    ref_it = int(g["gm_iters"])
    cells = []
    for form, env in FORMS:
        xs, rc, it, hist = with_env(env, lambda: a.solve(g["u"]))
        err = np.linalg.norm(xs - g["gm_x"]) / np.linalg.norm(g["gm_x"])
        cells.append("%d, %.2e, %.1e" % (it, hist[min(it, ref_it) - 1], err))
    print("%s block (K %d) | %d (%.3e) | %s" % (name, K, ref_it, float(g["gm_hist"][-1]), " | ".join(cells)), flush=True)
    a.delete()
