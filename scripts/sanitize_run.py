"""Small end-to-end run that touches every kernel family, for compute-sanitizer (run under gpurun):
    compute-sanitizer --tool memcheck --error-exitcode 3 python scripts/sanitize_run.py"""
import math, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso, BBFMMKernel

for name in ("a16", "b8d3", "nu6d5", "h32"):      # 16x16 symmetric, deg 3 cell walk, ragged non-symmetric, 36x36 blocks
    sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
    a = Aniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = a.nodes()
    a.setCoeff(*cases.sigma_fields(kind, x, y, sz, deg))
    f = cases.tilted_source(x, y)
    for kid in range(a.kernelSize):
        a.cache(kid)
        out = a.mapping(f, kid)
    a.mapping_stages(f, 0)
    xs, rc, it, hist = a.gmres(a.mapping(cases.gaussian_source(x, y), 0))
    u = cases.random_source((a.kernelSize // 2 + 1) * x.size, 5)
    a.mforward(u); a.forward(u)
    a.line_integral(cases.splitmix64(7, 400).reshape(100, 4)); a.line_integral_noalloc(cases.splitmix64(8, 400).reshape(100, 4))
    a.set_option("p2p", "onfly"); a.cache(0); a.mapping(f, 0)
    a.setCoeff(*cases.sigma_fields(kind, x, y, sz, deg)); a.mapping(f, 0)
    print(name, "ok: gmres", rc, it, "||out||", float(np.linalg.norm(out)), flush=True)
    a.delete()
n = 600
v = cases.splitmix64(3, 2 * n)
k = BBFMMKernel()
k.eval = lambda sx, sy, tx, ty: 0.0 if (sx == tx and sy == ty) else 0.5 * math.log((sx - tx) ** 2 + (sy - ty) ** 2)
k.initialize(4, v[:n], v[n:] ** 1.5, 2 * cases.splitmix64(9, n) - 1, 16, 20)
print("bbfmm ok", float(np.linalg.norm(k.runCache())), flush=True)
k.delete()
