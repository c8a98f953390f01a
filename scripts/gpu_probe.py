"""Development probe (run under gpurun): stage-wise errors vs the golden fixtures and first timings."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso

def make(name):
    sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
    a = Aniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    a.setCoeff(ss, st)
    return a, x, y, w

for name in (sys.argv[1:] or ["a16", "b8d3", "h32", "nu6d5", "ml12d2", "c2", "c1"]):
    if name.startswith("big"):
        continue
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))
    a, x, y, w = make(name)
    st = int(g["stride"])
    li = a.line_integral(g["li_pts"])
    print(name, "li max rel", np.max(np.abs(li - g["li"]) / np.abs(g["li"])))
    f = {"f1": cases.tilted_source(x, y), "f2": cases.random_source(a.numberOfNodes, 1)}
    for kid in g["ids"]:
        kid = int(kid)
        for part in (0, 1):
            k = a.kernel_eval(kid, 1 if part == 0 else 2, g["li_pts"])
            i = list(g["ids"]).index(kid)
            print("   kern id", kid, "part", part, np.max(np.abs(k - g["kern"][i, part]) / (np.abs(g["kern"][i, part]) + 1e-300)))
        t0 = time.time(); a.cache(kid); t1 = time.time()
        for tag in ("f1", "f2"):
            s = a.mapping_stages(f[tag], kid)
            s["fused"] = a.mapping(f[tag], kid)
            errs = {}
            for k in s:
                key = "%s_id%d_%s" % (tag, kid, "out" if k == "fused" else k)
                errs[k] = "%.1e" % (np.linalg.norm(s[k][::st] - g[key]) * np.sqrt(st) / float(g[key + "_norm"]))
            print("   id", kid, tag, "cache %.3fs" % (t1 - t0), errs)
    b = a.mapping(cases.gaussian_source(x, y), 0)
    t0 = time.time(); xs, rc, it, hist = a.gmres(b); t1 = time.time()
    print("   gmres rc", rc, "iters", it, "ref", int(g["gm_iters"]), "x err", np.linalg.norm(xs[::st] - g["gm_x"]) / np.linalg.norm(g["gm_x"]), "%.3fs" % (t1 - t0), "final", hist[-1])
    a.delete()

if "big" in "".join(sys.argv[1:]) or len(sys.argv) == 1:
    for (sz, np_) in [(512, 4), (1024, 4)]:
        t0 = time.time(); a = Aniso(sz, 1, 1, 0.9, 10, np_, 20); t1 = time.time()
        x, y, w = a.nodes(); ss, st_ = cases.sigma_fields("layers", x, y, sz, 1)
        a.setCoeff(ss, st_); t2 = time.time()
        a.cache(0); t3 = time.time()
        f = cases.gaussian_source(x, y)
        out = a.mapping(f, 0); t4 = time.time()
        for _ in range(5): out = a.mapping(f, 0)
        t5 = time.time()
        print("big sz", sz, "np", np_, "create %.2fs setcoeff %.2fs cache %.3fs first map %.3fs map(host) %.2f ms" % (t1 - t0, t2 - t1, t3 - t2, t4 - t3, (t5 - t4) / 5 * 1e3))
        print("   bytes", a.apply_bytes(), "m2l ms", a.time_kernel(0, 0, 20), "p2p ms", a.time_kernel(0, 1, 20))
        b = out
        t0 = time.time(); xs, rc, it, hist = a.gmres(b); t1 = time.time()
        print("   gmres rc", rc, "iters", it, "%.3fs" % (t1 - t0), "final", hist[-1], "launches", a.launch_count())
        a.delete()
