"""Development probe (run under gpurun): timing of the matvec pieces at a BASELINE size."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso

for spec in (sys.argv[1:] or ["1024,4"]):
    sz, np_ = (int(v) for v in spec.split(","))
    a = Aniso(sz, 1, 1, 0.9, 10, np_, 20)
    x, y, w = a.nodes(); ss, st_ = cases.sigma_fields("layers", x, y, sz, 1)
    a.setCoeff(ss, st_); a.cache(0)
    f = cases.gaussian_source(x, y)
    out = a.mapping(f, 0)
    ab = a.apply_bytes()
    m2l, p2p = a.time_kernel(0, 0, 20), a.time_kernel(0, 1, 20)
    print("sz", sz, "np", np_, "cache", a.cache_times(), "m2l %.4f ms %.0f GB/s  p2p %.4f ms %.0f GB/s" % (m2l, ab["m2l"] / m2l / 1e6, p2p, ab["p2p"] / p2p / 1e6), "||out||", np.linalg.norm(out))
    t0 = time.time(); xs, rc, it, hist = a.gmres(out); t1 = time.time()
    print("   gmres rc", rc, "iters", it, "%.3fs" % (t1 - t0), "final", hist[-1])
    a.delete()
