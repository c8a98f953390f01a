"""Profiling driver (run under gpurun/ncu): one cache phase and a few matvecs at a BASELINE size.
  python scripts/prof_matvec.py [sz] [np] [reps] [id]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso

sz = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
np_ = int(sys.argv[2]) if len(sys.argv) > 2 else 4
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
kid = int(sys.argv[4]) if len(sys.argv) > 4 else 0
a = Aniso(sz, 1, max(1, kid + 1), 0.9, 10, np_, 20)
x, y, w = a.nodes()
ss, st = cases.sigma_fields("layers", x, y, sz, 1)
a.setCoeff(ss, st)
a.cache(kid)
f = cases.gaussian_source(x, y)
for _ in range(reps):
    out = a.mapping(f, kid)
print("sz", sz, "np", np_, "id", kid, "||out||", np.linalg.norm(out), "cache", a.cache_times(), "launches", a.launch_count())
a.delete()
