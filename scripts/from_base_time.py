"""Time of deriving the operators of kernel id 1 from id 0 (fill_from_base) at BASELINE C5; run under gpurun, once plain and
once with ANISO_FROM_BASE_GENERIC=1.  Prints the event times of the two fills and a checksum of mapping_1."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from aniso_b200 import Aniso
a = Aniso(1024, 1, 2, 0.9, 10, 4, 20)
x, y, w = a.nodes()
a.setCoeff(*cases.sigma_fields("layers", x, y, 1024, 1))
a.cache(0)
a.cache(1)
ct = a.cache_times()
out = a.mapping(cases.tilted_source(x, y), 1)
nbytes = 16 * (a.cache_counts()["p2p"] + a.cache_counts()["m2l"])
print("generic" if os.environ.get("ANISO_FROM_BASE_GENERIC") else "16x16 kernel", "id 1 from id 0: P2P %.3f ms, M2L %.3f ms, %.2f TB/s; ||mapping_1|| %.15e"
      % (ct["p2p_fill_ms"], ct["m2l_fill_ms"], nbytes / (ct["p2p_fill_ms"] + ct["m2l_fill_ms"]) / 1e9, np.linalg.norm(out)))
a.delete()
