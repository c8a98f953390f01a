"""Generates tests/golden/*_block.npz: the multi-mode block operator (aniso.m:121-172) evaluated with the
COMPILED, UNMODIFIED REFERENCE `mapping` (oracle/_ref/libaniso_ref.so) under the restated aniso.m loops and the restated
gmres.cpp (oracle/block_oracle.py).  MATLAB cannot run here, so this is the strongest pin available for that layer.
Run in the build container only:   python tests/golden/make_golden_block.py [case ...]"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cases  # noqa: E402
from oracle import block_oracle  # noqa: E402
from oracle.refapi import RefAniso  # noqa: E402


def block_source(x, y, K):
    """demo.m:24-29 puts the Gaussian source in mode 0 and zeros elsewhere; a tilted copy in the other modes keeps every
    (mode, kernel id) pair of the loops active."""
    return np.concatenate([cases.gaussian_source(x, y) * (1.0 + 0.5 * a * (x - y)) for a in range(K)])


def run(name):
    sz, deg, K, g, sr, np_, ml, kind = cases.BLOCK_CASES[name]
    r = RefAniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = r.nodes()
    n = x.size
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    r.set_coeff(ss, st)
    for k in range(2 * K - 1):
        r.cache(k)
    u = block_source(x, y, K)
    fw = block_oracle.forward(r, u, K, n)
    mf = block_oracle.mforward(r, u, K, n, g, ss)
    xs, rc, it, hist = block_oracle.gmres(lambda v: v - block_oracle.mforward(r, v, K, n, g, ss), fw, m=400, max_iter=400, tol=1e-11)
    np.savez_compressed(os.path.join(HERE, name + "_block.npz"), cfg=np.array([sz, deg, K, sr, np_, ml], np.int64), g=np.array(g),
                        u=u, forward=fw, mforward=mf, gm_x=xs, gm_rc=np.array(rc), gm_iters=np.array(it), gm_hist=hist)
    print(name, "n", n, "K", K, "|forward|", np.linalg.norm(fw), "|mforward|", np.linalg.norm(mf), "gmres", rc, it, hist[-1])


if __name__ == "__main__":
    for nm in (sys.argv[1:] or list(cases.BLOCK_CASES)):
        run(nm)
