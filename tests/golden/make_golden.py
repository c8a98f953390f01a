"""Generates tests/golden/*.npz by RUNNING THE COMPILED, UNMODIFIED REFERENCE (oracle/_ref/libaniso_ref.so,
built by oracle/Makefile from /root/reference).  Run in the build container only:
    python tests/golden/make_golden.py [case ...]
The fixtures pin the C oracle (oracle/aniso_oracle.c) and the CUDA path on machines where
/root/reference does not exist."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cases  # noqa: E402
from oracle.refapi import RefAniso, gl_table  # noqa: E402

FULL = {"a16", "b8d3", "h32", "nu6d5", "ml12d2"}  # small: store full vectors
IDS = {"c2": [0, 1, 4], "c1": [0], "c3": [0, 3]}
STRIDE = 61  # subsampling stride for the large cases


def gl():
    out = {}
    for d in range(1, 129):
        x, w = gl_table(d)
        out["x%d" % d] = x
        out["w%d" % d] = w
    np.savez_compressed(os.path.join(HERE, "gl_tables.npz"), **out)


def run(name):
    sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
    r = RefAniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = r.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    r.set_coeff(ss, st)
    full = name in FULL
    sel = slice(None) if full else slice(0, None, STRIDE)
    out = dict(cfg=np.array([sz, deg, K, sr, np_, ml], np.int64), g=np.array(g), stride=np.array(1 if full else STRIDE))
    if full:
        out.update(x=x, y=y, w=w)
        gt = r.geometry_tables()
        out.update({"geo_" + k: v for k, v in gt.items()})
        sx, sy, sw = r.sing_tables()
        out.update(singX=sx, singY=sy, singW=sw)
        ct, cs = r.sigma_coeff()
        out.update(coeff_t=ct, coeff_s=cs)
    else:
        out.update(x=x[sel], y=y[sel], w=w[sel])
    # line integrals / kernel samples at seeded point pairs strictly inside the unit square
    u = cases.splitmix64(77, 4 * 64).reshape(64, 4) * 0.98 + 0.01
    out["li_pts"] = u
    out["li"] = np.array([r.line_integral(*p) for p in u])
    out["sig_eval"] = np.array([r.evaluate(p[0], p[1]) for p in u])
    kid_list = IDS.get(name, list(range(2 * K - 1)))
    out["ids"] = np.array(kid_list, np.int64)
    out["kern"] = np.array([[[r.kernel_eval(k, part, *p) for p in u] for part in (0, 1)] for k in kid_list])
    f1 = cases.tilted_source(x, y)
    f2 = cases.random_source(r.n, 1)
    for kid in kid_list:
        r.cache(kid)
        for tag, f in (("f1", f1), ("f2", f2)):
            s = r.mapping_stages(f, kid)
            for k, v in s.items():
                out["%s_id%d_%s" % (tag, kid, k)] = v[sel]
                out["%s_id%d_%s_norm" % (tag, kid, k)] = np.array(np.linalg.norm(v))
    ts = r.tree_stats(kid_list[0])
    out["tree_stats"] = np.array([ts[k] for k in ("nodes", "leaves", "max_level", "sumU", "sumV", "sumW", "sumX",
                                                  "p2p_pairs", "max_leaf")], np.int64)
    if full:
        geom, info = r.tree_nodes(kid_list[0])
        out.update(tree_geom=geom, tree_info=info)
    # GMRES of main.cpp: b = mapping_0(gaussian), A x = x - mapping_0(sigma_s x), GMRES(80, 400, 1e-12)
    b = r.mapping(cases.gaussian_source(x, y), 0)
    xs, rc, it, hist = r.gmres(b, m=80, max_iter=400, tol=1e-12)
    out.update(gm_b=b[sel], gm_x=xs[sel], gm_x_norm=np.array(np.linalg.norm(xs)), gm_rc=np.array(rc),
               gm_iters=np.array(it), gm_hist=hist)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "n", r.n, "iters", it, "last printed resid", hist[-1], flush=True)


if __name__ == "__main__":
    names = sys.argv[1:] or ["gl"] + list(cases.CASES)
    for n in names:
        gl() if n == "gl" else run(n)
