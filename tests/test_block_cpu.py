"""CPU tests of the checker for the multi-mode block operator (aniso.m:121-172): the restated aniso.m loops
and gmres.cpp (oracle/block_oracle.py) over the plain-C oracle `mapping`, against golden fixtures produced with the
compiled reference `mapping` under the same loops (tests/golden/make_golden_block.py)."""
import numpy as np
import pytest

import cases
from oracle import block_oracle, oracleapi


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def make(name):
    sz, deg, K, g, sr, np_, ml, kind = cases.BLOCK_CASES[name]
    o = oracleapi.OracleAniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = o.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    o.set_coeff(ss, st)
    return o, K, g, ss, x.size


def test_chi_weights():
    # aniso.m:148: chi = (g^id - g^N)/(1 - g^N): 1 for mode 0, decreasing, 0 would be reached at id = N
    for g, K in ((0.5, 2), (0.9, 5)):
        c = [block_oracle.chi(g, a, K) for a in range(K)]
        assert c[0] == 1.0 and all(c[i] > c[i + 1] > 0 for i in range(K - 1))


@pytest.mark.parametrize("name", ["a16", "k3h16"])
def test_block_operator_oracle_vs_reference_golden(golden, name):
    g = golden(name + "_block")
    o, K, gg, ss, n = make(name)
    assert rel(block_oracle.forward(o, g["u"], K, n), g["forward"]) < 1e-11
    assert rel(block_oracle.mforward(o, g["u"], K, n, gg, ss), g["mforward"]) < 1e-11


def test_restated_gmres_matches_c_oracle_gmres():
    """the numpy restatement of gmres.cpp is pinned on the C oracle's GMRES (itself pinned on the compiled reference) with
    the main.cpp operator"""
    sz, deg, K, g, sr, np_, ml, kind = cases.CASES["ml12d2"]
    o = oracleapi.OracleAniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = o.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    o.set_coeff(ss, st)
    b = o.mapping(cases.gaussian_source(x, y), 0)
    xo, rco, ito, histo = o.gmres(b)
    xs, rc, it, hist = block_oracle.gmres(lambda v: v - o.mapping(ss * v, 0), b, m=80, max_iter=400, tol=1e-12)
    assert (rc, it) == (rco, ito)
    assert rel(xs, xo) < 1e-10


def test_block_gmres_oracle_vs_reference_golden(golden):
    g = golden("a16_block")
    o, K, gg, ss, n = make("a16")
    xs, rc, it, hist = block_oracle.gmres(lambda v: v - block_oracle.mforward(o, v, K, n, gg, ss), g["forward"], m=400, max_iter=400, tol=1e-11)
    assert rc == int(g["gm_rc"]) == 0 and abs(it - int(g["gm_iters"])) <= 1
    assert rel(xs, g["gm_x"]) < 1e-8
