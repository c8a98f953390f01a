"""GPU parity of the multi-mode block operator (aniso.m:121-172 forward / mforward / solve) through the
C-ABI (aniso_block_forward, aniso_block_gmres) against golden fixtures made with the compiled reference `mapping`
(tests/golden/make_golden_block.py) and against the CPU oracle on a seeded input."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
TOL = 1e-10


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def make(name):
    from aniso_b200 import Aniso

    sz, deg, K, g, sr, np_, ml, kind = cases.BLOCK_CASES[name]
    a = Aniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    a.setCoeff(ss, st)
    return a, K, g, ss, x.size


@pytest.mark.parametrize("name", ["a16", "k3h16", "b8d3"])
def test_forward_mforward_vs_reference_golden(golden, name):
    g = golden(name + "_block")
    a, K, gg, ss, n = make(name)
    assert rel(a.forward(g["u"]), g["forward"]) < TOL
    assert rel(a.mforward(g["u"]), g["mforward"]) < TOL
    a.delete()


def test_block_operator_vs_oracle_seeded():
    from oracle import block_oracle, oracleapi

    name = "k3h16"
    a, K, gg, ss, n = make(name)
    sz, deg, _, _, sr, np_, ml, kind = cases.BLOCK_CASES[name]
    o = oracleapi.OracleAniso(sz, deg, K, gg, sr, np_, ml)
    x, y, w = o.nodes()
    o.set_coeff(*cases.sigma_fields(kind, x, y, sz, deg))
    u = cases.random_source(K * n, 5)
    assert rel(a.forward(u), block_oracle.forward(o, u, K, n)) < TOL
    assert rel(a.mforward(u), block_oracle.mforward(o, u, K, n, gg, ss)) < TOL
    # linearity of the block operator
    v = cases.random_source(K * n, 6)
    assert rel(a.mforward(2.0 * u - 3.0 * v), 2.0 * a.mforward(u) - 3.0 * a.mforward(v)) < 1e-12
    a.delete()


@pytest.mark.parametrize("name", ["a16", "k3h16", "b8d3"])
def test_block_gmres_vs_reference_golden(golden, name):
    g = golden(name + "_block")
    a, K, gg, ss, n = make(name)
    charge = g["u"]
    xs, rc, it, hist = a.solve(charge)  # rhs = forward(charge); GMRES(400, 1e-11, 400) on I - mforward
    ref_it = int(g["gm_iters"])
    assert rc == int(g["gm_rc"]) == 0
    assert it == ref_it, (it, ref_it, hist[-3:], g["gm_hist"][-3:])  # 43 / 44 / 55 with every orthogonalisation form (profiles/r02_gmres_gate.txt)
    assert rel(xs, g["gm_x"]) < 1e-8
    m = min(it, ref_it) - 2
    np.testing.assert_allclose(hist[:m], g["gm_hist"][:m], rtol=5e-3, atol=1e-8)
    # the solution satisfies the block equation
    assert rel(xs - a.mforward(xs), g["forward"]) < 1e-9
    a.delete()


def test_block_operator_at_c2():
    """BASELINE config 2 (N = 2^14, K = 5, g = 0.5): properties only (the oracle needs minutes there): out_0 of forward with
    the source in mode 0 contains mapping_0, and linearity."""
    from aniso_b200 import Aniso

    sz, deg, K, g, sr, np_, ml, kind = cases.CASES["c2"]
    a = Aniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = a.nodes()
    n = x.size
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    a.setCoeff(ss, st)
    f = cases.gaussian_source(x, y)
    u = np.zeros(K * n)
    u[:n] = f  # demo.m:24-29
    fw = a.forward(u)
    for i in range(K):  # only j = 0 contributes: rhs_i = mapping(f, i)
        assert rel(fw[i * n:(i + 1) * n], a.mapping(f, i)) < 1e-13
    a.delete()


@pytest.mark.parametrize("cfg", [
    (128, 1, 5, 0.5, 10, 4, 20, "const"),     # BASELINE config 2: K = 5, kernel ids 0..8, 16 x 16 blocks, 4 tree levels
    (64, 1, 3, 0.9, 10, 8, 20, "hetero"),     # 64 x 64 blocks
    (256, 1, 2, 0.9, 10, 4, 20, "layers"),    # N = 2^16, K = 2 as in config 5
])
def test_batched_block_operator_matches_the_loop(cfg, monkeypatch):
    """The batched form (each cached operator streamed once for all K modes, one downward pass per output mode folding the
    2K-1 terms) against the plain loop of K(2K-1) `mapping` calls of the same handle (the form checked against the
    reference fixtures above)."""
    from aniso_b200 import Aniso

    sz, deg, K, g, sr, np_, ml, kind = cfg
    a = Aniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = a.nodes()
    n = x.size
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    a.setCoeff(ss, st)
    u = cases.random_source(K * n, 12) * np.tile(cases.gaussian_source(x, y) + 0.05, K)
    monkeypatch.delenv("ANISO_BLOCK_LOOP", raising=False)
    fw, mf = a.forward(u), a.mforward(u)
    launches_batched = a.launch_count()
    monkeypatch.setenv("ANISO_BLOCK_LOOP", "1")
    fw_loop, mf_loop = a.forward(u), a.mforward(u)
    launches_loop = a.launch_count() - launches_batched
    assert rel(fw, fw_loop) < 1e-12 and rel(mf, mf_loop) < 1e-12
    assert launches_loop > 2 * (launches_batched - 2 * (2 * K - 1) * 4)  # the batched form really is the shorter one
    # 16 x 16 blocks: the batched form above ran on the FP64 tensor cores (mma.m8n8k4); the DFMA / butterfly form of the
    # same multi-RHS stream must agree with it
    monkeypatch.delenv("ANISO_BLOCK_LOOP", raising=False)
    monkeypatch.setenv("ANISO_NO_DMMA", "1")
    assert rel(a.forward(u), fw) < 1e-13 and rel(a.mforward(u), mf) < 1e-13
    monkeypatch.setenv("ANISO_STREAM_REGS", "1")  # ... and with register-held loads instead of the TMA ring
    assert rel(a.forward(u), fw) < 1e-13
    a.delete()


def _diag_prec(n, K=1):
    """a fixed diagonal left preconditioner (stand-in for the FEM diffusion solve of aniso.m:111-119, which is outside the path)"""
    d = 1.0 / (1.0 + 0.5 * np.cos(np.arange(K * n) * 0.37) ** 2)
    return lambda r: d * r


@pytest.mark.parametrize("name", ["a16", "k3h16", "b8d3"])
def test_block_gmres_preconditioner_hook(golden, name):
    """aniso.m:165 `gmres(A, rhs, 400, 1e-11, 400, @obj.prec)`: the identity callback reproduces the unpreconditioned solve
    step for step; a diagonal callback follows the oracle's left-preconditioned GMRES (restated gmres.cpp with M^-1 applied
    to b, r0 and every A v) on the compiled reference's operator values (golden forward / the oracle mforward)."""
    from oracle import block_oracle, oracleapi

    g = golden(name + "_block")
    a, K, gg, ss, n = make(name)
    charge = g["u"]
    x0, rc0, it0, h0 = a.solve(charge)
    x1, rc1, it1, h1 = a.solve(charge, prec=lambda r: r)
    assert rc0 == rc1 == 0 and it1 == it0
    assert rel(x1, x0) < 1e-10
    np.testing.assert_allclose(h1[: it0 - 2], h0[: it0 - 2], rtol=1e-6, atol=1e-12)
    P = _diag_prec(n, K)
    x2, rc2, it2, h2 = a.solve(charge, prec=P)
    sz, deg, _, _, sr, np_, ml, kind = cases.BLOCK_CASES[name]
    o = oracleapi.OracleAniso(sz, deg, K, gg, sr, np_, ml)
    xo, yo, wo = o.nodes()
    o.set_coeff(*cases.sigma_fields(kind, xo, yo, sz, deg))
    A = lambda v: v - block_oracle.mforward(o, v, K, n, gg, ss)  # noqa: E731
    xr, rcr, itr, hr = block_oracle.gmres(A, g["forward"], m=400, max_iter=400, tol=1e-11, prec=P)
    assert rc2 == rcr == 0 and abs(it2 - itr) <= 1, (it2, itr)
    assert rel(x2, xr) < 1e-8
    m = min(it2, itr) - 2
    np.testing.assert_allclose(h2[:m], hr[:m], rtol=5e-3, atol=1e-8)
    assert rel(x2, x0) < 1e-7  # same solution as the unpreconditioned solve
    a.delete()


def test_gmres_preconditioner_hook_single_mode(golden):
    """aniso_gmres_prec on the main.cpp operator: identity = plain solve, diagonal = oracle's left-preconditioned GMRES;
    a failing callback surfaces as an error instead of crossing the C boundary."""
    from aniso_b200 import Aniso, AnisoError
    from oracle import block_oracle, oracleapi

    g = golden("b8d3")
    sz, deg, K, gg, sr, np_, ml, kind = cases.CASES["b8d3"]
    a = Aniso(sz, deg, K, gg, sr, np_, ml)
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    a.setCoeff(ss, st)
    a.cache(0)
    b = g["gm_b"]
    x0, rc0, it0, h0 = a.gmres(b)
    x1, rc1, it1, h1 = a.gmres(b, prec=lambda r: r)
    assert rc0 == rc1 == 0 and it0 == it1 == int(g["gm_iters"]) and rel(x1, x0) < 1e-10
    P = _diag_prec(x.size)
    x2, rc2, it2, h2 = a.gmres(b, prec=P)
    o = oracleapi.OracleAniso(sz, deg, K, gg, sr, np_, ml)
    o.set_coeff(ss, st)
    A = lambda v: v - o.mapping(ss * v, 0)  # noqa: E731
    xr, rcr, itr, hr = block_oracle.gmres(A, b, m=80, max_iter=400, tol=1e-12, prec=P)
    assert rc2 == rcr == 0 and abs(it2 - itr) <= 1
    assert rel(x2, xr) < 1e-8
    with pytest.raises(AnisoError):
        a.gmres(b, prec=lambda r: 1 / 0)
    a.delete()
