"""GPU test of the allocation-free entry point aniso_line_integral_noalloc (include/aniso_b200.h; no counterpart in the
reference): same kernel as aniso_line_integral on a fixed scratch array in the
library image instead of a cudaMalloc / cudaFree pair per call.  Through the C-ABI, against the existing entry point, on two
calls in a row that share the scratch array.  Collected last on purpose."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

RTOL = 1e-12  # the tolerance test_gpu_parity.py uses for line integrals; the two entry points run the same kernel


def make(name):
    from aniso_b200 import Aniso

    sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
    a = Aniso(sz, deg, K, g, sr, np_, ml)
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields(kind, x, y, sz, deg)
    return a, ss, st


def segments(n, seed):
    return cases.splitmix64(seed, 4 * n).reshape(n, 4) * 0.98 + 0.01  # strictly inside the unit square


@pytest.mark.parametrize("name", ["a16", "b8d3"])  # deg 1: table line integral; deg 3: cell walk
def test_noalloc_matches_existing_entry_point_twice(golden, name):
    a, ss, st = make(name)
    a.setCoeff(ss, st)
    # first call: many segments; second call: fewer and different ones, so that stale contents of the shared array
    # (inputs or results of the first call) would show up as wrong values
    first, second = segments(5000, 101), segments(37, 202)
    want1, want2 = a.line_integral(first), a.line_integral(second)
    got1 = a.line_integral_noalloc(first)
    got2 = a.line_integral_noalloc(second)
    np.testing.assert_allclose(got1, want1, rtol=RTOL, atol=0)
    np.testing.assert_allclose(got2, want2, rtol=RTOL, atol=0)
    assert np.all(got1 > 0) and np.all(got2 > 0)
    # and against the compiled reference's values at its own seeded segments
    g = golden(name)
    np.testing.assert_allclose(a.line_integral_noalloc(g["li_pts"]), g["li"], rtol=RTOL)
    # a third call repeating the first: the array carries nothing over
    assert np.array_equal(a.line_integral_noalloc(first), got1)
    a.delete()


def test_noalloc_limits_and_call_order():
    from aniso_b200 import AnisoError, _lib

    a, ss, st = make("a16")
    with pytest.raises(AnisoError) as e:  # before setCoeff: no lazy set-up (it would allocate)
        a.line_integral_noalloc(segments(4, 1))
    assert e.value.code == -4
    a.setCoeff(ss, st)
    nmax = _lib.NOALLOC_MAX_SEGMENTS
    full = segments(nmax, 7)
    np.testing.assert_allclose(a.line_integral_noalloc(full), a.line_integral(full), rtol=RTOL, atol=0)  # the maximum itself
    with pytest.raises(AnisoError) as e:
        a.line_integral_noalloc(segments(nmax + 1, 8))
    assert e.value.code == -1 and "exceed the fixed scratch" in str(e.value)
    assert a.line_integral_noalloc(np.zeros((0, 4))).size == 0
    # a second handle on the same GPU shares the array: used one after the other, both stay correct
    b, ss2, st2 = make("b8d3")
    b.setCoeff(ss2, st2)
    p = segments(300, 9)
    rb, ra = b.line_integral_noalloc(p), a.line_integral_noalloc(p)
    np.testing.assert_allclose(rb, b.line_integral(p), rtol=RTOL, atol=0)
    np.testing.assert_allclose(ra, a.line_integral(p), rtol=RTOL, atol=0)
    a.delete()
    b.delete()
