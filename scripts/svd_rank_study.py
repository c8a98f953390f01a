"""NS1 (north_star: "M2L ... batched FP64 tensor-core GEMM over SVD-compressed operators"): singular-value decay of the M2L
blocks the matvec streams.  CPU only (numpy); run here, no GPU needed:
    python scripts/svd_rank_study.py > profiles/r02_m2l_svd_ranks.txt
For BASELINE C5 (sz 1024, np 4: 16 x 16 blocks, sigma layers) and C4 np 8 (sz 512: 64 x 64 blocks, heterogeneous sigma) it
samples V-list pairs on every tree level, evaluates the block the library caches, K[t, s] = exp(-tau(s, t)) / |s - t|
between the Chebyshev nodes of the two boxes (bbfmm.h:782-804 with the fused kernel of DESIGN 3.3), and reports the
eps-rank k(eps) = #{sigma_i > eps * sigma_1}.  A rank-k factorisation U V^T of an r x r block stores 2 r k values and serves
both directions of the symmetric pair like the full block does, so it saves bytes only for k < r / 2; the truncation has to
stay near 1e-12 per block to keep the 1e-10 matvec gate (the reference has no compression: SURVEY 7 'Hard parts')."""
import sys

import numpy as np

sys.path.insert(0, "tests")
import cases  # noqa: E402


def line_integral(sig, sz, a, b):
    """exact integral of the piecewise-constant sigma (cell [i, j] = x-cell i, y-cell j) from a to b"""
    h = 1.0 / sz
    d = b - a
    ts = [0.0, 1.0]
    for ax in (0, 1):
        if d[ax] != 0.0:
            lo, hi = sorted((a[ax], b[ax]))
            k = np.arange(np.ceil(lo / h), np.floor(hi / h) + 1)
            t = (k * h - a[ax]) / d[ax]
            ts.extend(t[(t > 0) & (t < 1)])
    ts = np.unique(np.array(ts))
    mid = 0.5 * (ts[1:] + ts[:-1])
    px, py = a[0] + mid * d[0], a[1] + mid * d[1]
    ci = np.clip((px * sz).astype(int), 0, sz - 1)
    cj = np.clip((py * sz).astype(int), 0, sz - 1)
    return float(np.sum(sig[ci, cj] * np.diff(ts)) * np.hypot(d[0], d[1]))


def block(sig, sz, np_, cA, cB, rad):
    t = -np.cos((np.arange(np_) + 0.5) * np.pi / np_)
    pts = lambda c: np.array([[c[0] + rad * t[i], c[1] + rad * t[j]] for j in range(np_) for i in range(np_)])  # noqa: E731
    A, B = pts(cA), pts(cB)
    K = np.empty((len(A), len(B)))
    for it_, pt in enumerate(A):
        for is_, ps in enumerate(B):
            K[it_, is_] = np.exp(-line_integral(sig, sz, ps, pt)) / np.hypot(*(ps - pt))
    return K


def study(name, nblocks):
    sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
    h = 1.0 / sz
    xs = (np.arange(sz) + 0.5) * h
    X, Y = np.meshgrid(xs, xs, indexing="ij")
    _, st = cases.sigma_fields(kind, X.reshape(-1), Y.reshape(-1), sz, 1)
    sig = st.reshape(sz, sz)
    r = np_ * np_
    leaf_pts = 64 if np_ == 8 else 16
    deepest = int(round(np.log(sz * sz / leaf_pts) / np.log(4)))
    radius0 = (1.0 - h) / 2  # root box = bounding box of the nodes (bbfmm.h:231-248)
    rng = np.random.default_rng(20261105)
    print("\n%s: sz %d, np %d (%d x %d blocks), sigma %s, tree levels 2..%d, %d sampled V-list pairs per level" %
          (name, sz, np_, r, r, kind, deepest, nblocks))
    print("level | box width (cells) | eps-rank mean at 1e-6 | 1e-8 | mean / max at 1e-10 | 1e-12 | 1e-13 | sigma_r/sigma_1 (median) | bytes vs full block at 1e-12 (2k/r)")
    tot = {}
    for L in range(2, deepest + 1):
        nb = 1 << L
        rad = radius0 / nb
        ranks = {1e-6: [], 1e-8: [], 1e-10: [], 1e-12: [], 1e-13: []}
        tail = []
        for _ in range(nblocks):
            while True:
                i, j = rng.integers(0, nb, 2)
                di, dj = rng.integers(-3, 4, 2)
                i2, j2 = i + di, j + dj
                if not (0 <= i2 < nb and 0 <= j2 < nb) or max(abs(di), abs(dj)) < 2:
                    continue
                if abs(i // 2 - i2 // 2) > 1 or abs(j // 2 - j2 // 2) > 1:
                    continue
                break
            cA = np.array([0.5 - radius0 + (2 * i + 1) * rad, 0.5 - radius0 + (2 * j + 1) * rad])
            cB = np.array([0.5 - radius0 + (2 * i2 + 1) * rad, 0.5 - radius0 + (2 * j2 + 1) * rad])
            s = np.linalg.svd(block(sig, sz, np_, cA, cB, rad), compute_uv=False)
            for eps in ranks:
                ranks[eps].append(int(np.sum(s > eps * s[0])))
            tail.append(s[-1] / s[0])
        k12 = np.mean(ranks[1e-12])
        print("%5d | %8.1f | %5.1f | %5.1f | %5.1f / %2d | %5.1f / %2d | %5.1f / %2d | %.1e | %.2f" %
              (L, 2 * rad * sz, np.mean(ranks[1e-6]), np.mean(ranks[1e-8]), np.mean(ranks[1e-10]), max(ranks[1e-10]), k12, max(ranks[1e-12]), np.mean(ranks[1e-13]),
               max(ranks[1e-13]), np.median(tail), 2 * k12 / r))
        tot[L] = (k12, 27 * 4 ** L)  # weight: blocks on this level (interior boxes have 27 V entries)
    wsum = sum(w for _, w in tot.values())
    kbar = sum(k * w for k, w in tot.values()) / wsum
    print("block-count-weighted mean eps-rank at 1e-12: %.1f of %d -> low-rank storage would take %.2f x the bytes of the full blocks" %
          (kbar, r, 2 * kbar / r))


if __name__ == "__main__":
    # self-check of the line integral against a brute-force midpoint sum
    rng = np.random.default_rng(1)
    sg = 1 + rng.random((16, 16))
    a, b = np.array([0.13, 0.71]), np.array([0.82, 0.29])
    tt = (np.arange(400000) + 0.5) / 400000
    brute = np.mean(sg[np.minimum((a[0] + tt * (b[0] - a[0])) * 16, 15).astype(int), np.minimum((a[1] + tt * (b[1] - a[1])) * 16, 15).astype(int)]) * np.hypot(*(b - a))
    assert abs(line_integral(sg, 16, a, b) - brute) < 1e-5
    study("c5", int(sys.argv[1]) if len(sys.argv) > 1 else 12)
    study("c4n8", int(sys.argv[2]) if len(sys.argv) > 2 else 6)
