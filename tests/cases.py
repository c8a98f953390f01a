"""Shared synthetic inputs for tests / bench (SURVEY 8 'Synthetic inputs').  Pure numpy."""
import numpy as np

MASK = (1 << 64) - 1


def splitmix64(seed, n):
    """n uniform doubles in [0,1) from splitmix64(seed); the state after i draws is
    seed + i * golden (mod 2^64), so the sequence vectorises exactly."""
    with np.errstate(over="ignore"):
        s = np.uint64(seed & MASK) + np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
        z = (s ^ (s >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z ^= z >> np.uint64(31)
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def gaussian_source(x, y):
    """main.cpp:29-32"""
    return np.exp(-25 * ((x - 0.5) ** 2 + (y - 0.5) ** 2))


def tilted_source(x, y):
    """anchor source of SURVEY 8(c): e^{-25 rho^2} (1 + x + 2y)"""
    return gaussian_source(x, y) * (1 + x + 2 * y)


def random_source(n, seed=1):
    return 2 * splitmix64(seed, n) - 1


def sigma_fields(kind, x, y, sz, deg):
    """returns (sigma_s, sigma_t) per node."""
    if kind == "main":  # main.cpp:34-40
        ss = 16 * 0.5 * (1 - np.cos(2 * np.pi * x))
        return ss, ss + 0.2
    if kind == "const":  # demo.m:15-16
        ss = np.full_like(x, 20.0)
        return ss, ss + 0.2
    if kind == "hetero":  # C3: sigma_s = 4 + 12 u(cell), sigma_a = 0.2
        u = splitmix64(0xA715020, sz * sz)
        ss = np.repeat(4 + 12 * u, deg * deg)
        return ss, ss + 0.2
    if kind == "layers":  # C5: 4 bands in y
        band = np.minimum((y * 4).astype(int), 3)
        ss = np.array([2.0, 16.0, 6.0, 12.0])[band]
        sa = np.array([0.1, 0.5, 0.2, 0.3])[band]
        return ss, ss + sa
    raise ValueError(kind)


# name -> (sz, deg, K, g, sr, np, maxLevel, sigma kind)
CASES = {
    "a16": (16, 1, 2, 0.5, 10, 4, 20, "main"),      # SURVEY 8(c) anchor
    "b8d3": (8, 3, 2, 0.5, 8, 4, 20, "main"),       # SURVEY 8(c) anchor, deg 3
    "h32": (32, 1, 2, 0.9, 10, 6, 20, "hetero"),    # C3-like, np 6
    "nu6d5": (6, 5, 1, 0.5, 6, 2, 20, "layers"),    # non-uniform tree (W/X lists)
    "ml12d2": (12, 2, 1, 0.5, 4, 3, 2, "main"),     # maxLevel cap, non power-of-two sz
    "c2": (128, 1, 5, 0.5, 10, 4, 20, "const"),     # BASELINE config 2
    "c1": (64, 3, 1, 0.95, 8, 4, 20, "main"),       # BASELINE config 1 (data.cfg)
    "c3": (256, 1, 5, 0.9, 10, 6, 20, "hetero"),    # BASELINE config 3 (N = 2^16, np 6, heterogeneous sigma)
    "c4n4": (512, 1, 1, 0.9, 10, 4, 20, "hetero"),  # BASELINE config 4 (N = 2^18), Chebyshev order 4
    "c4n8": (512, 1, 1, 0.9, 10, 8, 20, "hetero"),  # BASELINE config 4, Chebyshev order 8 (64-point leaves)
    "c5": (1024, 1, 1, 0.9, 10, 4, 20, "layers"),   # BASELINE config 5 (N = 2^20): the headline workload of bench.py
}

# multi-mode block operator (aniso.m forward / mforward / solve): name -> same tuple
BLOCK_CASES = {
    "a16": CASES["a16"],                              # K 2, g 0.5
    "k3h16": (16, 1, 3, 0.9, 10, 4, 20, "hetero"),    # K 3 (kernel ids 0..4), strong forward scattering, heterogeneous sigma
    "b8d3": CASES["b8d3"],                            # deg 3
}
