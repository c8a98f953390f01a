"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/aniso_b200.h declares, the host-side
set-up (geometry tables, Gauss-Legendre rules, quadtree and interaction lists, leaf partition) reproduces the golden
fixtures generated from the compiled reference, and compute entry points fail loudly without a device."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["a16", "b8d3", "h32", "nu6d5", "ml12d2"]


@pytest.fixture(scope="module")
def lib():
    from aniso_b200 import _lib

    if not os.path.exists(_lib.LIB_PATH):
        subprocess.check_call(["make", "-s", "-j4", "-C", os.path.join(ROOT, "aniso_b200", "csrc")])
    return _lib.load()


def make(name):
    from aniso_b200 import Aniso

    sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
    return Aniso(sz, deg, K, g, sr, np_, ml)


def test_header_symbols_exported(lib):
    from aniso_b200 import _lib

    hdr = open(os.path.join(ROOT, "include", "aniso_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(aniso_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 40
    for name in declared:
        assert hasattr(lib, name), "symbol %s declared in aniso_b200.h but not exported" % name
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_gl_tables_bit_exact(lib, golden):
    from aniso_b200.api import gl_table
    from aniso_b200 import AnisoError

    g = golden("gl_tables")
    for d in range(1, 129):
        x, w = gl_table(d)
        assert np.array_equal(x, g["x%d" % d]) and np.array_equal(w, g["w%d" % d]), d
    with pytest.raises(AnisoError):  # "The degree is not implemented." Quadrature.cpp:22189-22191
        gl_table(129)


@pytest.mark.parametrize("name", SMALL)
def test_geometry_tables(lib, golden, name):
    g = golden(name)
    a = make(name)
    x, y, w = a.nodes()
    assert np.array_equal(x, g["x"]) and np.array_equal(y, g["y"]) and np.array_equal(w, g["w"])
    assert np.array_equal(a.getNodes(), np.stack([g["x"], g["y"]], axis=1))
    for k, v in a.geometry_tables().items():
        np.testing.assert_allclose(v, g["geo_" + k], rtol=0, atol=1e-15)
    for got, want in zip(a.sing_tables(), (g["singX"], g["singY"], g["singW"])):
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-15)
    a.delete()


@pytest.mark.parametrize("name", SMALL + ["c2", "c1"])
def test_tree_matches_reference(lib, golden, name):
    g = golden(name)
    a = make(name)
    ts = a.tree_stats()
    got = [ts[k] for k in ("nodes", "leaves", "max_level", "sumU", "sumV", "sumW", "sumX", "p2p_pairs", "max_leaf")]
    assert got == list(g["tree_stats"])
    if "tree_geom" in g:  # same boxes as the reference tree (numbering differs: compare as sorted sets)
        geom, info = a.tree_nodes()
        key = lambda G, I: sorted(zip(I[:, 0], np.round(G[:, 1], 12), np.round(G[:, 0], 12), I[:, 1], I[:, 2]))
        assert key(geom, info) == key(g["tree_geom"], g["tree_info"])
        # parent links are consistent with the geometry
        for i in range(1, len(geom)):
            p = info[i, 3]
            assert info[p, 0] == info[i, 0] - 1
            assert abs(geom[p, 2] - 2 * geom[i, 2]) < 1e-15 and abs(abs(geom[p, 0] - geom[i, 0]) - geom[i, 2]) < 1e-12
    a.delete()


def test_interaction_lists_against_oracle(lib):
    """U/V/W/X lists node by node (non-uniform tree with W and X entries) against the CPU oracle."""
    from oracle import oracleapi

    sz, deg, K, g, sr, np_, ml, kind = cases.CASES["nu6d5"]
    a = make("nu6d5")
    o = oracleapi.OracleAniso(sz, deg, K, g, sr, np_, ml)
    x, y, _ = o.nodes()
    o.set_coeff(*cases.sigma_fields(kind, x, y, sz, deg))  # builds the oracle tree
    ga, ia = a.tree_nodes()
    go, io = o.tree_nodes()
    box = lambda G, I, i: (int(I[i, 0]), round(float(G[i, 0]), 12), round(float(G[i, 1]), 12))
    omap = {box(go, io, i): i for i in range(len(go))}
    seen_w = seen_x = 0
    for i in range(len(ga)):
        ids, kinds = a.node_lists(i)
        oid, okinds = o.node_lists(omap[box(ga, ia, i)])
        mine = sorted((int(k), box(ga, ia, int(j))) for j, k in zip(ids, kinds))
        theirs = sorted((int(k), box(go, io, int(j))) for j, k in zip(oid, okinds))
        assert mine == theirs, i
        seen_w += int((kinds == 3).sum())
        seen_x += int((kinds == 1).sum())
    assert seen_w == 704 and seen_x == 704
    a.delete()


@pytest.mark.parametrize("name,nranks", [("c2", 2), ("c2", 4), ("c2", 8), ("nu6d5", 3), ("a16", 8)])
def test_partition_covers_and_balances(lib, name, nranks):
    a = make(name)
    n = a.numberOfNodes
    ranges = [a.partition(nranks, r) for r in range(nranks)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (b0, e0, _), (b1, e1, _) in zip(ranges, ranges[1:]):
        assert e0 == b1 and b0 <= e0
    perm = a.slot_permutation()
    assert sorted(perm) == list(range(n))
    if name == "c2":  # uniform tree: near-equal shares
        sizes = [e - b for b, e, _ in ranges]
        assert max(sizes) <= 1.25 * n / nranks
    a.delete()


def test_compute_fails_loudly_without_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from aniso_b200 import AnisoError

    a = make("a16")
    with pytest.raises(AnisoError) as e:
        a.setCoeff(np.ones(a.numberOfNodes), np.ones(a.numberOfNodes))
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(AnisoError):
        a.mapping(np.ones(a.numberOfNodes), 0)
    a.delete()


def test_bad_arguments(lib):
    from aniso_b200 import Aniso, AnisoError

    for args in [(0, 1, 1, 0.5, 4, 2, 20), (8, 0, 1, 0.5, 4, 2, 20), (8, 1, 0, 0.5, 4, 2, 20), (8, 1, 1, 0.5, 4, 0, 20),
                 (8, 1, 1, 0.5, 4, 17, 20)]:
        with pytest.raises(AnisoError):
            Aniso(*args)


def test_product_does_not_import_oracle():
    """the product package must never route through the CPU oracle"""
    for root, _, files in os.walk(os.path.join(ROOT, "aniso_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".hpp", ".cuh", ".h")):
                src = open(os.path.join(root, f)).read()
                assert "oracle" not in src.lower(), os.path.join(root, f)


@pytest.mark.skipif(not os.path.isdir("/root/reference/mexplus"), reason="needs the reference's mexplus headers (build container only)")
def test_mex_gateway_compiles_against_stub_mex():
    """the MEX gateway rebuilt on the C-ABI (examples/AnisoWrapper_b200.cpp) compiles against a
    declarations-only mex.h and the reference's own mexplus (MATLAB itself is not installed)."""
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run(["make", "-s", "-C", os.path.join(root, "oracle"), "mexcheck"], capture_output=True, text=True)
    assert out.returncode == 0 and "mexcheck ok" in out.stdout, out.stderr[-2000:]


def test_bench_reference_arm_contract():
    """`bench.py --impl reference`: runs on the host cores without a GPU, prints one JSON line with the contract
    keys, uses every available core even when the launcher exported OMP_NUM_THREADS=1 (torchrun does), and only on rank 0."""
    import json
    import subprocess
    import sys

    from oracle import refapi

    if not refapi.available(fast=True):
        pytest.skip("oracle/_ref not built")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-sz", "32"]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "config",
                "cpu_baseline", "e2e"):
        assert key in line
    assert line["impl"] == "reference" and line["value"] > 0 and line["gpu_launches"] == 0
    assert line["cpu_baseline"]["kind"] == "reference"
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["value"] == line["value"]
    other = subprocess.run(cmd, env=dict(env, RANK="1", WORLD_SIZE="2"), capture_output=True, text=True, timeout=600, cwd=root)
    assert other.returncode == 0 and other.stdout.strip() == ""


@pytest.mark.parametrize("name", ["p1500n4", "p900n6", "p700cap"])
def test_bbfmm_point_set_tree_matches_reference(lib, golden, name):
    """bbfmm::kernel::initialize on an arbitrary (clustered) point set (bbfmm.h:476-505 -> tree::populate): the host plan of
    aniso_fmm_create reproduces the reference engine's tree - node, leaf and list counts, P2P pairs, largest leaf - with a
    leaf capacity that differs from np^2 (p900n6) and a depth cap (p700cap).  Host only."""
    import ctypes as C

    from aniso_b200 import _lib

    g = golden("bbfmm_points")
    n, np_, rank, ml = (int(v) for v in g[name + "_cfg"])
    h = C.c_void_p()
    _lib.check(lib.aniso_fmm_create(np_, n, g[name + "_x"], g[name + "_y"], rank, ml, C.byref(h)))
    st = np.zeros(9, np.int64)
    _lib.check(lib.aniso_tree_stats(h, st))
    assert list(st) == list(g[name + "_stats"])
    assert lib.aniso_num_nodes(h) == n
    with pytest.raises(_lib.AnisoError) as e:  # an Aniso entry point on a point-set handle
        _lib.check(lib.aniso_set_coeff(h, np.ones(n), np.ones(n)))
    assert e.value.code == -4
    lib.aniso_destroy(h)
    with pytest.raises(_lib.AnisoError):
        _lib.check(lib.aniso_fmm_create(np_, 0, g[name + "_x"], g[name + "_y"], rank, ml, C.byref(h)))


def test_group_handle_host_side(lib):
    """aniso_create_multi (one process, several GPUs) without a GPU: creating the group is host-only; the host getters answer
    from rank 0, the group owns every row, compute entry points fail loudly on the ranks (no CPU fallback), device-pointer
    entry points and preconditioner callbacks are rejected, bad device lists are refused."""
    import ctypes as C

    from aniso_b200 import Aniso, AnisoError, _lib

    a = Aniso(16, 1, 2, 0.5, 10, 4, 20, devices=[0, 1, 2])
    one = Aniso(16, 1, 2, 0.5, 10, 4, 20)
    assert a.numberOfNodes == one.numberOfNodes == 256 and a.kernelSize == 3
    assert a.owned_range() == (0, 256)
    assert a.tree_stats() == one.tree_stats()
    for u, v in zip(a.nodes(), one.nodes()):
        assert np.array_equal(u, v)
    assert np.array_equal(a.slot_permutation(), one.slot_permutation())
    with pytest.raises(AnisoError) as e:
        a.setCoeff(np.ones(256), np.ones(256))
    assert e.value.code == -2 and "rank" in str(e.value) and "no CPU fallback" in str(e.value)
    with pytest.raises(AnisoError) as e:
        _lib.check(lib.aniso_forward_dev(a.h, C.c_void_p(8), C.c_void_p(8)))
    assert e.value.code == -1 and "single-process multi-GPU" in str(e.value)
    with pytest.raises(AnisoError) as e:
        a.gmres(np.ones(256), prec=lambda r: r)
    assert e.value.code == -1
    assert lib.aniso_stream(a.h) is None
    a.delete()
    one.delete()
    h = C.c_void_p()
    with pytest.raises(AnisoError):  # the same device twice
        _lib.check(lib.aniso_create_multi(16, 1, 2, 0.5, 10, 4, 20, 2, np.array([1, 1], np.int32), C.byref(h)))
    _lib.check(lib.aniso_create_multi(16, 1, 2, 0.5, 10, 4, 20, 1, np.array([0], np.int32), C.byref(h)))  # ndev 1: a plain handle
    assert lib.aniso_num_nodes(h) == 256
    lib.aniso_destroy(h)


def test_rank_major_multipole_layout_properties(tmp_path):
    """The rank-major packing of the sharded upward pass (mc_node / DevTree::pack_* in aniso_b200/csrc/device.cuh; the
    reference's upPass, bbfmm.h:817-861, has one address space and no such layout): a bijection, one contiguous block per rank
    (so one all-gather moves every level), Morton order kept inside a block.  tests/data/mc_node_check.cu includes the
    product header and runs the mapping on the host."""
    import shutil

    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found")
    exe = str(tmp_path / "mc_node_check")
    c = subprocess.run([nvcc, "-std=c++17", "-I", os.path.join(ROOT, "aniso_b200", "csrc"), "-o", exe,
                        os.path.join(ROOT, "tests", "data", "mc_node_check.cu")], capture_output=True, text=True, timeout=600)
    assert c.returncode == 0, c.stderr[-2000:]
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "mc_node ok" in out.stdout, out.stdout[-2000:]


def test_symmetric_block_storage_invariants(tmp_path):
    """The one-block-per-pair storage of the V/X and U/W lists (symmetrize in aniso_b200/csrc/plan.cpp; the reference applies
    both blocks of a pair, bbfmm.h:863-946): every list entry is covered exactly once, as a stored block or as the transposed
    product of its partner's block; mirror rows lie in the partner's scratch range and no scratch row has two writers; for
    complete, ragged, depth-capped and sharded plans.  tests/data/plan_sym_check.cpp is compiled with the product's plan.cpp."""
    exe = str(tmp_path / "plan_sym_check")
    src = os.path.join(ROOT, "aniso_b200", "csrc")
    c = subprocess.run(["g++", "-O2", "-std=c++17", "-fopenmp", "-I", src, "-o", exe,
                        os.path.join(ROOT, "tests", "data", "plan_sym_check.cpp"), os.path.join(src, "plan.cpp")],
                       capture_output=True, text=True, timeout=600)
    assert c.returncode == 0, c.stderr[-2000:]
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "plan symmetric storage ok" in out.stdout, out.stdout[-3000:]
