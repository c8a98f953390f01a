"""Drop-in tests of the C++ boundary (GPU):
  * oracle/_ref/main_dropin = the reference's own main.cpp, compiled unmodified against include/compat and linked to
    libaniso_b200.so (oracle/Makefile `dropin`, built in the build container where /root/reference exists).  It must
    reproduce the known answers recorded from the reference binary: iteration counts, printed residuals and
    result.csv entries.
  * examples/solve = our own driver on the same headers using the fused matvec and the device-resident GMRES."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "oracle", "_ref", "main_dropin")
SOLVE = os.path.join(ROOT, "examples", "solve")

# SURVEY 8(c) anchors recorded from the compiled reference (-O2): cfg -> (iterations, {iteration: printed residual},
# last printed residual, leading entries of result.csv)
ANCHORS = {
    "c1.cfg": (29, {2: 4.532870e-01, 3: 2.509016e-01, 28: 4.932291e-12, 29: 1.525963e-12},
               [1.98398813154482e-02, 1.93556984034883e-02, 2.03310830765627e-02]),
    "s8d3.cfg": (48, {48: 1.811987e-12}, [-6.89602090507962e-01]),
}


def run(binary, cfg, tmp_path, *extra, **more_env):
    import time

    env = dict(os.environ, **more_env)
    t0 = time.perf_counter()
    p = subprocess.run([binary, os.path.join(ROOT, "tests", "data", cfg), *extra], cwd=tmp_path, capture_output=True,
                       text=True, timeout=600, env=env)
    wall = time.perf_counter() - t0
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    out_dir = os.path.join(ROOT, "gpurun_out")  # wall times of the drop-in lanes, for DESIGN.md (whole process: set-up included)
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "dropin_times.txt"), "a") as f:
            f.write("%s %s %s wall %.3f s\n" % (os.path.basename(binary), cfg, " ".join("%s=%s" % kv for kv in more_env.items()), wall))
    rows = {int(m.group(1)): float(m.group(2)) for m in re.finditer(r"^\s*(\d+)\s+([0-9.]+e[-+]\d+)\s+[0-9.]+\s*$", p.stdout, re.M)}
    return p.stdout, rows


@pytest.mark.skipif(not os.path.exists(DROPIN), reason="oracle/_ref/main_dropin not built (needs /root/reference)")
@pytest.mark.parametrize("cfg", sorted(ANCHORS))
def test_reference_main_cpp_on_b200_library(cfg, tmp_path):
    iters, resid, head = ANCHORS[cfg]
    out, rows = run(DROPIN, cfg, tmp_path)
    assert "CONVERGED" in out and "NOT CONVERGED" not in out
    assert max(rows) == iters, (max(rows), iters)
    for j, r in resid.items():
        # early residuals to the printed 7 digits; the last ones sit at rounding level (1e-12)
        assert abs(rows[j] / r - 1) < (1e-5 if r > 1e-9 else 0.5), (j, rows[j], r)
    x = np.loadtxt(os.path.join(tmp_path, "result.csv"))
    for k, v in enumerate(head):
        assert abs(x[k] / v - 1) < 1e-9, (k, x[k], v)
    assert os.path.exists(os.path.join(tmp_path, "points.csv"))


@pytest.mark.skipif(not os.path.exists(DROPIN), reason="oracle/_ref/main_dropin not built (needs /root/reference)")
def test_reference_main_cpp_on_two_gpus(tmp_path):
    """The same unmodified main.cpp, one process, two GPUs: ANISO_DEVICES=0,1 makes the compat Aniso object a group handle
    (aniso_create_multi).  Same anchors as on one GPU."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    iters, resid, head = ANCHORS["c1.cfg"]
    out, rows = run(DROPIN, "c1.cfg", tmp_path, ANISO_DEVICES="0,1")
    assert "CONVERGED" in out and "NOT CONVERGED" not in out
    assert max(rows) == iters, (max(rows), iters)
    x = np.loadtxt(os.path.join(tmp_path, "result.csv"))
    for k, v in enumerate(head):
        assert abs(x[k] / v - 1) < 1e-9, (k, x[k], v)


@pytest.mark.skipif(not os.path.exists(SOLVE), reason="examples/solve not built")
def test_example_driver_fused_path(tmp_path):
    iters, resid, head = ANCHORS["c1.cfg"]
    out, rows = run(SOLVE, "c1.cfg", tmp_path, "x.csv")
    assert "CONVERGED" in out and "NOT CONVERGED" not in out
    assert max(rows) == iters, (max(rows), iters)
    x = np.loadtxt(os.path.join(tmp_path, "x.csv"))
    for k, v in enumerate(head):
        assert abs(x[k] / v - 1) < 1e-9


def test_blackbox_kernel_source_compiles_against_both_headers(golden, tmp_path):
    """bbfmm::kernel as a black box (bbfmm.h:451-592): ONE source file written against the reference API only
    (tests/data/bbfmm_blackbox.cpp: its own unsymmetric eval lambda, initialize, runCache, re-initialize, runFast) was
    compiled against the reference's bbfmm.h and run on the CPU for the fixture (tests/golden/make_golden_bbfmm.py); the same
    file compiled against include/compat/bbfmm/bbfmm.h + libaniso_b200.so must print the same potentials on the B200."""
    import subprocess

    exe = os.path.join(ROOT, "oracle", "_ref", "bbfmm_dropin")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/bbfmm_dropin not built (make -C oracle bbfmm)")
    g = golden("bbfmm_blackbox")
    n, np_, rank, ml = (int(v) for v in g["cfg"])
    inp = tmp_path / "pts.txt"
    inp.write_text("%d %d %d %d\n" % (n, np_, rank, ml) + "".join("%.17g %.17g %.17g\n" % t for t in zip(g["x"], g["y"], g["q"])))
    out = subprocess.run([exe, str(inp)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    got = np.array([float(t) for t in out.stdout.split()])
    assert got.size == 2 * n
    for part in (slice(0, n), slice(n, 2 * n)):  # runCache, then runFast with doubled charges
        assert np.linalg.norm(got[part] - g["pot"][part]) / np.linalg.norm(g["pot"][part]) < 1e-10
