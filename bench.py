#!/usr/bin/env python
"""bench.py - headline benchmark of the aniso hot path (BASELINE.json: FMM matvec time & near-field interactions/s;
GMRES time-to-solve at N = 2^20 on 1/2/4/8 GPUs).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--sz 1024]

Workload (config C5 of SURVEY 8): unit square, sz x sz cells, deg 1 => N = sz^2 nodes (2^20), Chebyshev order 4,
g = 0.9, piecewise sigma_s / sigma_t layers, Gaussian source; kernel id 0.  One "step" = one application of the
GMRES operator y = x - mapping_0(sigma_s .* x) (main.cpp:123-136), i.e. exactly one FMM matvec `mapping`.
  value  device-resident throughput, target points per second (N / time of one matvec), CUDA events on the
         library's stream, max over ranks.  Every step streams the cached operators (5.9 GB at C5 >> 126 MB L2), so
         no L2 flush is needed between iterations.
  e2e    the same through the reference-facing C-ABI call aniso_mapping() with HOST buffers (H2D + D2H inside).
  --impl reference  times the compiled, unmodified reference (oracle/_ref/libaniso_ref_fast.so, OpenMP on all host
         cores) on the SAME configuration (N = 2^20) when a short probe at N = 2^16 projects the whole run under
         REF_BUDGET_S and the host has the memory (about 30 GB); otherwise on the largest of N = 2^18 / 2^16 / 2^14 that
         fits, and says which in config.sample.
  parity   the line carries the comparison of this very run with the reference: against the committed fixture of the
         compiled reference at N = 2^20 (tests/golden/c5.npz) and, in cpu_baseline, against the reference run beside it.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402

METRIC = "fmm_matvec_points_per_s"
UNIT = "points/s"
NP_ORDER, SING_RULE, G_ANISO, MAX_LEVEL = 4, 10, 0.9, 20
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the M2L stream kernel, from the committed `ncu --set full`
# capture profiles/r02_ncu_raw.csv (cannot be measured live); keyed by (sz, n_gpus)
# (TMA-ring kernel: the two launches of a matvec, levels 2-7 and level 8, read 0.590 + 1.813 GB and write 0.007 + 0.077 GB)
NCU_TRAFFIC = {(1024, 1): 590148000 + 1813174000 + 8699000 + 78852000}
NCU_TRAFFIC_SOURCE = "profiles/r02_ncu_raw.csv"


def workload_name(sz):
    return ("C5: N=%d (sz %d, deg 1), np %d, g %.1f, sigma layers, kernel id 0; step = y = x - mapping_0(sigma_s x)"
            % (sz * sz, sz, NP_ORDER, G_ANISO))


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the measured regions (B200_PROFILING.md): one `nvidia-smi -lms 50`
    child for the whole measurement section (device-timed steps, e2e steps, roofline launches), rows read by a thread."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.t, self.start = index, [], None, None, 0

    def mark(self):
        """Rows before this point (sampler start-up, warm-up) are not part of the summary."""
        self.start = len(self.rows)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.strip().split(",")])

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
            self.t.join(timeout=5)

    def summary(self):
        ok = [r for r in self.rows[self.start:] if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        sm = [float(r[0]) for r in ok]
        mx = [float(r[1]) for r in ok if r[1].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({n for r in ok for n, v in zip(names, r[2:6]) if v.lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


REF_BUDGET_S = 600.0  # the reference arm must end within a few minutes (the driver's per-N limit is ~870 s)


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is one process that must use every host core.
    The variable only matters until libgomp has read it, so the run-time call is made as well."""
    import ctypes

    cores = host_cores()
    os.environ["OMP_NUM_THREADS"] = str(cores)
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(cores)
    except OSError:
        pass
    return cores


def make_reference(sz):
    """(object, kind, cores): the compiled reference with its own flags when oracle/_ref travelled, else the scalar C port"""
    from oracle import refapi  # bench.py executes oracle/ only in run_reference and cpu_baseline

    if refapi.available(fast=True):
        r = refapi.RefAniso(sz, 1, 1, G_ANISO, SING_RULE, NP_ORDER, MAX_LEVEL, fast=True)
        return r, "reference", int(r.num_threads())
    from oracle import oracleapi

    return oracleapi.OracleAniso(sz, 1, 1, G_ANISO, SING_RULE, NP_ORDER, MAX_LEVEL), "port", 1


def time_reference(r, sz, warmup, steps):
    """cache phase + `steps` applications of the GMRES operator on the reference object r"""
    x, y, w = r.nodes()
    ss, st = cases.sigma_fields("layers", x, y, sz, 1)
    t0 = time.perf_counter()
    r.set_coeff(ss, st)
    times = r.cache(0)
    t_cache = time.perf_counter() - t0
    f = cases.gaussian_source(x, y)
    for _ in range(warmup):
        r.mapping(ss * f, 0)
    t0 = time.perf_counter()
    out = None
    for _ in range(steps):
        out = f - r.mapping(ss * f, 0)  # same operator as our step
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return {"t_cache": t_cache, "cache_times": times, "dt": dt, "out": out, "f": f, "ss": ss, "x": x, "y": y}


def choose_reference_sz(args, kind):
    """Same configuration (sz 1024) when it fits the time and memory budget, judged from a probe at sz 256 (the FMM apply
    is O(N); the cache phase grows a little faster because coarse-level rays cross more cells: measured 1.3x per 4x points)."""
    if args.ref_sz:
        return args.ref_sz, None
    if kind != "reference":
        return 64, None
    r, _, _ = make_reference(256)
    p = time_reference(r, 256, 0, 2)
    r.close()
    mem = mem_available_gb()
    probe = {"sz": 256, "cache_s": p["t_cache"], "matvec_s": p["dt"], "mem_available_gb": mem}
    napply = args.warmup + args.steps + (30 if args.gpus == 1 else 0)  # + the GMRES solve reported at N = 1
    for sz, need_gb in ((args.sz, 48.0), (512, 12.0)):
        scale = (sz / 256.0) ** 2
        growth = 1.3 ** (np.log(scale) / np.log(4.0)) if scale > 1 else 1.0
        projected = p["t_cache"] * scale * max(growth, 1.0) + p["dt"] * scale * napply
        probe["projected_s_sz%d" % sz] = projected
        if sz >= 256 and projected <= REF_BUDGET_S and mem >= need_gb:
            return sz, probe
    return 256, probe


def run_reference(args):
    """The reference's own CPU implementation (compiled unmodified, its own -Ofast/-ffast-math flags, OpenMP on all
    host cores); rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    use_all_host_threads()
    from oracle import refapi

    kind = "reference" if refapi.available(fast=True) else "port"
    sz, probe = choose_reference_sz(args, kind)
    r, kind, cores = make_reference(sz)
    t = time_reference(r, sz, args.warmup, args.steps)
    n = sz * sz
    val = n / t["dt"]
    same = sz == args.sz
    sample = ("the full workload, N=%d (sz %d): same configuration as the GPU arm" % (n, sz)) if same else \
        ("N=%d (sz %d) sub-problem of the same workload (%s); per-point cost of the O(N) FMM"
         % (n, sz, "size chosen with --ref-sz" if args.ref_sz else
            "the full size does not fit the %d s / memory budget of this arm on this host" % int(REF_BUDGET_S)))
    sample += "; cache phase %.1f s" % t["t_cache"]
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t["dt"] * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.sz), "sample": sample, "same_config": same},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                             "cache_phase_s": t["t_cache"], "ms_per_matvec": t["dt"] * 1e3},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "checksum": float(np.linalg.norm(t["out"]))}
    if probe:
        line["probe"] = probe
    ct = t["cache_times"]
    if ct is not None and len(ct) >= 4:
        line["cpu_baseline"]["cache_labels_s"] = {"Non Singular Cache": float(ct[0]), "Singular Cache": float(ct[1]),
                                                  "NearAddOn Cache": float(ct[2]), "SingularAddOn Cache": float(ct[3])}
        ts = r.tree_stats(0)
        evals = ts["p2p_pairs"] + (ts["sumV"] + ts["sumX"]) * NP_ORDER ** 4
        if ct[0] > 0:
            line["cpu_baseline"]["kernel_evals_per_s"] = evals / float(ct[0])
    if args.gpus == 1 and not args.no_ref_gmres:
        # GMRES time-to-solve of the reference on this host (main.cpp:141: GMRES(80, 400, 1e-12), x0 = 0)
        b = r.mapping(t["f"], 0)
        t0 = time.perf_counter()
        xs, rc, it, hist = r.gmres(b, m=80, max_iter=400, tol=1e-12)
        line["gmres"] = {"rc": int(rc), "iters": int(it), "time_to_solve_s": time.perf_counter() - t0,
                         "last_printed_resid": float(hist[-1]) if len(hist) else None, "m": 80, "tol": 1e-12}
    print(json.dumps(line), flush=True)


def cpu_baseline(args):
    """Bounded CPU sample (10-30 s of CPU work) timed beside the GPU run: rank 0, N=1 only.  The compiled reference
    (its own -Ofast flags, OpenMP on all host cores) when oracle/_ref is present, else the scalar C port.  Its output is
    not thrown away: the GPU library runs the same sub-problem and the line reports how far the two are apart."""
    from oracle import refapi

    cores = host_cores()
    sz = (256 if cores >= 12 else 128) if refapi.available(fast=True) else 64
    r, kind, cores = make_reference(sz)
    t = time_reference(r, sz, 1, 5)
    ts = r.tree_stats(0)
    evals = ts["p2p_pairs"] + (ts["sumV"] + ts["sumX"]) * NP_ORDER ** 4
    n = sz * sz
    out = {"value": n / t["dt"], "unit": UNIT, "cores": cores, "kind": kind,
           "sample": "N=%d (sz %d) sub-problem of the same workload, 5 matvecs after a %.1f s cache phase" % (n, sz, t["t_cache"]),
           "ms_per_matvec": t["dt"] * 1e3, "cache_phase_s": t["t_cache"]}
    ct = t["cache_times"]
    if ct is not None and len(ct) and ct[0] > 0:
        out["kernel_evals_per_s"] = evals / float(ct[0])  # 'Non Singular Cache': P2P + M2L evaluations with line integrals
    # parity of the GPU path against this very reference run (same sub-problem, same inputs)
    from aniso_b200 import Aniso

    f, ss = t["f"], t["ss"]
    b_ref = r.mapping(f, 0)
    t0 = time.perf_counter()
    xr, rcr, itr, histr = r.gmres(b_ref, m=80, max_iter=400, tol=1e-12)
    t_ref_gmres = time.perf_counter() - t0
    a = Aniso(sz, 1, 1, G_ANISO, SING_RULE, NP_ORDER, MAX_LEVEL)
    a.setCoeff(ss, cases.sigma_fields("layers", t["x"], t["y"], sz, 1)[1])
    a.cache(0)
    got = f - a.mapping(ss * f, 0)
    xg, rcg, itg, histg = a.gmres(a.mapping(f, 0), m=80, max_iter=400, tol=1e-12)
    a.delete()
    rel = lambda u, v: float(np.linalg.norm(u - v) / np.linalg.norm(v))  # noqa: E731
    out["parity"] = {"config": "N=%d (sz %d), same inputs through the %s and through libaniso_b200" % (n, sz, kind),
                     "operator_rel_l2": rel(got, t["out"]), "gmres_iters_ref": int(itr), "gmres_iters_gpu": int(itg),
                     "gmres_solution_rel_l2": rel(xg, xr), "ref_gmres_s": t_ref_gmres}
    return out


def block_section(sz, K=2, world=1, rank=0, new_uid=None, device=0):
    """the multi-mode block operator of the MATLAB layer (aniso.m forward / mforward / solve) on the same
    geometry with K = 2 angular modes (kernel ids 0..2, g = 0.9): device-resident mforward (regrouped tensor-core form and
    the loop of K(2K-1) mappings) and the block GMRES of aniso.m:158-172 (restart 400, tol 1e-11).  On world > 1 ranks the
    vectors are K blocks of owned rows; the checksums (norm of mforward's result, of the solution) let a sharded run be
    compared with the one-GPU line."""
    import ctypes as C

    import torch
    import torch.distributed as dist

    from aniso_b200 import Aniso, _lib

    a = Aniso(sz, 1, K, G_ANISO, SING_RULE, NP_ORDER, MAX_LEVEL, nranks=world, rank=rank, nccl_uid=new_uid() if world > 1 else None,
              device=device)
    L = a.L
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields("layers", x, y, sz, 1)
    a.setCoeff(ss, st)
    t0 = time.perf_counter()
    for k in range(2 * K - 1):
        a.cache(k)
    a.synchronize()
    t_cache = time.perf_counter() - t0
    slot0, slot1 = a.owned_range()
    nown = slot1 - slot0
    perm = a.slot_permutation()
    src_own = torch.from_numpy(cases.gaussian_source(x, y)[perm][slot0:slot1].copy()).cuda()
    du = src_own.repeat(K).contiguous()
    dout = torch.empty_like(du)

    def norm_all(t):
        v = (t * t).sum().reshape(1)
        if world > 1:
            dist.all_reduce(v)
        return float(v.sqrt().item())

    res = {"K": K, "kernel_ids": 2 * K - 1, "cache_all_ids_s": t_cache, "ranks": world}
    for key, env in (("mforward_ms", None), ("mforward_loop_ms", "1")):
        if env:
            os.environ["ANISO_BLOCK_LOOP"] = env
        else:
            os.environ.pop("ANISO_BLOCK_LOOP", None)
        for _ in range(3):
            _lib.check(L.aniso_block_forward_dev(a.h, du.data_ptr(), 1, dout.data_ptr()))
        a.synchronize()
        l0, t0, reps = a.launch_count(), time.perf_counter(), 10
        for _ in range(reps):
            _lib.check(L.aniso_block_forward_dev(a.h, du.data_ptr(), 1, dout.data_ptr()))
        a.synchronize()
        res[key] = (time.perf_counter() - t0) / reps * 1e3
        res[key.replace("_ms", "_launches")] = (a.launch_count() - l0) // reps
        res[key.replace("_ms", "_checksum")] = norm_all(dout)
    os.environ.pop("ANISO_BLOCK_LOOP", None)
    drhs, dx = torch.empty_like(du), torch.empty_like(du)
    it = C.c_int(0)
    hist = np.zeros(402)
    for _ in range(2):  # the first solve allocates the Krylov work space (restart 400: 403 vectors of K*N doubles)
        dx.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _lib.check(L.aniso_block_forward_dev(a.h, du.data_ptr(), 0, drhs.data_ptr()))  # rhs = forward(charge), aniso.m:160
        rc = _lib.check(L.aniso_block_gmres_dev(a.h, drhs.data_ptr(), dx.data_ptr(), 400, 400, 1e-11, C.byref(it), hist, hist.size))
        a.synchronize()
        t_solve = time.perf_counter() - t0
    res.update({"solve_s": t_solve, "solve_iters": int(it.value), "solve_rc": int(rc), "solve_final_resid": float(hist[it.value]),
                "solve_checksum": norm_all(dx),
                "solve_note": "device-resident: rhs = forward(charge), then GMRES(400, 1e-11, 400) on I - mforward (aniso.m:158-172)"})
    a.delete()
    return res


def n2_section(sz):
    """what an inverse-problem loop pays per setCoeff, and the memory / time trade of the cache-free near field.
    For both P2P modes (aniso_set_option "p2p" = "cached" | "onfly"): wall time from the setCoeff call to the first mapping
    result through the host API (sigma upload + tables + evaluator + stencils + first matvec + read-back), operator bytes
    held per kernel id, and the time of one application of the GMRES operator afterwards (CUDA events)."""
    from aniso_b200 import Aniso

    a = Aniso(sz, 1, 1, G_ANISO, SING_RULE, NP_ORDER, MAX_LEVEL)
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields("layers", x, y, sz, 1)
    f = cases.gaussian_source(x, y)
    a.setCoeff(ss, st)
    a.mapping(f, 0)  # first use of the handle: stream / table set-up and kernel loading stay out of the numbers below
    res = {}
    ref = None
    for mode in ("cached", "onfly"):
        a.set_option("p2p", mode)
        t, parts = [], None
        for _ in range(3):
            a.synchronize()
            t0 = time.perf_counter()
            a.setCoeff(ss, st)
            t1 = time.perf_counter()
            a.cache(0)
            t2 = time.perf_counter()
            out = a.mapping(f, 0)
            t3 = time.perf_counter()
            t.append((t3 - t0) * 1e3)
            if t[-1] == min(t):
                parts = {"set_coeff_ms": (t1 - t0) * 1e3, "cache_ms": (t2 - t1) * 1e3, "first_mapping_ms": (t3 - t2) * 1e3}
        ct = a.cache_times()
        res[mode] = {"set_coeff_to_first_mapping_ms": min(t), "split": parts, "operator_bytes_per_id": int(a.L.aniso_cache_bytes(a.h, 0)),
                     "matvec_ms": a.time_forward(20), "p2p_fill_ms": ct["p2p_fill_ms"], "m2l_fill_ms": ct["m2l_fill_ms"],
                     "stencil_ms": ct["stencil_ms"], "p2p_apply_ms": a.time_kernel(0, 1, 20)}
        if ref is None:
            ref = out
        else:
            res[mode]["rel_l2_vs_cached"] = float(np.linalg.norm(out - ref) / np.linalg.norm(ref))
    a.delete()
    return res


CONFIG_ROWS = [("C1", "c1", None), ("C2", "c2", None), ("C3", "c3", None), ("C4 np4", "c4n4", None),
               ("C4 np5", None, (512, 1, 1, 0.9, 10, 5, 20, "hetero")), ("C4 np6", None, (512, 1, 1, 0.9, 10, 6, 20, "hetero")),
               ("C4 np7", None, (512, 1, 1, 0.9, 10, 7, 20, "hetero")), ("C4 np8", "c4n8", None)]


def configs_section():
    """The other BASELINE configurations (C1 - C4, Chebyshev order sweep 4 - 8) on this GPU, kernel id 0: time of one
    application of the GMRES operator (CUDA events, device-resident), the M2L operator stream against the HBM roofline,
    the evaluator (cache phase) and the GMRES solve of main.cpp:141.  Where a fixture of the compiled reference exists
    (tests/golden/<name>.npz) the same run reports its distance from it."""
    from aniso_b200 import Aniso

    peak, _ = measured_peak()
    rows = {}
    for label, name, cfg in CONFIG_ROWS:
        sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name] if name else cfg
        t0 = time.perf_counter()
        a = Aniso(sz, deg, K, g, sr, np_, ml)
        setup = time.perf_counter() - t0
        x, y, w = a.nodes()
        ss, st = cases.sigma_fields(kind, x, y, sz, deg)
        a.setCoeff(ss, st)
        a.cache(0)
        ct, cnt, ab, tr = a.cache_times(), a.cache_counts(), a.apply_bytes(), a.tree_stats()
        r = np_ * np_
        m2l_ms, p2p_ms = a.time_kernel(0, 0, 20), a.time_kernel(0, 1, 20)
        m2l_alg = ab["m2l"] + 8 * r * tr["nodes"]
        b = a.mapping(cases.gaussian_source(x, y), 0)
        a.gmres(b)  # allocates the Krylov work space
        t0 = time.perf_counter()
        xs, rc, it, hist = a.gmres(b)
        t_gm = time.perf_counter() - t0
        row = {"N": int(x.size), "np": np_, "deg": deg, "setup_s": setup, "matvec_ms": a.time_forward(20),
               "m2l_stream_ms": m2l_ms, "m2l_stream_gbs": m2l_alg / m2l_ms / 1e6, "m2l_stream_frac_of_peak": m2l_alg / m2l_ms / 1e6 / peak,
               "p2p_stream_ms": p2p_ms, "p2p_fill_ms": ct["p2p_fill_ms"], "m2l_fill_ms": ct["m2l_fill_ms"], "stencil_ms": ct["stencil_ms"],
               "near_field_interactions_per_s": cnt["p2p"] / (ct["p2p_fill_ms"] * 1e-3) if ct["p2p_fill_ms"] else None,
               "gmres_iters": int(it), "gmres_rc": int(rc), "gmres_ms_host_api": t_gm * 1e3, "gmres_final_resid": float(hist[it])}
        gpath = os.path.join(ROOT, "tests", "golden", "%s.npz" % name) if name else None
        if gpath and os.path.exists(gpath):
            gd = np.load(gpath)
            stride = int(gd["stride"])
            row["parity"] = {"fixture": "tests/golden/%s.npz (compiled reference)" % name,
                             "gm_b_rel_l2": float(np.linalg.norm(b[::stride] - gd["gm_b"]) / np.linalg.norm(gd["gm_b"])),
                             "gmres_iters_ref": int(gd["gm_iters"]), "gmres_iters_gpu": int(it),
                             "gm_x_rel_l2": float(np.linalg.norm(xs[::stride] - gd["gm_x"]) / np.linalg.norm(gd["gm_x"]))}
        rows[label] = row
        a.delete()
    return rows


def run_ours(args):
    import torch
    import torch.distributed as dist

    from aniso_b200 import Aniso, _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    torch.cuda.set_device(local)
    L = _lib.load()
    uid = None

    def new_uid():  # a fresh NCCL unique id from rank 0, agreed by all ranks (one per communicator / handle)
        import ctypes

        buf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = ctypes.create_string_buffer(128)
            _lib.check(L.aniso_nccl_unique_id(raw))
            buf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        buf = buf.cuda()
        dist.broadcast(buf, 0)
        return bytes(buf.cpu().numpy().tobytes())

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = new_uid()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sz = args.sz
    t0 = time.perf_counter()
    a = Aniso(sz, 1, 1, G_ANISO, SING_RULE, NP_ORDER, MAX_LEVEL, nranks=world, rank=rank, nccl_uid=uid, device=local)
    t_setup = time.perf_counter() - t0
    n = a.numberOfNodes
    x, y, w = a.nodes()
    ss, st = cases.sigma_fields("layers", x, y, sz, 1)
    t0 = time.perf_counter()
    a.setCoeff(ss, st)
    a.cache(0)
    a.synchronize()
    t_cache = max_over_ranks(time.perf_counter() - t0)
    ct = a.cache_times()
    counts = a.cache_counts()
    slot0, slot1 = a.owned_range()
    nown = slot1 - slot0
    perm = a.slot_permutation()
    f_host = cases.gaussian_source(x, y)
    stream = torch.cuda.ExternalStream(L.aniso_stream(a.h))
    x_full = torch.from_numpy(f_host[perm]).cuda()          # slot order
    x_own = x_full[slot0:slot1].contiguous()
    y_own = torch.empty(nown, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()

    def step():
        _lib.check(L.aniso_forward_dev(a.h, x_own.data_ptr(), y_own.data_ptr()))

    # ---- device-resident timed region ----
    clk = ClockSampler(local)
    clk.__enter__()  # started before the warm-up: nvidia-smi needs up to a second before its first row on an 8-GPU box
    for _ in range(max(args.warmup, 3)):
        step()
    a.synchronize()
    barrier()
    l0 = a.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    time.sleep(0.12)
    clk.mark()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    e1.synchronize()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = a.launch_count() - l0
    tot = torch.tensor([float(counts["p2p"]), float(counts["m2l"]), ct["p2p_fill_ms"], ct["m2l_fill_ms"]], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = tot.clone()
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        tot[2:] = mx[2:]
    all_p2p, all_m2l, p2p_fill_ms, m2l_fill_ms = (float(v) for v in tot.tolist())
    ms_step = ms_total / args.steps
    value = n / (ms_step * 1e-3)

    # ---- end-to-end through the C-ABI with host buffers ----
    # host buffers in page-locked memory (aniso_host_alloc), as the e2e contract asks: the library DMAs them directly
    out_host = a.pinned_empty(n)
    charge = a.pinned_empty(n)
    charge[:] = f_host * ss
    _lib.check(L.aniso_mapping(a.h, charge, 0, out_host))
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _lib.check(L.aniso_mapping(a.h, charge, 0, out_host))
    a.synchronize()
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)

    # ---- parity of this run against the compiled reference at the same configuration (tests/golden/c5.npz) ----
    parity = None
    gpath = os.path.join(ROOT, "tests", "golden", "c5.npz")
    if sz == 1024 and os.path.exists(gpath):
        gd = np.load(gpath)
        stride = int(gd["stride"])
        f1 = cases.tilted_source(x, y)

        def gather_rows(own_dev):  # owned rows (slot order) of every rank -> full vector in reference order, on every rank
            full = torch.zeros(n, dtype=torch.float64, device="cuda")
            full[slot0:slot1] = own_dev
            if world > 1:
                dist.all_reduce(full)
            ref_order = np.empty(n)
            ref_order[perm] = full.cpu().numpy()
            return ref_order

        def sampled_err(got, key):
            return float(np.linalg.norm(got[::stride] - gd[key]) * np.sqrt(stride) / float(gd[key + "_norm"]))

        # (a) device-resident sharded matvec  (b) the host API (every rank fills its owned rows)
        f1_slots = torch.from_numpy(f1[perm]).cuda()
        o_own = torch.empty(nown, dtype=torch.float64, device="cuda")
        _lib.check(L.aniso_mapping_dev(a.h, f1_slots.data_ptr(), 0, o_own.data_ptr()))
        a.synchronize()
        dev_res = gather_rows(o_own)
        host_out = np.zeros(n)
        _lib.check(L.aniso_mapping(a.h, f1, 0, host_out))
        if world > 1:
            t = torch.from_numpy(host_out).cuda()
            dist.all_reduce(t)
            host_out = t.cpu().numpy()
        parity = {"fixture": "tests/golden/c5.npz: the compiled reference at this configuration (N=2^20), every %dth row" % stride,
                  "mapping_dev_rel_l2": sampled_err(dev_res, "f1_id0_out"), "mapping_host_api_rel_l2": sampled_err(host_out, "f1_id0_out"),
                  "gather_rows": gather_rows}

    # ---- roofline of the dominant kernel: the M2L streaming apply (bsr_apply_kernel over the V/X blocks) ----
    ab = a.apply_bytes()
    tr = a.tree_stats()
    m2l_ms = max_over_ranks(a.time_kernel(0, 0, 50))
    p2p_ms = max_over_ranks(a.time_kernel(0, 1, 50))
    # keep the device under the same load until the sampler has seen it on every rank (a 30-step timed region is 6 - 20 ms,
    # one sampler period is 50 ms); the number of extra steps is agreed between the ranks, step() holds a collective
    for batch in range(300):
        for _ in range(100):
            step()
        a.synchronize()
        have = torch.tensor([len(clk.rows) - clk.start], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(have, op=dist.ReduceOp.MIN)
        if batch >= 1 and int(have.item()) >= 3:
            break
    clk.__exit__()
    r = NP_ORDER * NP_ORDER
    m2l_alg = ab["m2l"] + 8 * r * tr["nodes"]  # stored blocks + scratch rows written + multipoles read
    peak, peak_src = measured_peak()
    roof = {"bound": "hbm", "kernel": "bsr_apply_sym_ring_kernel<16,1,3> (M2L stream over the V/X blocks, symmetric storage, TMA-fed ring)", "achieved": m2l_alg / (m2l_ms * 1e-3) / 1e9,
            "peak": peak, "unit": "GB/s", "frac": m2l_alg / (m2l_ms * 1e-3) / 1e9 / peak, "traffic": NCU_TRAFFIC.get((sz, world)),
            "peak_source": peak_src, "algorithmic_bytes_per_launch": m2l_alg, "ms_per_launch": m2l_ms,
            # the peak is a COPY (read + write); this kernel reads 97 % of its bytes, and a read-only stream on this pool reaches
            # 7177 GB/s (profiles/r01_fp64_peaks.json, scripts/fp64_peak.cu), so frac may slightly exceed 1
            "frac_note": "read-dominated kernel against a copy-measured peak; frac_of_read_only_stream uses 7177 GB/s measured with scripts/fp64_peak.cu",
            "frac_of_read_only_stream": m2l_alg / (m2l_ms * 1e-3) / 1e9 / 7177.0,
            "p2p_stream": {"algorithmic_bytes": ab["p2p"] + 8 * n, "ms_per_launch": p2p_ms,
                           "achieved": (ab["p2p"] + 8 * n) / (p2p_ms * 1e-3) / 1e9}}

    # ---- GMRES time-to-solve (main.cpp:141: GMRES(80, 400, 1e-12), x0 = 0) ----
    b_own = torch.empty(nown, dtype=torch.float64, device="cuda")
    _lib.check(L.aniso_mapping_dev(a.h, x_full.data_ptr(), 0, b_own.data_ptr()))
    xs = torch.zeros(nown, dtype=torch.float64, device="cuda")
    a.synchronize()
    barrier()
    import ctypes

    it = ctypes.c_int(0)
    hist = np.zeros(402)
    t_first = None
    for rep in range(2):  # the first solve allocates the Krylov work space (kept in the handle); the second is reported
        xs.zero_()
        a.synchronize()
        barrier()
        l1 = a.launch_count()
        t0 = time.perf_counter()
        rc = _lib.check(L.aniso_gmres_dev(a.h, b_own.data_ptr(), xs.data_ptr(), 80, 400, 1e-12, ctypes.byref(it), hist, hist.size))
        a.synchronize()
        barrier()
        t_gmres = max_over_ranks(time.perf_counter() - t0)
        gm_launches = a.launch_count() - l1
        if t_first is None:
            t_first = t_gmres

    if parity is not None:
        gather_rows = parity.pop("gather_rows")
        xs_ref = gather_rows(xs)
        b_ref = gather_rows(b_own)
        gd = np.load(gpath)
        stride = int(gd["stride"])
        parity.update({"gm_b_rel_l2": float(np.linalg.norm(b_ref[::stride] - gd["gm_b"]) / np.linalg.norm(gd["gm_b"])),
                       "gmres_iters_ref": int(gd["gm_iters"]), "gmres_iters_gpu": int(it.value),
                       "gmres_x_rel_l2": float(np.linalg.norm(xs_ref[::stride] - gd["gm_x"]) / np.linalg.norm(gd["gm_x"])),
                       "gmres_last_printed_resid_ref": float(gd["gm_hist"][-1]),
                       "gmres_same_step_resid_gpu": float(hist[min(int(gd["gm_iters"]), int(it.value)) - 1])})

    line = None
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "vs_baseline_note": "BASELINE.md holds no published number for this metric (the reference publishes none)",
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(sz), "l2": "operator stream %.2f GB per step > 126 MB L2, no flush needed"
                           % ((ab["p2p"] + ab["m2l"]) / 1e9), "partition": "target leaves by quadtree ownership, %d rank(s)" % world},
                "e2e": {"value": n / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 8 * ((n + world - 1) // world), "d2h_bytes_per_step": 8 * nown,
                        "ms_per_step": e2e_s * 1e3,
                        "note": "per rank: each rank uploads 1/n_gpus of the source (the pieces are all-gathered over NVLink) and reads back its owned rows"},
                "gpu_launches": int(launches), "clocks": clk.summary(), "roofline": roof,
                "near_field_interactions_per_s": all_p2p / (p2p_fill_ms * 1e-3) if p2p_fill_ms else None,
                "cache": {"wall_s": t_cache, "p2p_pairs": int(all_p2p), "m2l_pairs": int(all_m2l), "p2p_fill_ms": p2p_fill_ms,
                          "m2l_fill_ms": m2l_fill_ms, "stencil_ms": ct["stencil_ms"],
                          "m2l_evals_per_s": all_m2l / (m2l_fill_ms * 1e-3) if m2l_fill_ms else None},
                "gmres": {"rc": int(rc), "iters": int(it.value), "final_resid": float(hist[it.value]), "time_to_solve_s": t_gmres,
                          "first_call_s": t_first, "m": 80, "tol": 1e-12, "gpu_launches": int(gm_launches)},
                "setup_s": t_setup, "library_config": a.config()}
        if parity is not None:
            line["parity" if world == 1 else "parity_sharded"] = parity
        roof["traffic_source"] = ("dram__bytes_read.sum + dram__bytes_write.sum of the same kernel from the committed `ncu --set full` "
                                  "capture %s (cannot be measured live)" % NCU_TRAFFIC_SOURCE) if roof["traffic"] else None
    if not args.no_block and (world == 1 or args.block_sharded):
        a.delete()
        a = None
        blk = block_section(sz, world=world, rank=rank, new_uid=new_uid, device=local)
        if rank == 0:
            line["block_operator"] = blk
    if rank == 0:
        if world == 1 and not args.no_configs:
            if a is not None:
                a.delete()
                a = None
            line["configs"] = configs_section()
        if world == 1 and not args.no_configs:
            line["n2_variants"] = n2_section(sz)
        if world == 1 and not args.no_cpu:
            try:
                line["cpu_baseline"] = cpu_baseline(args)
            except Exception as e:  # the checker is optional equipment; never fail the GPU number for it
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                                        "sample": "failed: %r" % (e,)}
        print(json.dumps(line), flush=True)
    if a is not None:
        a.delete()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sz", type=int, default=1024, help="cells per side (N = sz^2); 1024 = BASELINE C5")
    ap.add_argument("--ref-sz", type=int, default=0, help="sample size of the reference arm (0 = by core count)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-block", action="store_true", help="skip the multi-mode block operator section")
    ap.add_argument("--block-sharded", action="store_true", help="run the block operator section on N > 1 ranks too")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1-C4 rows (N = 1 only)")
    ap.add_argument("--no-ref-gmres", action="store_true", help="reference arm: skip its GMRES solve")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
