#!/usr/bin/env python
"""bench.py -- single-wall fp64 Mf pair-interactions/s at N = 1M (BASELINE.json config 4).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--n N_BLOBS]

A step = one product u = M f over the whole synthetic blob cloud.  With N > 1
ranks the targets are sharded by rank, every rank owns the force slice of its
targets, and a step = NCCL all-gather of the slices + this rank's share of the
unordered pairs + all-reduce of the raw sums (strong scaling: total work N^2
fixed).  One JSON line is printed by rank 0.  Besides the headline it carries
sub-records for the other BASELINE configs: c2 (GMRES mobility solve, with the
reference's seconds on the same host), c3 (100k rollers, four blocks + fused),
c5 (Brownian step, 8 GPUs only), small_n (NumPy-boundary latency at N = 1,000 and
15,000 next to numba) and dense (the C1-style clouds).

--impl reference times the REFERENCE's own CPU implementation of the path:
mobility.single_wall_mobility_trans_times_force_numba (mobility/mobility.py:1132
-> mobility_numba.py:124) from the staged copy oracle/_ref, all host threads.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_PAIR = 200.0   # algorithmic flop of one single-wall UF pair as the oracle writes it (the figure
                        # the headline uses; tools/count_flops.py counts 211 from the reference source, frozen in
                        # profiles/flops_per_pair.json and reported as roofline.flop_per_pair_counted)


def counted_flop_per_pair():
    try:
        with open(os.path.join(ROOT, "profiles", "flops_per_pair.json")) as f:
            return json.load(f)["headline"]["single_wall_UF_counted"]
    except (OSError, KeyError, ValueError):
        return None


def sass_counts():
    """FP64-pipe instructions of the flagship's pair loop, counted from the built library's SASS by
    tools/loop_count.py at build time (profiles/sass_counts.json; tests/test_sass_counts.py re-counts the binary and
    fails when file and binary disagree)."""
    with open(os.path.join(ROOT, "profiles", "sass_counts.json")) as f:
        return json.load(f)["flagship"]


METRIC = "single-wall fp64 Mf pair-interactions/s at N=1M"


def make_cloud(n, seed=1234):
    """10% volume fraction random cloud, all z >= a; a, eta as mobility/test_blobs.py:30-44."""
    rng = np.random.default_rng(seed)
    a, eta = 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1.0 / 3.0) * a
    r = side * rng.random((n, 3))
    r[:, 2] += a
    f = rng.standard_normal((n, 3))
    return r, f, a, eta


def make_c2(n_bodies=1000, seed=77):
    """BASELINE config 2: boomerang_N_15 bodies on a grid with spacing 4.0, heights in [3,4),
    random unit quaternions, a = 0.25, eta = 1, tolerance 1e-8 (multi_bodies/inputfile.dat)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "rigid_golden.npz"))
    rng = np.random.default_rng(seed)
    side = int(np.ceil(np.sqrt(n_bodies)))
    k = np.arange(n_bodies)
    loc = np.stack([4.0 * (k % side), 4.0 * (k // side), 3.0 + rng.random(n_bodies)], axis=1)
    quat = rng.standard_normal((n_bodies, 4))
    quat /= np.linalg.norm(quat, axis=1)[:, None]
    FT = rng.standard_normal(6 * n_bodies)
    return g["struct/boomerang_N_15"], loc, quat, FT, 0.25, 1.0, 1e-8


def make_c3(n=100000, seed=3):
    """BASELINE config 3: single-blob rollers above the wall, pseudo-periodic in xy, in-plane packing
    fraction 0.4, heights in [a, 4.5a); a, eta of multi_bodies/examples/Mobility_Prod_Timing/Multiblob.inputfile:17-30."""
    rng = np.random.default_rng(seed)
    a, eta = 1.0155, 0.957e-3
    Lxy = np.sqrt(np.pi * a * a * n / 0.4)
    L = np.array([Lxy, Lxy, 0.0])
    r = np.stack([Lxy * rng.random(n), Lxy * rng.random(n), a + 3.5 * a * rng.random(n)], axis=1)
    return r, rng.standard_normal((n, 3)), rng.standard_normal((n, 3)), a, eta, L


def make_c5(n_bodies=10000, seed=55):
    """BASELINE config 5: shell_N_162 spheres (blob radius 0.26201755) on a square grid with spacing 3.0
    at z in [1.5, 2.5), random unit quaternions, random force-torque, eta = 1, kT = 0.0041, dt = 0.01.
    Returns (shell, a, eta, kT, dt, loc, quat, FT, z) -- the same arrays on every rank."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "rigid_golden.npz"))
    shell = g["struct/shell_N_162"]
    a = float(g["struct/shell_N_162_radius"][0])
    rng = np.random.default_rng(seed)
    side = int(np.ceil(np.sqrt(n_bodies)))
    k = np.arange(n_bodies)
    loc = np.stack([3.0 * (k % side), 3.0 * (k // side), 1.5 + rng.random(n_bodies)], axis=1)
    quat = rng.standard_normal((n_bodies, 4))
    quat /= np.linalg.norm(quat, axis=1)[:, None]
    FT = rng.standard_normal((n_bodies, 6))
    z = rng.standard_normal(3 * n_bodies * len(shell))
    return shell, a, 1.0, 0.0041, 0.01, loc, quat, FT, z


def make_dense_cloud(n, seed=4321):
    """The C1 cloud of mobility/test_blobs.py:30-34 (N = 1000 in a 5a box: ~16 % overlapping pairs, ~20 % of the
    blobs below z = a) scaled to n blobs at the SAME number density: box side 5a (n/1000)^(1/3), z from 0."""
    rng = np.random.default_rng(seed)
    a, eta = 0.13, 7.0
    side = 5.0 * a * (n / 1000.0) ** (1.0 / 3.0)
    return side * rng.random((n, 3)), rng.standard_normal((n, 3)), a, eta


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}




def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def timed_events(fn, reps, torch):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = None
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def sampled_rows(n, m, seed):
    """m random target indices and the permutation that moves them to the front (the product is equivariant under
    a permutation of the blobs, so the oracle's contiguous `targets` range serves random rows)"""
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(n, min(m, n), replace=False))
    rest = np.setdiff1d(np.arange(n), idx, assume_unique=True)
    return idx, np.concatenate([idx, rest])


# ------------------------------------------------------------------------------------------- CPU legs ----
def reference_engine():
    """The reference's own mobility module (numba kernels) from oracle/_ref, or None + the reason."""
    try:
        from oracle import make_ref, ref_driver
        ref_driver.use_all_cores()
        mb_ref = make_ref.import_mobility()
        if not hasattr(mb_ref, "mobility_numba"):
            return None, "reference mobility.py imported without its numba kernels (numba not importable)"
        return mb_ref, None
    except Exception as e:                              # noqa: BLE001 -- recorded in the JSON line
        return None, "%s: %s" % (type(e).__name__, e)


def cpu_rate_sample(mb_ref, n_sample, reps=1):
    """pairs/s of the CPU arm on make_cloud(n_sample) (same density as the workload): the reference's numba through
    its public wrapper when available, else the oracle's C/OpenMP port."""
    if mb_ref is not None:
        from oracle import ref_driver
        rate, sec = ref_driver.numba_uf_rate(n_sample, make_cloud, reps=reps, mb_ref=mb_ref)
        return rate, sec, ref_driver.numba_threads()
    import oracle
    cores = oracle.set_threads()
    r, f, a, eta = make_cloud(n_sample)
    oracle.uf_wall_fast(r[:256], f[:256], eta, a)
    best = 1e300
    for _ in range(reps):
        t0 = time.perf_counter()
        oracle.uf_wall_fast(r, f, eta, a)
        best = min(best, time.perf_counter() - t0)
    return float(n_sample) ** 2 / best, best, cores


def cpu_baseline_record(n_sample=100000):
    """cpu_baseline of the b200 line: the reference's numba at N = 1,000 / 10,000 / n_sample on this host."""
    mb_ref, why = reference_engine()
    rates = {}
    for n, reps in ((1000, 5), (10000, 3), (n_sample, 1)):
        rate, sec, cores = cpu_rate_sample(mb_ref, n, reps)
        rates[str(n)] = {"pairs_per_s": rate, "seconds": sec}
    rec = {"value": rates[str(n_sample)]["pairs_per_s"], "unit": "pairs/s", "cores": cores,
           "kind": "reference" if mb_ref is not None else "port",
           "engine": "mobility.single_wall_mobility_trans_times_force_numba (mobility/mobility.py:1132 -> "
                     "mobility_numba.py:124), unmodified, from oracle/_ref" if mb_ref is not None
                     else "oracle/mobility_oracle.c uf_wall_fast (C + OpenMP port)",
           "sample": "full product at N = %d blobs of the same 10 %% volume-fraction cloud (%.1f s); the rate is flat in "
                     "N above ~10k (rates)" % (n_sample, rates[str(n_sample)]["seconds"]),
           "rates": rates}
    if why:
        rec["reference_unavailable"] = why
    return rec, mb_ref


def c2_cpu_seconds():
    """The reference's own mobility solve of config 2 on this host (oracle/ref_driver.mobility_problem)."""
    try:
        from oracle import ref_driver
        ref_cfg, loc, quat, FT, a, eta, tol = make_c2(1000)
        out = ref_driver.mobility_problem(ref_cfg, loc, quat, FT, a, eta, tol, implementation="numba")
        return {"cpu_seconds": out["seconds_pc"] + out["seconds_gmres"], "cpu_seconds_pc": out["seconds_pc"],
                "cpu_seconds_gmres": out["seconds_gmres"], "cpu_iterations": out["iterations"],
                "cpu_true_residual": out["residual"], "cpu_cores": ref_driver.numba_threads(),
                "cpu_kind": "reference: multi_bodies.linear_operator_rigid + build_block_diagonal_preconditioner + "
                            "numba product under scipy.sparse.linalg.gmres (oracle/_ref)"}, out["solution"]
    except Exception as e:                              # noqa: BLE001
        return {"cpu_seconds": None, "cpu_unavailable": "%s: %s" % (type(e).__name__, e)}, None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    mb_ref, why = reference_engine()
    n_sample = min(args.n, args.ref_sample)
    rates, secs, cores = [], [], 0
    for it in range(args.warmup + args.steps):
        rate, sec, cores = cpu_rate_sample(mb_ref, n_sample, reps=1)
        if it >= args.warmup:
            rates.append(rate)
            secs.append(sec)
    v = float(np.mean(rates))
    sample = "each step = the full product at N = %d blobs of the same cloud (%.2f s per step)" % (n_sample, np.mean(secs))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(args.n) ** 2 / v,
        "extrapolated": n_sample != args.n, "sample_n": n_sample,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C4: single-wall UF, N=%d random blobs, 10%% volume fraction, open boundaries" % args.n,
                   "note": "ms_per_step = N^2 / value, extrapolated from the sample (the CPU would need minutes per "
                           "product at N = 1M)"},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores,
                         "kind": "reference" if mb_ref is not None else "port",
                         "engine": "numba: mobility.single_wall_mobility_trans_times_force_numba, unmodified (oracle/_ref)"
                         if mb_ref is not None else "oracle/mobility_oracle.c uf_wall_fast (C + OpenMP port)",
                         "sample": sample},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if why:
        line["cpu_baseline"]["reference_unavailable"] = why
    if not args.no_solve:
        rec, _ = c2_cpu_seconds()
        line["gmres_mobility_solve"] = {"seconds": rec.get("cpu_seconds"), "unit": "s", "bodies": 1000, "blobs": 15000,
                                        "iterations": rec.get("cpu_iterations"), "tolerance": 1e-8,
                                        "true_residual": rec.get("cpu_true_residual"), "cores": rec.get("cpu_cores"),
                                        "includes": "block-diagonal preconditioner build + SciPy GMRES(60), reference code"}
    emit(line)


# ------------------------------------------------------------------------------------- GPU sub-records ----
def gmres_mobility_solve(n_bodies=1000, repeats=3):
    """Secondary metric: seconds of the config-2 mobility solve (preconditioner build + GMRES), device resident,
    timed with CUDA events after one warm-up solve; parity against the reference's own answer at this size."""
    import torch
    from rigidmultiblobswall_b200 import rigid
    ref_cfg, loc, quat, FT, a, eta, tol = make_c2(n_bodies)
    bodies = rigid.RigidBodies(ref_cfg, loc, quat)
    out = None
    times = []
    for it in range(repeats + 1):
        ms, out = timed_events(lambda: rigid.solve_mobility_problem(bodies, FT, eta, a, tolerance=tol), 1, torch)
        if it > 0:
            times.append(ms * 1e-3)
    vel, lam, its, info = out
    r = bodies.get_r_vectors()
    x = torch.cat([lam.reshape(-1), vel.reshape(-1)])
    rhs = torch.cat([torch.zeros(3 * bodies.n_blobs, dtype=torch.float64, device=x.device),
                     -torch.from_numpy(FT).to(x.device)])
    res = float(torch.linalg.norm(rigid.linear_operator_rigid(x, bodies, r, eta, a) - rhs) / torch.linalg.norm(rhs))
    rec = {"seconds": float(np.median(times)), "unit": "s", "bodies": n_bodies, "blobs": bodies.n_blobs,
           "iterations": int(its), "info": int(info), "tolerance": tol, "true_residual": res,
           "includes": "block-diagonal preconditioner build + GMRES(60), device resident"}
    gpath = os.path.join(ROOT, "tests", "golden", "configs_golden.npz")
    if n_bodies == 1000 and os.path.exists(gpath):
        g = np.load(gpath)
        v = vel.reshape(-1).cpu().numpy()
        rec["parity_velocity_rel_err_vs_reference_golden"] = float(np.linalg.norm(v - g["c2/velocity"]) /
                                                                   np.linalg.norm(g["c2/velocity"]))
        rec["reference_golden_iterations"] = int(g["c2/iterations"][0])
    return rec, vel.reshape(-1).cpu().numpy()


def c3_record(n=100000):
    """BASELINE config 3: the four blocks + the fused UF f + UT t for n rollers, 9 xy image boxes; CUDA-event ms and
    256 random target rows of each block against the oracle."""
    import torch
    import oracle
    from rigidmultiblobswall_b200 import mobility as mb
    r, f, t, a, eta, L = make_c3(n)
    idx, perm = sampled_rows(n, 256, seed=31)
    rd, fd, td = (torch.from_numpy(x).cuda().reshape(-1) for x in (r, f, t))
    pairs = 9.0 * n * n
    rec = {"config": "C3: %d rollers, a=%g, L=[%.1f,%.1f,0], 9 image boxes, fp64" % (n, a, L[0], L[1]), "blocks": {}}
    refs = {}
    oracle.set_threads()
    for name, blk, vec, vh in (("UF", mb.UF, fd, f), ("UT", mb.UT, td, t), ("WF", mb.WF, fd, f), ("WT", mb.WT, td, t)):
        fn = lambda: mb.mobility_product(blk, mb.SINGLE_WALL, rd, vec, eta, a, periodic_length=L)
        fn()
        ms, u = timed_events(fn, 3, torch)
        refs[name] = oracle.mobility_product(name, "single_wall", r[perm], vh[perm], eta, a, L, targets=(0, len(idx)))
        got = u.reshape(-1, 3)[torch.from_numpy(idx).cuda()].reshape(-1).cpu().numpy()
        rec["blocks"][name] = {"ms": ms, "pairs_per_s": pairs / ms * 1e3,
                               "rel_err_256_random_rows_vs_oracle": float(np.linalg.norm(got - refs[name]) /
                                                                          np.linalg.norm(refs[name]))}
    fn = lambda: mb.mobility_product(mb.UF_UT, mb.SINGLE_WALL, rd, fd, eta, a, vec_b=td, periodic_length=L)
    fn()
    ms, u = timed_events(fn, 3, torch)
    ref = refs["UF"] + refs["UT"]
    got = u.reshape(-1, 3)[torch.from_numpy(idx).cuda()].reshape(-1).cpu().numpy()
    rec["blocks"]["UF+UT fused"] = {"ms": ms, "pairs_per_s": pairs / ms * 1e3,
                                    "rel_err_256_random_rows_vs_oracle": float(np.linalg.norm(got - ref) / np.linalg.norm(ref))}
    rec["four_blocks_seconds"] = sum(rec["blocks"][k]["ms"] for k in ("UF", "UT", "WF", "WT")) * 1e-3
    # the caller of these blocks: the rollers' deterministic step (quaternion_integrator_rollers.py:905-979), device resident
    from rigidmultiblobswall_b200 import rollers
    omega, tol = np.array([0.0, 62.8, 0.0]), 1e-8
    step = lambda: rollers.compute_deterministic_velocity_and_torque(rd, fd, omega, eta, a, periodic_length=L, tolerance=tol)
    step()
    ms, (v, T, its, info) = timed_events(step, 1, torch)
    res = torch.from_numpy(np.tile(omega, n)).cuda() - mb.mobility_product(mb.WF, mb.SINGLE_WALL, rd, fd, eta, a, periodic_length=L) \
        - mb.mobility_product(mb.WT, mb.SINGLE_WALL, rd, T, eta, a, periodic_length=L)
    rec["deterministic_step"] = {"seconds": ms * 1e-3, "gmres_iterations": int(its), "info": int(info), "tolerance": tol,
                                 "true_residual": float(torch.linalg.norm(res) / np.linalg.norm(np.tile(omega, n))),
                                 "what": "solve M_rr T = omega - M_rt f (GMRES(20), no preconditioner), then v = M_tt f + M_tr T "
                                         "in one fused pass; parity against the reference's own method: tests/test_rollers_gpu.py"}
    rec["parity_max_rel_err"] = max(b["rel_err_256_random_rows_vs_oracle"] for b in rec["blocks"].values())
    return rec


def small_n_record(mb_ref):
    """Latency of ONE call through the NumPy drop-in boundary (what SciPy GMRES makes at every iteration) at the sizes
    of configs 1 and 2, next to the reference's numba wrapper on this host."""
    import torch
    from rigidmultiblobswall_b200 import mobility as mb
    rec = {}
    for n in (1000, 15000):
        r, f, a, eta = make_cloud(n)
        L = np.zeros(3)
        fn = lambda: mb.single_wall_mobility_trans_times_force_b200(r, f, eta, a, periodic_length=L)
        for _ in range(5):
            fn()
        ts = []
        for _ in range(50):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
        dev_fn = lambda: mb.single_wall_mobility_trans_times_force_b200(rd, fd, eta, a)
        dev_fn()
        ms_dev, _ = timed_events(dev_fn, 50, torch)
        e = {"b200_numpy_boundary_us": 1e6 * float(np.median(ts)), "b200_device_resident_us": 1e3 * ms_dev}
        if mb_ref is not None:
            cf = lambda: mb_ref.single_wall_mobility_trans_times_force_numba(r, f, eta, a, periodic_length=L)
            cf()
            tc = []
            for _ in range(7):
                t0 = time.perf_counter()
                cf()
                tc.append(time.perf_counter() - t0)
            e["reference_numba_us"] = 1e6 * float(np.median(tc))
        rec[str(n)] = e
    return rec


def dense_record(n=200000):
    """Throughput where the near-field branch matters: the C1 cloud of mobility/test_blobs.py:30-34 at the same
    number density (`same_density`), and all n blobs inside the C1 5a box (`c1_box`: 16 % of ALL pairs overlap and
    20 % of the blobs lie below z = a, so the near-field select path runs in every warp at every step)."""
    import torch
    from rigidmultiblobswall_b200 import mobility as mb
    rec = {"n": n}
    rng = np.random.default_rng(7)
    clouds = {"same_density": make_dense_cloud(n),
              "c1_box": (5 * 0.13 * rng.random((n, 3)), rng.standard_normal((n, 3)), 0.13, 7.0)}
    for key, (r, f, a, eta) in clouds.items():
        rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
        fn = lambda: mb.single_wall_mobility_trans_times_force_b200(rd, fd, eta, a)
        fn()
        ms, _ = timed_events(fn, 3, torch)
        rec[key + "_pairs_per_s"] = float(n) * n / ms * 1e3
    return rec


def fp32_record(n=400000):
    """The fp32 variant (north_star): single-wall UF with fp32 pair arithmetic (pair-symmetric kernel by default, row
    kernel with RMB_DETERMINISTIC), their rates, the FP32 FMA probe of this GPU, and 256 random rows against the fp64
    oracle (bar: 1e-5)."""
    import torch
    import oracle
    from rigidmultiblobswall_b200 import _lib, mobility as mb
    r, f, a, eta = make_cloud(n)
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    fn = lambda: mb.mobility_product(mb.UF, mb.SINGLE_WALL, rd, fd, eta, a, precision=mb.FP32)
    fn()
    ms, u = timed_events(fn, 3, torch)
    idx, perm = sampled_rows(n, 256, seed=51)
    oracle.set_threads()
    ref = oracle.uf_wall_fast(r[perm], f[perm], eta, a, targets=(0, len(idx)))
    got = u.reshape(-1, 3)[torch.from_numpy(idx).cuda()].reshape(-1).cpu().numpy()
    peak32, _ = _lib.fma_peak(_lib.FP32, 200000)
    row = lambda: mb.mobility_product(mb.UF, mb.SINGLE_WALL, rd, fd, eta, a, precision=mb.FP32, flags=mb.DETERMINISTIC)
    row()
    ms_row, _ = timed_events(row, 3, torch)
    return {"n": n, "kernel": "sym_kernel<UF, ..., float> (fp32 pair arithmetic, every unordered pair once, fp64 sums)",
            "pairs_per_s": float(n) * n / ms * 1e3, "row_kernel_pairs_per_s": float(n) * n / ms_row * 1e3,
            "rel_err_256_random_rows_vs_oracle": float(np.linalg.norm(got - ref) / np.linalg.norm(ref)),
            "fp32_fma_probe_tflops": peak32 / 1e12,
            "ncu": "row kernel, profiles/r02b_uf_f32_ncu_summary.txt: 76.3 warp instructions per pair, issue slots "
                   "88.7 % busy (the bound: one instruction per scheduler per clock), FMA pipe 74.4 %"}


def c5_record(world, rank, dev, n_bodies=10000, tol=1e-6, lanczos_tol=1e-4):
    """BASELINE config 5: one Brownian step (preconditioner, Lanczos M^{1/2} z, GMRES with the random slip) for
    n_bodies shell_N_162 spheres sharded by body over the ranks.  Seconds are CUDA-event times, max over ranks."""
    import torch
    import torch.distributed as dist
    from rigidmultiblobswall_b200 import parallel
    shell, a, eta, kT, dt, loc, quat, FT, z_full = make_c5(n_bodies)

    def timed(fn):
        if world > 1:
            dist.barrier()
        ms, out = timed_events(fn, 1, torch)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return out, float(t.item()) * 1e-3

    system, _ = timed(lambda: parallel.ShardedRigidSystem(shell, loc, quat, eta, a,
                                                          group=dist.group.WORLD if world > 1 else False))
    o, c = system.ranges[system.rank]
    z_loc = torch.from_numpy(z_full[3 * o:3 * (o + c)].copy()).to(dev)
    system.mobility(z_loc)
    _, t_mf = timed(lambda: system.mobility(z_loc))
    _, t_pc = timed(system.build_preconditioner)
    (noise, it_l), t_l = timed(lambda: system.stochastic_noise(z_loc, factor=np.sqrt(2 * kT / dt), tolerance=lanczos_tol))
    sol, t_g = timed(lambda: system.solve_mobility(system.local_slice(FT), tolerance=tol, slip_loc=-noise))
    U_loc, lam_loc, its, info = sol
    x = torch.cat([lam_loc.reshape(-1), U_loc.reshape(-1)])
    rhs = torch.cat([-noise.reshape(-1), -system.local_slice(FT)])
    rr = system.linear_operator(x) - rhs
    nd = torch.stack([torch.dot(rr, rr), torch.dot(rhs, rhs)])
    if world > 1:
        dist.all_reduce(nd)
    return {"config": "C5: Brownian step, %d spheres x 162 blobs = %d blobs, single wall, fp64" % (n_bodies, system.n_blobs_total),
            "n_gpus": world, "seconds": t_pc + t_l + t_g, "seconds_preconditioner": t_pc, "seconds_lanczos": t_l,
            "seconds_gmres": t_g, "seconds_one_Mf": t_mf, "pairs_per_s_one_Mf": float(system.n_blobs_total) ** 2 / t_mf,
            "lanczos_iterations": int(it_l), "lanczos_tolerance": lanczos_tol, "gmres_iterations": int(its),
            "gmres_info": int(info), "gmres_tolerance": tol, "true_residual": float(torch.sqrt(nd[0] / nd[1]).item()),
            "parity": "reference answers at 128 and 1,000 spheres: tests/test_configs_gpu.py (the reference needs "
                      "~1 h at 1,000 spheres, tests/golden/make_golden_configs.py)"}


# ------------------------------------------------------------------------------------------- B200 arm ----
def run_b200(args):
    import torch
    import torch.distributed as dist
    from rigidmultiblobswall_b200 import _lib, mobility as mb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n = args.n
    r_np, f_np, a, eta = make_cloud(n)
    # contiguous target ranges, one per rank
    cuts = [(n * k) // world for k in range(world + 1)]
    off, cnt = cuts[rank], cuts[rank + 1] - cuts[rank]

    r_d = torch.from_numpy(r_np).to(dev).reshape(-1)
    f_full = torch.from_numpy(f_np).to(dev).reshape(-1)
    f_loc = f_full[3 * off:3 * (off + cnt)].clone()
    u_loc = torch.empty(3 * cnt, dtype=torch.float64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    # targets sharded by rank; the pair-symmetric schedule deals the unordered pairs to the ranks, every rank
    # all-gathers the force slices, accumulates raw sums for all blobs, all-reduces them and finishes its shard
    from rigidmultiblobswall_b200 import parallel
    sm = parallel.ShardedMobility(r_d, eta, a, group=dist.group.WORLD if world > 1 else False,
                                  ranges=[(cuts[k], cuts[k + 1] - cuts[k]) for k in range(world)])
    assert sm.symmetric or n < 8000

    def step():
        flush.zero_()                       # L2 flush between iterations (inside the timed region; ~0.1 ms)
        u_loc.copy_(sm(f_loc))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # FP64 FMA-pipe peak of this GPU, measured now (MEASURED_PEAKS.json has no FP64 figure)
    peak_flops, _ = _lib.fma_peak(_lib.FP64, 100000)

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(step, args.steps)
    clocks = sampler.stop()
    pairs_per_step = float(n) * float(n)
    value = pairs_per_step * args.steps / (ms * 1e-3)

    # parity in the same run: 1,000 random targets spread over EVERY rank's shard against the oracle
    parity = None
    if not args.no_parity:
        idx, perm = sampled_rows(n, 1000, seed=41)
        mine = idx[(idx >= off) & (idx < off + cnt)]
        rows = u_loc.reshape(-1, 3)[torch.from_numpy(mine - off).to(dev)].reshape(-1)
        if world > 1:
            counts = [int(((idx >= cuts[k]) & (idx < cuts[k + 1])).sum()) for k in range(world)]
            pad = torch.zeros(3 * max(counts), dtype=torch.float64, device=dev)
            pad[:rows.numel()] = rows
            parts = [torch.empty_like(pad) for _ in range(world)]
            dist.all_gather(parts, pad)
            rows = torch.cat([parts[k][:3 * counts[k]] for k in range(world)])
        if rank == 0:
            import oracle
            oracle.set_threads()
            ref = oracle.uf_wall_fast(r_np[perm], f_np[perm], eta, a, targets=(0, len(idx)))
            got = rows.cpu().numpy()
            parity = {"rel_err": float(np.linalg.norm(got - ref) / np.linalg.norm(ref)), "targets": int(len(idx)),
                      "targets_per_rank": [int(((idx >= cuts[k]) & (idx < cuts[k + 1])).sum()) for k in range(world)],
                      "against": "oracle (C restatement of mobility_numba.py:124-287), all %d sources" % n}

    # end to end through the public NumPy API: host buffers in, host result out, every step
    r_host = torch.from_numpy(r_np).pin_memory().numpy()
    f_host = torch.from_numpy(f_np).pin_memory().numpy()

    acc_e2e = [None]

    def e2e_step():
        if world == 1:
            return mb.mobility_product(mb.UF, mb.SINGLE_WALL, r_host, f_host, eta, a)     # NumPy in, NumPy out
        r_t = torch.from_numpy(r_host).to(dev, non_blocking=True)
        f_t = torch.from_numpy(f_host).to(dev, non_blocking=True)
        acc_e2e[0] = mb.symmetric_partial(r_t, f_t, eta, a, rank=rank, world=world, acc=acc_e2e[0])
        dist.all_reduce(acc_e2e[0])
        return mb.symmetric_finish(r_t, f_t, acc_e2e[0], eta, a, targets=(off, cnt)).cpu().numpy()

    e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    e2e_value = pairs_per_step * args.steps / (ms_e2e * 1e-3)

    single = rank == 0 and world == 1
    cpu, mb_ref = (None, None)
    if single and not args.no_cpu:
        cpu, mb_ref = cpu_baseline_record(min(n, args.ref_sample))

    solve = c3 = small = dense = fp32 = None
    if single and not args.no_solve:
        solve, _ = gmres_mobility_solve()
        if not args.no_cpu:
            solve.update(c2_cpu_seconds()[0])
    if single and not args.no_extra:
        c3 = c3_record()
        small = small_n_record(mb_ref)
        dense = dense_record()
        fp32 = fp32_record()
    c5 = None
    if world == 8 and not args.no_extra:
        c5 = c5_record(world, rank, dev)

    if rank == 0:
        # dominant kernel = sym_kernel<UF, TMA>; one launch per step per rank; the step's other launches
        # (L2 flush memset, prepare_kernel, accumulator memset, sym_finish_kernel) are < 0.1% of it
        launch_ms = ms / args.steps
        sc = sass_counts()
        per_int = float(sc["fp64_per_interaction"])       # FP64-pipe instructions per interaction (far pair / 2)
        executed = value / world * per_int * 2 / 1e12       # this rank: executed FP64 instr/s as TFLOP/s at 2 flop each
        algorithmic = (float(n) * n / world * FLOP_PER_PAIR) / (launch_ms * 1e-3) / 1e12
        plan_launches = int(_lib.load().rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP64, 0, n, n, None))
        sms = _lib.load().rmb_device_sm_count()
        nominal = sms * 64 * 2 * (clocks.get("sm_mhz") or 1965.0) * 1e6 / 1e12   # 64 FP64 FMA lanes / clk / SM
        traffic, traffic_note = None, None
        tpath = os.path.join(ROOT, "profiles", "flagship_traffic.json")
        if os.path.exists(tpath) and n == 1000000 and world == 1:
            tj = json.load(open(tpath))
            if tj.get("sass_loop") == {"total": sc["total"], "fp64": sc["fp64"]}:
                traffic = tj["dram_bytes_total"]
            else:
                traffic_note = "profiles/flagship_traffic.json was captured for another build of the kernel (stale)"
        frac = executed / (peak_flops / 1e12)
        # (under a profiler the probe and the kernel are replayed and serialised: their ratio means nothing)
        assert frac <= 1.0 or args.profiling, "roofline.frac = %.3f > 1: executed FP64 rate above the measured FMA peak" % frac
        line = {
            "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C4: single-wall UF, N=%d random blobs, 10%% volume fraction, open boundaries" % n,
                       "kernel": "pair-symmetric (every unordered pair once, both blobs updated)",
                       "sharding": "unordered pairs dealt to ranks by row group; per step NCCL all-gather of force "
                                   "slices + all-reduce of the raw sums (24 N bytes); each rank finishes its targets"
                       if world > 1 else "single GPU",
                       "l2": "flushed between iterations (256 MiB memset inside the timed region)"},
            "roofline": {"bound": "fp64", "achieved": executed, "peak": peak_flops / 1e12, "unit": "TFLOP/s",
                         "frac": frac, "traffic": traffic,
                         "fp64_instr_per_interaction": per_int,
                         "fp64_instr_source": "profiles/sass_counts.json (tools/loop_count.py from the built library's "
                                              "SASS: %d FP64 of %d instructions per unordered far pair)" %
                                              (sc["fp64_far_pair"], sc["total"] - sc["fp64_near_path"]),
                         "peak_source": "rmb_fma_peak (FP64 FMA probe, measured in this run on this GPU)",
                         "nominal_peak": nominal, "frac_of_nominal": executed / nominal,
                         "algorithmic_tflops": algorithmic, "algorithmic_frac": algorithmic / (peak_flops / 1e12),
                         "flop_per_pair": FLOP_PER_PAIR, "flop_per_pair_counted": counted_flop_per_pair(),
                         "note": "achieved = FP64-pipe instructions the kernel EXECUTES per second x 2 flop (an FMA), "
                                 "this rank; frac = achieved / the FMA probe's rate, <= 1 by construction.  "
                                 "algorithmic_* credit the reference's 200 flop per ordered pair (SURVEY 8d) to a "
                                 "kernel that evaluates each unordered pair once: an algorithmic saving, not a "
                                 "hardware fraction.  traffic = DRAM bytes of one launch from ncu "
                                 "(profiles/flagship_traffic.json); algorithmic bytes are 72 B per blob"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "pairs/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(world * 2 * 24 * n), "d2h_bytes_per_step": int(24 * n)},
            "gpu_launches": int(args.steps * plan_launches * world),
            "clocks": clocks,
            "parity": parity,
            "gmres_mobility_solve": solve,
            "c3": c3, "c5": c5, "small_n": small, "dense": dense, "fp32": fp32,
        }
        if traffic_note:
            line["roofline"]["traffic_note"] = traffic_note
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def emit(line):
    """The ONE JSON line, on the real stdout (fd 1 is pointed at stderr while the run is in progress so
    that library chatter such as NCCL's version banner cannot land in front of it)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=1000000)
    ap.add_argument("--ref-sample", type=int, default=100000, help="blobs of the CPU arm's bounded sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-solve", action="store_true", help="skip the config-2 GMRES mobility solve")
    ap.add_argument("--no-extra", action="store_true", help="skip the c3 / c5 / small_n / dense / fp32 sub-records")
    ap.add_argument("--profiling", action="store_true", help="run under ncu: timings are not measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
