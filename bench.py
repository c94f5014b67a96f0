#!/usr/bin/env python
"""bench.py -- single-wall fp64 Mf pair-interactions/s at N = 1M (BASELINE.json config 4).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--n N_BLOBS]

A step = one product u = M f over the whole synthetic blob cloud.  With N > 1
ranks the targets are sharded by rank, every rank owns the force slice of its
targets, and a step = NCCL all-gather of the slices + the local product
(strong scaling: total work N^2 fixed).  One JSON line is printed by rank 0.
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
FP64_INSTR_PER_PAIR = 43.5  # FP64-pipe instructions per interaction of the pair-symmetric kernel: 87 per unordered
                            # far pair = two interactions (SASS count); the row kernel executes 66
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


def gmres_mobility_solve(n_bodies=1000, repeats=3):
    """Secondary metric: wall seconds of the config-2 mobility solve (preconditioner build + GMRES), device
    resident, timed with CUDA events after one warm-up solve."""
    import torch
    from rigidmultiblobswall_b200 import rigid
    ref_cfg, loc, quat, FT, a, eta, tol = make_c2(n_bodies)
    bodies = rigid.RigidBodies(ref_cfg, loc, quat)
    out = None
    times = []
    for it in range(repeats + 1):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        vel, lam, its, info = rigid.solve_mobility_problem(bodies, FT, eta, a, tolerance=tol)
        e1.record()
        torch.cuda.synchronize()
        if it > 0:
            times.append(e0.elapsed_time(e1) * 1e-3)
        out = (vel, lam, its, info)
    vel, lam, its, info = out
    r = bodies.get_r_vectors()
    x = torch.cat([lam.reshape(-1), vel.reshape(-1)])
    rhs = torch.cat([torch.zeros(3 * bodies.n_blobs, dtype=torch.float64, device=x.device),
                     -torch.from_numpy(FT).to(x.device)])
    res = float(torch.linalg.norm(rigid.linear_operator_rigid(x, bodies, r, eta, a) - rhs) / torch.linalg.norm(rhs))
    return {"seconds": float(np.median(times)), "unit": "s", "bodies": n_bodies, "blobs": bodies.n_blobs,
            "iterations": int(its), "info": int(info), "tolerance": tol, "true_residual": res,
            "includes": "block-diagonal preconditioner build + GMRES(60), device resident"}


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


def cpu_reference_rate(r, f, a, eta, budget_s=12.0):
    """The oracle's vectorised single-wall UF (C/OpenMP port of mobility_numba.py:124-287, all host threads) on a
    bounded sample: the first m targets against ALL sources."""
    import oracle
    n = r.shape[0]
    m = min(n, 256)
    oracle.uf_wall_fast(r, f, eta, a, targets=(0, m))          # thread-pool warm-up
    t0 = time.perf_counter()
    oracle.uf_wall_fast(r, f, eta, a, targets=(0, m))
    dt = time.perf_counter() - t0
    rate = m * n / dt
    m2 = int(min(n, max(m, budget_s * rate / n)))
    t0 = time.perf_counter()
    oracle.uf_wall_fast(r, f, eta, a, targets=(0, m2))
    dt = time.perf_counter() - t0
    return m2 * n / dt, oracle.threads(), "targets [0,%d) x all %d sources, %.1f s" % (m2, n, dt)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r, f, a, eta = make_cloud(args.n)
    times, sample, cores = [], "", 0
    for it in range(args.warmup + args.steps):
        rate, cores, sample = cpu_reference_rate(r, f, a, eta, budget_s=args.ref_budget)
        if it >= args.warmup:
            times.append(rate)
    v = float(np.mean(times))
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * args.n * args.n / v,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C4: single-wall UF, N=%d random blobs, 10%% volume fraction, open boundaries" % args.n,
                   "note": "CPU port of mobility_numba (oracle/, C+OpenMP); ms_per_step extrapolated from the sample"},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


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
    equal = all(cuts[k + 1] - cuts[k] == cnt for k in range(world))

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
    assert sm.symmetric or n < 16384

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

    # parity spot check in the same run: 48 rows of this rank's shard against the oracle
    parity = None
    if rank == 0 and not args.no_parity:
        import oracle
        b = off + min(cnt - 48, 1234) if cnt > 48 else off
        e = min(b + 48, off + cnt)
        ref = oracle.mobility_product("UF", "single_wall", r_np, f_np, eta, a, targets=(b, e))
        got = u_loc[3 * (b - off):3 * (e - off)].cpu().numpy()
        parity = float(np.linalg.norm(got - ref) / np.linalg.norm(ref))

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

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, cores, sample = cpu_reference_rate(r_np, f_np, a, eta, budget_s=args.ref_budget)
        cpu = {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": sample}

    solve = None
    if rank == 0 and world == 1 and not args.no_solve:
        solve = gmres_mobility_solve()

    if rank == 0:
        # dominant kernel = sym_kernel<UF, TMA>; one launch per step per rank; the step's other launches
        # (L2 flush memset, prepare_kernel, accumulator memset, sym_finish_kernel) are < 0.1% of it
        launch_ms = ms / args.steps
        achieved = (float(n) * n / world * FLOP_PER_PAIR) / (launch_ms * 1e-3) / 1e12   # this rank's share of the pairs
        plan_launches = int(_lib.load().rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP64, 0, n, n, None))
        sms = _lib.load().rmb_device_sm_count()
        nominal = sms * 64 * 2 * (clocks.get("sm_mhz") or 1965.0) * 1e6 / 1e12   # 64 FP64 FMA lanes / clk / SM
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "flagship_traffic.json")
        if os.path.exists(tpath) and n == 1000000 and world == 1:
            traffic = json.load(open(tpath))["dram_bytes_total"]
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
            "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak_flops / 1e12, "unit": "TFLOP/s",
                         "frac": achieved / (peak_flops / 1e12), "traffic": traffic,
                         "flop_per_pair": FLOP_PER_PAIR, "flop_per_pair_counted": counted_flop_per_pair(),
                         "peak_source": "rmb_fma_peak (FP64 FMA probe, measured in this run on this GPU)",
                         "note": "achieved counts the reference's 200 flop per pair (SURVEY 8d); the kernel executes "
                                 "fp64_instr_per_pair FP64-pipe instructions per pair, pipe_frac = those lanes/s over "
                                 "the nominal issue rate (SMs x 64 lanes x SM clock); traffic = DRAM bytes of one launch "
                                 "from ncu (profiles/flagship_traffic.json), algorithmic bytes are 72 B per blob",
                         "fp64_instr_per_pair": FP64_INSTR_PER_PAIR,
                         "pipe_frac": value / world * FP64_INSTR_PER_PAIR * 2 / 1e12 / nominal,
                         "nominal_peak": nominal},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "pairs/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(world * 2 * 24 * n), "d2h_bytes_per_step": int(24 * n)},
            "gpu_launches": int(args.steps * plan_launches * world),
            "clocks": clocks,
            "parity_rel_err_vs_oracle": parity,
            "gmres_mobility_solve": solve,
        }
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
    ap.add_argument("--ref-budget", type=float, default=12.0, help="seconds of CPU work per reference sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-solve", action="store_true", help="skip the config-2 GMRES mobility solve")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
