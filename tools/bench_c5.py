"""BASELINE config 5: one Brownian step for 10,000 162-blob spheres (1.62 M blobs) sharded by body over the ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_c5.py
    python tools/bench_c5.py --bodies 400            (one GPU, smaller)

Step (the stochastic slip + mobility solve of quaternion_integrator_multi_bodies.py, first part):
  1. block-diagonal preconditioner of the local bodies,
  2. Lanczos  W = sqrt(2 kT / dt) M^{1/2} z   (preconditioned: P^T M P, mapped back with P^-1),
  3. GMRES    [[M, -K], [-K^T, 0]] [lambda; U] = [-W; -F]   (slip = -W: the random blob velocities).
Input: shell_N_162 (blob radius 0.26201755), centres on a 100 x 100 grid with spacing 3.0 at z in [1.5, 2.5),
random unit quaternions, random force/torque, eta = 1, kT = 0.0041, dt = 0.01.  One JSON line on stdout (rank 0);
all times are CUDA-event times, max over ranks.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bodies", type=int, default=10000)
    ap.add_argument("--tol", type=float, default=1e-6)
    ap.add_argument("--lanczos-tol", type=float, default=1e-4)
    args = ap.parse_args()
    from rigidmultiblobswall_b200 import parallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    g = np.load(os.path.join(ROOT, "tests", "golden", "rigid_golden.npz"))
    shell = g["struct/shell_N_162"]
    a = float(g["struct/shell_N_162_radius"][0])
    eta, kT, dt = 1.0, 0.0041, 0.01
    nb = args.bodies
    rng = np.random.default_rng(55)                       # same on every rank
    side = int(np.ceil(np.sqrt(nb)))
    k = np.arange(nb)
    loc = np.stack([3.0 * (k % side), 3.0 * (k // side), 1.5 + rng.random(nb)], axis=1)
    quat = rng.standard_normal((nb, 4))
    quat /= np.linalg.norm(quat, axis=1)[:, None]
    FT = rng.standard_normal((nb, 6))
    z_full = rng.standard_normal(3 * nb * 162)

    def timed(fn):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return out, float(ms.item()) * 1e-3

    system, t_setup = timed(lambda: parallel.ShardedRigidSystem(shell, loc, quat, eta, a,
                                                                group=dist.group.WORLD if world > 1 else False))
    o, c = system.ranges[system.rank]
    z_loc = torch.from_numpy(z_full[3 * o:3 * (o + c)].copy()).to(dev)
    # warm-up: one sharded product (NCCL communicator, kernel load)
    system.mobility(z_loc)
    _, t_mf = timed(lambda: system.mobility(z_loc))
    _, t_pc = timed(system.build_preconditioner)
    (noise, it_l), t_l = timed(lambda: system.stochastic_noise(z_loc, factor=np.sqrt(2 * kT / dt),
                                                                 tolerance=args.lanczos_tol))
    res_hist = []
    (sol, t_g) = timed(lambda: system.solve_mobility(system.local_slice(FT), tolerance=args.tol, slip_loc=-noise,
                                                     callback=res_hist.append))
    U_loc, lam_loc, its, info = sol
    # true residual of the sharded solve
    x = torch.cat([lam_loc.reshape(-1), U_loc.reshape(-1)])
    rhs = torch.cat([-noise.reshape(-1), -system.local_slice(FT)])
    rr = system.linear_operator(x) - rhs
    num = torch.dot(rr, rr).reshape(1)
    den = torch.dot(rhs, rhs).reshape(1)
    if world > 1:
        dist.all_reduce(num)
        dist.all_reduce(den)
    n_blobs = system.n_blobs_total
    if rank == 0:
        print(json.dumps({
            "config": "C5: Brownian step, %d spheres x 162 blobs = %d blobs, single wall, fp64" % (nb, n_blobs),
            "n_gpus": world, "sharding": "by body; lambda slices all-gathered per product; dots all-reduced",
            "seconds_total": t_pc + t_l + t_g, "seconds_preconditioner": t_pc, "seconds_lanczos": t_l,
            "seconds_gmres": t_g, "seconds_one_Mf": t_mf, "seconds_setup": t_setup,
            "pairs_per_s_one_Mf": float(n_blobs) ** 2 / t_mf,
            "lanczos_iterations": int(it_l), "lanczos_tolerance": args.lanczos_tol,
            "gmres_iterations": int(its), "gmres_info": int(info), "gmres_tolerance": args.tol,
            "true_residual": float(torch.sqrt(num / den).item()),
            "mf_products": int(it_l) + 1 + int(its) + 2,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
