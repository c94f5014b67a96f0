"""Dev tool: breakdown of the config-2 mobility solve (1,000 boomerangs)."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import make_c2
from rigidmultiblobswall_b200 import rigid, solvers, mobility as mb
ref_cfg, loc, quat, FT, a, eta, tol = make_c2(1000)
bodies = rigid.RigidBodies(ref_cfg, loc, quat)
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3, out
r = bodies.get_r_vectors()
res = {}
res["pc_build_ms"], PC = timed(lambda: rigid.BlockDiagonalPreconditioner(bodies, r, eta, a))
x = torch.randn(3 * bodies.n_blobs + 6 * bodies.n_bodies, dtype=torch.float64, device="cuda")
res["pc_apply_ms"], _ = timed(lambda: PC(x), 20)
res["operator_ms"], _ = timed(lambda: rigid.linear_operator_rigid(x, bodies, r, eta, a), 20)
res["mf_ms"], _ = timed(lambda: mb.single_wall_mobility_trans_times_force_b200(r, x[:3 * bodies.n_blobs], eta, a), 20)
res["solve_ms"], out = timed(lambda: rigid.solve_mobility_problem(bodies, FT, eta, a, tolerance=tol))
res["solve_with_pc_given_ms"], out = timed(lambda: rigid.solve_mobility_problem(bodies, FT, eta, a, tolerance=tol, PC=PC))
res["iterations"] = int(out[2])
print(json.dumps(res))
