"""BASELINE config 3: force+torque (trans/rot) Mf for 100k minimally-resolved microrollers above a wall,
pseudo-periodic in xy (parameters of multi_bodies/examples/Mobility_Prod_Timing/Multiblob.inputfile:17-30).
Times the four blocks and the fused UF+UT product (CUDA events, device-resident inputs), and checks sampled rows
against the CPU oracle.  One JSON line."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rigidmultiblobswall_b200 import mobility as mb
import oracle

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
rng = np.random.default_rng(3)
a, eta = 1.0155, 0.957e-3
Lxy = np.sqrt(np.pi * a * a * n / 0.4)
L = np.array([Lxy, Lxy, 0.0])
r = np.stack([Lxy * rng.random(n), Lxy * rng.random(n), a + 3.5 * a * rng.random(n)], axis=1)
f, t = rng.standard_normal((n, 3)), rng.standard_normal((n, 3))
rd, fd, td = (torch.from_numpy(x).cuda().reshape(-1) for x in (r, f, t))
out = {"config": "C3: %d rollers, a=%g, L=[%.1f,%.1f,0], 9 image boxes, fp64" % (n, a, Lxy, Lxy), "blocks": {}}
pairs = float(n) * n * 9

def timed(fn, reps=3):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        res = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, res

b, e = 4242, 4242 + 16
for name, blk, vec in (("UF", mb.UF, fd), ("UT", mb.UT, td), ("WF", mb.WF, fd), ("WT", mb.WT, td)):
    ms, u = timed(lambda: mb.mobility_product(blk, mb.SINGLE_WALL, rd, vec, eta, a, periodic_length=L))
    ref = oracle.mobility_product(name, "single_wall", r, f if vec is fd else t, eta, a, L, targets=(b, e))
    err = float(np.linalg.norm(u[3 * b:3 * e].cpu().numpy() - ref) / np.linalg.norm(ref))
    out["blocks"][name] = {"ms": ms, "Gpairs_per_s": pairs / ms / 1e6, "rel_err_sampled_rows": err}
ms, u = timed(lambda: mb.mobility_product(mb.UF_UT, mb.SINGLE_WALL, rd, fd, eta, a, vec_b=td, periodic_length=L))
ref = (oracle.mobility_product("UF", "single_wall", r, f, eta, a, L, targets=(b, e))
       + oracle.mobility_product("UT", "single_wall", r, t, eta, a, L, targets=(b, e)))
err = float(np.linalg.norm(u[3 * b:3 * e].cpu().numpy() - ref) / np.linalg.norm(ref))
out["blocks"]["UF+UT fused"] = {"ms": ms, "Gpairs_per_s": pairs / ms / 1e6, "rel_err_sampled_rows": err}
out["four_blocks_ms"] = sum(out["blocks"][k]["ms"] for k in ("UF", "UT", "WF", "WT"))
print(json.dumps(out))
