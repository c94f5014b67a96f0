"""Rates of the source-target kernels (per-blob radii) on one GPU: G pairs/s per wall mode and precision."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from rigidmultiblobswall_b200 import mobility as mb

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
rng = np.random.default_rng(3)
a = 0.13
side = (n * 4 * np.pi / 3 / 0.1) ** (1.0 / 3.0) * a
dev = torch.device("cuda")
r = side * rng.random((n, 3)); r[:, 2] += a
rad = a * (0.6 + 0.8 * rng.random(n))
r_d, f_d, rad_d = (torch.from_numpy(x).to(dev) for x in (r, rng.standard_normal((n, 3)), rad))
res = {"n": n}
for wname, w in (("single_wall", mb.SINGLE_WALL), ("no_wall", mb.NO_WALL), ("free_surface", mb.FREE_SURFACE)):
    for pname, p in (("fp64", mb.FP64), ("fp32", mb.FP32)):
        out = torch.empty(3 * n, dtype=torch.float64, device=dev)
        for it in range(2):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            mb.mobility_source_target(w, r_d, r_d, f_d, rad_d, rad_d, 7.0, precision=p, out=out)
            e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        res["%s_%s_Gpairs_s" % (wname, pname)] = n * n / ms / 1e6
print(json.dumps(res))
