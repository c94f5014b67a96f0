"""Dev tool: row kernel (RMB_DETERMINISTIC) vs pair-symmetric kernel; all blocks, open and pseudo-periodic (xy)
boundaries, wall (and no wall for UF).   python tools/time_sym.py 100000 400000"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rigidmultiblobswall_b200 import mobility as mb
from bench import make_cloud
res = {}
only = os.environ.get("TIME_SYM_ONLY")          # e.g. "xy_UF_wall,open_UF_wall"
for n in [int(x) for x in sys.argv[1:]] or [400000]:
    r, f, a, eta = make_cloud(n)
    side = float(r[:, :2].max())
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    for bname, L, boxes in (("open", None, 1), ("xy", np.array([side, side, 0.0]), 9)):
        for wname, wall, block in (("UF_wall", mb.SINGLE_WALL, mb.UF), ("UF_nowall", mb.NO_WALL, mb.UF),
                                   ("UT_wall", mb.SINGLE_WALL, mb.UT), ("WF_wall", mb.SINGLE_WALL, mb.WF),
                                   ("WT_wall", mb.SINGLE_WALL, mb.WT)):
            if only and "%s_%s" % (bname, wname) not in only.split(","):
                continue
            outs = {}
            for name, fl in (("row", mb.DETERMINISTIC), ("symmetric", mb.SYMMETRIC)):
                reps = 3 if boxes * n <= 400000 else 1
                for _ in range(2 if boxes * n <= 400000 else 1):
                    out = mb.mobility_product(block, wall, rd, fd, eta, a, periodic_length=L, flags=fl)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    out = mb.mobility_product(block, wall, rd, fd, eta, a, periodic_length=L, flags=fl)
                e1.record(); torch.cuda.synchronize()
                res["%s_%s_%s_n%d_Gpairs_s" % (name, bname, wname, n)] = boxes * n * n / (e0.elapsed_time(e1) / reps) / 1e6
                outs[name] = out
            res["diff_%s_%s_n%d" % (bname, wname, n)] = float((outs["row"] - outs["symmetric"]).norm() / outs["row"].norm())
print(json.dumps(res))
