"""Dev tool: row kernel (RMB_DETERMINISTIC) vs pair-symmetric kernel for the no_wall rotation blocks and the fused
no_wall product, open and pseudo-periodic (xy).   python tools/time_sym_free.py [n_open] [n_xy]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rigidmultiblobswall_b200 import mobility as mb
from bench import make_cloud
n_open = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
n_xy = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
res = {}
for bname, n, boxes in (("open", n_open, 1), ("xy", n_xy, 9)):
    r, f, a, eta = make_cloud(n)
    side = float(r[:, :2].max())
    L = None if boxes == 1 else np.array([side, side, 0.0])
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    td = torch.from_numpy(f[::-1].copy()).cuda().reshape(-1)
    for wname, block in (("UT", mb.UT), ("WT", mb.WT), ("UF_UT", mb.UF_UT)):
        outs = {}
        for name, fl in (("row", mb.DETERMINISTIC), ("symmetric", mb.SYMMETRIC)):
            vb = td if block == mb.UF_UT else None
            out = mb.mobility_product(block, mb.NO_WALL, rd, fd, eta, a, periodic_length=L, flags=fl, vec_b=vb)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(2):
                out = mb.mobility_product(block, mb.NO_WALL, rd, fd, eta, a, periodic_length=L, flags=fl, vec_b=vb)
            e1.record(); torch.cuda.synchronize()
            res["%s_%s_%s_nowall_n%d_Gpairs_s" % (name, bname, wname, n)] = boxes * n * n / (e0.elapsed_time(e1) / 2) / 1e6
            outs[name] = out
        res["diff_%s_%s_n%d" % (bname, wname, n)] = float((outs["row"] - outs["symmetric"]).norm() / outs["row"].norm())
print(json.dumps(res))
