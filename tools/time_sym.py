"""Dev tool: default vs pair-symmetric flagship product."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rigidmultiblobswall_b200 import mobility as mb
from bench import make_cloud
res = {}
for n in [int(x) for x in sys.argv[1:]] or [400000]:
    r, f, a, eta = make_cloud(n)
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    for name, fl in (("default", 0), ("symmetric", mb.SYMMETRIC)):
        for _ in range(2):
            out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, rd, fd, eta, a, flags=fl)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, rd, fd, eta, a, flags=fl)
        e1.record(); torch.cuda.synchronize()
        res["%s_n%d_Gpairs_s" % (name, n)] = n * n / (e0.elapsed_time(e1) / 3) / 1e6
        res["%s_n%d_norm" % (name, n)] = float(out.norm())
print(json.dumps(res))
