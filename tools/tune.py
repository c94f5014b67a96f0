"""Dev tool: time the flagship product (single-wall UF fp64, open) for the library in $RMB_LIB."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rigidmultiblobswall_b200 import mobility as mb
from bench import make_cloud
out = {"lib": os.path.basename(os.environ.get("RMB_LIB", "default"))}
for n, blk, wall, prec, tag in [(int(a.split(":")[0]), *a.split(":")[1:]) for a in sys.argv[1:]]:
    r, f, a, eta = make_cloud(n)
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    L = np.array([r[:, 0].max() * 1.01, r[:, 1].max() * 1.01, 0.0]) if tag == "xy" else None
    args = dict(periodic_length=L, precision=int(prec))
    B = getattr(mb, blk)
    vb = fd if blk == "UF_UT" else None
    for _ in range(2):
        mb.mobility_product(B, int(wall), rd, fd, eta, a, vec_b=vb, **args)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        mb.mobility_product(B, int(wall), rd, fd, eta, a, vec_b=vb, **args)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    boxes = 9 if tag == "xy" else 1
    out["%s_w%s_p%s_%s_n%d" % (blk, wall, prec, tag, n)] = round(n * n * boxes / ms / 1e6, 2)   # Gpairs/s
print(json.dumps(out))
