"""Dev tool: norm-wise error of the fp32 variant against the fp64 oracle on random rows, as a function of N (the
coordinates / a grow like N^(1/3), and with them the rounding of the fp32 positions).  python tools/fp32_error.py 100000 1000000"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, oracle
from rigidmultiblobswall_b200 import mobility as mb
out = {}
oracle.set_threads()
for n in [int(x) for x in sys.argv[1:]] or [100000, 1000000]:
    r, f, a, eta = bench.make_cloud(n)
    idx, perm = bench.sampled_rows(n, 512, seed=3)
    ref = oracle.uf_wall_fast(r[perm], f[perm], eta, a, targets=(0, len(idx)))
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    u = mb.mobility_product(mb.UF, mb.SINGLE_WALL, rd, fd, eta, a, precision=mb.FP32)
    got = u.reshape(-1, 3)[torch.from_numpy(idx).cuda()].reshape(-1).cpu().numpy()
    out[str(n)] = {"rel_err_norm": float(np.linalg.norm(got - ref) / np.linalg.norm(ref)),
                   "max_row_rel_err": float(np.max(np.linalg.norm((got - ref).reshape(-1, 3), axis=1) /
                                                   np.linalg.norm(ref.reshape(-1, 3), axis=1))),
                   "max_coordinate_over_a": float(np.abs(r).max() / a)}
print(json.dumps(out))
