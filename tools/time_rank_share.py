"""Dev tool: on ONE GPU, time the pair-symmetric partial product of every rank of an N-rank job (rmb_symmetric_partial
with rank = k, world = N); the slowest rank bounds the multi-GPU step.   python tools/time_rank_share.py 8 1000000"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rigidmultiblobswall_b200 import mobility as mb
from bench import make_cloud
world, n = int(sys.argv[1]), int(sys.argv[2])
r, f, a, eta = make_cloud(n)
rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
acc = None
ms = []
for k in range(world):
    acc = mb.symmetric_partial(rd, fd, eta, a, rank=k, world=world, acc=acc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(2):
        acc = mb.symmetric_partial(rd, fd, eta, a, rank=k, world=world, acc=acc)
    e1.record(); torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1) / 2)
print(json.dumps({"lib": os.path.basename(os.environ.get("RMB_LIB", "default")), "world": world, "n": n,
                  "ms_per_rank": [round(x, 2) for x in ms], "max_ms": round(max(ms), 2),
                  "ideal_ms": round(sum(ms) / world, 2), "Gpairs_s_at_max": round(n * n / max(ms) / 1e6, 1)}))
