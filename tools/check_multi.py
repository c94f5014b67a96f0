"""torchrun check (N ranks, NCCL): the sharded pair-symmetric product equals the single-device row kernel."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from rigidmultiblobswall_b200 import mobility as mb, parallel
world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rng = np.random.default_rng(5)
n, a, eta = 70001, 0.13, 7.0
side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
r = torch.from_numpy(side * rng.random((n, 3))).to(dev)
f = torch.from_numpy(rng.standard_normal((n, 3))).to(dev).reshape(-1)
sm = parallel.ShardedMobility(r, eta, a)
o, c = sm.offset, sm.count
u = sm(f[3 * o:3 * (o + c)].clone())
ref = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.DETERMINISTIC)[3 * o:3 * (o + c)]
err = torch.linalg.norm(u - ref) / torch.linalg.norm(ref)
errs = [torch.zeros_like(err) for _ in range(world)]
dist.all_gather(errs, err)
if rank == 0:
    print(json.dumps({"world": world, "n": n, "symmetric": bool(sm.symmetric), "rel_err_per_rank": [float(e) for e in errs]}))
dist.destroy_process_group()
