"""torchrun check (N ranks, NCCL): the sharded pair-symmetric products (all four blocks, open and periodic xy) equal
the single-device row kernel."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from rigidmultiblobswall_b200 import mobility as mb, parallel
world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rng = np.random.default_rng(5)
n, a, eta = 40001, 0.13, 7.0
side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
r = torch.from_numpy(side * rng.random((n, 3))).to(dev)
f = torch.from_numpy(rng.standard_normal((n, 3))).to(dev).reshape(-1)
res = {"world": world, "n": n}
for bname, block in (("UF", mb.UF), ("UT", mb.UT), ("WF", mb.WF), ("WT", mb.WT)):
    for lname, L in (("open", None), ("xy", np.array([side, side, 0.0]))):
        sm = parallel.ShardedMobility(r, eta, a, block=block, periodic_length=L)
        o, c = sm.offset, sm.count
        u = sm(f[3 * o:3 * (o + c)].clone())
        ref = mb.mobility_product(block, mb.SINGLE_WALL, r, f, eta, a, periodic_length=L, flags=mb.DETERMINISTIC)[3 * o:3 * (o + c)]
        err = torch.linalg.norm(u - ref) / torch.linalg.norm(ref)
        errs = [torch.zeros_like(err) for _ in range(world)]
        dist.all_gather(errs, err)
        res["%s_%s" % (bname, lname)] = {"symmetric": bool(sm.symmetric), "max_rel_err_over_ranks": max(float(e) for e in errs)}
if rank == 0:
    print(json.dumps(res))
dist.destroy_process_group()
