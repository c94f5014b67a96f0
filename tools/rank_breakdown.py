"""Per-rank CUDA-event breakdown of one multi-GPU product step (what parallel.ShardedMobility.__call__ does):
all-gather of the force slices | this rank's share of the unordered pairs (prepare_kernel + memset + sym_kernel) |
all-reduce of the raw sums | sym_finish_kernel.  One JSON line on rank 0: ms per phase per rank (mean of --steps).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/rank_breakdown.py"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from rigidmultiblobswall_b200 import mobility as mb, parallel

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1000000)
ap.add_argument("--steps", type=int, default=5)
args = ap.parse_args()
world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = args.n
r, f, a, eta = bench.make_cloud(n)
cuts = [(n * k) // world for k in range(world + 1)]
off, cnt = cuts[rank], cuts[rank + 1] - cuts[rank]
rd = torch.from_numpy(r).to(dev).reshape(-1)
f_loc = torch.from_numpy(f).to(dev).reshape(-1)[3 * off:3 * (off + cnt)].clone()
sv = parallel.ShardedVector([(cuts[k], cuts[k + 1] - cuts[k]) for k in range(world)], 3, dev,
                            group=dist.group.WORLD if world > 1 else None)
acc = None
phases = ("all_gather", "pairs", "all_reduce", "finish", "step")
tot = {p: 0.0 for p in phases}
for it in range(args.steps + 2):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record()
    full = sv.gather(f_loc)
    ev[1].record()
    acc = mb.symmetric_partial(rd, full, eta, a, rank=rank, world=world, acc=acc)
    ev[2].record()
    if world > 1:
        dist.all_reduce(acc)
    ev[3].record()
    out = mb.symmetric_finish(rd, full, acc, eta, a, targets=(off, cnt))
    ev[4].record()
    torch.cuda.synchronize()
    if it >= 2:
        for k, p in enumerate(phases[:4]):
            tot[p] += ev[k].elapsed_time(ev[k + 1])
        tot["step"] += ev[0].elapsed_time(ev[4])
mine = torch.tensor([tot[p] / args.steps for p in phases], dtype=torch.float64, device=dev)
allr = [torch.empty_like(mine) for _ in range(world)]
if world > 1:
    dist.all_gather(allr, mine)
else:
    allr = [mine]
if rank == 0:
    tab = torch.stack(allr).cpu().numpy()
    print(json.dumps({"n": n, "world": world, "steps": args.steps, "unit": "ms per step, per rank",
                      **{p: [round(float(x), 3) for x in tab[:, k]] for k, p in enumerate(phases)},
                      "note": "all_reduce includes the wait for the slowest rank's pair kernel (the collective "
                              "synchronises the ranks); pairs = prepare_kernel + accumulator memset + sym_kernel"}))
if world > 1:
    dist.destroy_process_group()
