"""BASELINE config 5 from the command line: one Brownian step for --bodies shell_N_162 spheres sharded by body over
the ranks (bench.c5_record; the `c5` key of the 8-GPU bench line is the same function at 10,000 spheres).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_c5.py
    python tools/bench_c5.py --bodies 400            (one GPU, smaller)"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--bodies", type=int, default=10000)
ap.add_argument("--tol", type=float, default=1e-6)
ap.add_argument("--lanczos-tol", type=float, default=1e-4)
args = ap.parse_args()
world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
rec = bench.c5_record(world, rank, dev, n_bodies=args.bodies, tol=args.tol, lanczos_tol=args.lanczos_tol)
if rank == 0:
    print(json.dumps(rec))
if world > 1:
    dist.destroy_process_group()
