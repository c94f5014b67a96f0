"""Dev tool: time ONE product configuration (CUDA events, device-resident inputs), print one JSON line.  Also the
command ncu is pointed at (`ncu ... python tools/run_case.py <case> --n N --reps 1 --warm 1`).

  python tools/run_case.py CASE[,CASE...] [--n N] [--reps R] [--warm W] [--cloud dilute|dense|c1box|c3]

CASE = <block>_<wall>_<boundary>_<precision>[_row|_sym]   block UF|UT|WF|WT|UFUT, wall wall|free|fs,
       boundary open|xy|x, precision f64|f32;  or st_<wall>_<boundary>_<precision> (source-target, radii +-40 %)
RMB_LIB=<variant .so> selects a tuning build (tools/build_tune.py)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from rigidmultiblobswall_b200 import mobility as mb

ap = argparse.ArgumentParser()
ap.add_argument("cases")
ap.add_argument("--n", type=int, default=400000)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--warm", type=int, default=2)
ap.add_argument("--cloud", default="dilute")
args = ap.parse_args()
n = args.n
Lc3 = None
if args.cloud == "dilute":
    r, f, a, eta = bench.make_cloud(n)
elif args.cloud == "dense":
    r, f, a, eta = bench.make_dense_cloud(n)
elif args.cloud == "c1box":                       # every blob in the 5a box of mobility/test_blobs.py:30-34
    rng = np.random.default_rng(7); a, eta = 0.13, 7.0
    r, f = 5 * a * rng.random((n, 3)), rng.standard_normal((n, 3))
else:
    r, f, _, a, eta, Lc3 = bench.make_c3(n)
t = np.random.default_rng(1).standard_normal((n, 3))
rd, fd, td = (torch.from_numpy(x).cuda().reshape(-1) for x in (r, f, t))
rad = torch.from_numpy(a * (0.6 + 0.8 * np.random.default_rng(2).random(n))).cuda()
side = float(r[:, :2].max()) * 1.0001
out = {"lib": os.path.basename(os.environ.get("RMB_LIB", "default")), "n": n, "cloud": args.cloud}
for case in args.cases.split(","):
    p = case.split("_")
    st = p[0] == "st"
    blk, wall, bnd, prec = p[0], p[1], p[2], p[3]
    mode = p[4] if len(p) > 4 else ""
    W = {"wall": mb.SINGLE_WALL, "free": mb.NO_WALL, "fs": mb.FREE_SURFACE}[wall]
    L = {"open": None, "xy": np.array([side, side, 0.0]), "x": np.array([side, 0.0, 0.0])}[bnd]
    if Lc3 is not None and bnd == "xy":
        L = Lc3
    boxes = {"open": 1, "xy": 9, "x": 3}[bnd]
    P = mb.FP64 if prec == "f64" else mb.FP32
    fl = {"": 0, "row": mb.DETERMINISTIC, "sym": mb.SYMMETRIC}[mode]
    if st:
        fn = lambda: mb.mobility_source_target(W, rd, rd, fd, rad, rad, eta, periodic_length=L, precision=P)
    else:
        B = {"UF": mb.UF, "UT": mb.UT, "WF": mb.WF, "WT": mb.WT, "UFUT": mb.UF_UT}[blk]
        vec = fd if blk in ("UF", "WF", "UFUT") else td
        fn = lambda: mb.mobility_product(B, W, rd, vec, eta, a, periodic_length=L, vec_b=td if blk == "UFUT" else None,
                                         precision=P, flags=fl)
    for _ in range(args.warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    out[case] = {"ms": round(ms, 3), "Gpairs_s": round(boxes * float(n) * n / ms / 1e6, 2)}
print(json.dumps(out))
