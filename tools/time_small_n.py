"""Dev tool: latency of one single-wall UF product at small N (device-resident and through the NumPy boundary) for
the library in $RMB_LIB.   python tools/time_small_n.py 1000 5000 15000 50000"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from rigidmultiblobswall_b200 import mobility as mb
out = {"lib": os.path.basename(os.environ.get("RMB_LIB", "default"))}
for n in [int(x) for x in sys.argv[1:]] or [1000, 15000]:
    r, f, a, eta = bench.make_cloud(n)
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    fn = lambda: mb.single_wall_mobility_trans_times_force_b200(rd, fd, eta, a)
    fn()
    ms, _ = bench.timed_events(fn, 100, torch)
    hn = lambda: mb.single_wall_mobility_trans_times_force_b200(r, f, eta, a, periodic_length=np.zeros(3))
    for _ in range(5):
        hn()
    ts = []
    for _ in range(100):
        t0 = time.perf_counter(); hn(); ts.append(time.perf_counter() - t0)
    out[str(n)] = {"device_resident_us": round(1e3 * ms, 1), "numpy_boundary_us": round(1e6 * float(np.median(ts)), 1),
                   "Gpairs_s_device": round(n * n / ms / 1e6, 1)}
print(json.dumps(out))
