"""Build-container tool: the REFERENCE's own numba single-wall UF (mobility_numba.py:124) timed on this machine's
host cores at N = 1,000 / 10,000 / 100,000, next to the oracle port that bench.py's CPU legs
use on the GPU box (where /root/reference does not exist).   python tools/time_numba_reference.py > profiles/<name>.json"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import numba
import oracle
from oracle import make_ref
mobility_numba = make_ref.import_mobility().mobility_numba        # the staged, byte-identical copy (oracle/_ref)
from bench import make_cloud

res = {"numba_threads": numba.get_num_threads(), "oracle_threads": oracle.threads(), "cpu_count": os.cpu_count(), "rates": {}}
L = np.zeros(3)
for n in (1000, 10000, 100000):
    r, f, a, eta = make_cloud(n)
    reps = 5 if n <= 10000 else 1
    mobility_numba.single_wall_mobility_trans_times_force_numba(r[:64].copy(), f[:64].copy(), eta, a, L)   # JIT
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        u = mobility_numba.single_wall_mobility_trans_times_force_numba(r, f, eta, a, L)
        best = min(best, time.perf_counter() - t0)
    t0 = time.perf_counter()
    uo = oracle.uf_wall_fast(r, f, eta, a)
    to = time.perf_counter() - t0
    res["rates"][str(n)] = {"numba_reference_Gpairs_s": n * n / best / 1e9, "oracle_port_fast_Gpairs_s": n * n / to / 1e9,
                            "rel_diff": float(np.linalg.norm(u - uo) / np.linalg.norm(u))}
print(json.dumps(res, indent=1))
