"""The frozen algorithmic flop counts (profiles/flops_per_pair.json) are what tools/count_flops.py
derives from the reference source; re-derived when /root/reference is present (build container), else only read."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FROZEN = os.path.join(ROOT, "profiles", "flops_per_pair.json")


def test_frozen_counts_are_consistent():
    d = json.load(open(FROZEN))
    k = d["kernels"]
    uf = k["single_wall_mobility_trans_times_force_numba"]
    assert uf["pair_far"] == {"flops": 211, "sqrt": 2, "div": 9}
    assert d["headline"]["single_wall_UF_counted"] == 211 and d["headline"]["single_wall_UF_used_by_bench"] == 200
    # UT and WF are transposes of each other: same count; every kernel wraps with the same 8 flops per direction
    assert k["single_wall_mobility_trans_times_torque_numba"]["pair_far"] == \
        k["single_wall_mobility_rot_times_force_numba"]["pair_far"]
    assert {e["wrap_per_periodic_direction"] for e in k.values()} == {8}
    import bench
    assert bench.counted_flop_per_pair() == 211 and bench.FLOP_PER_PAIR <= 211


@pytest.mark.skipif(not os.path.exists("/root/reference/mobility/mobility_numba.py"), reason="reference checkout absent")
def test_counts_regenerate_identically(tmp_path):
    before = open(FROZEN).read()
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "count_flops.py")], stdout=subprocess.DEVNULL)
    assert open(FROZEN).read() == before
