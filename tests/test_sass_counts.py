"""The instruction counts bench.py's roofline uses (profiles/sass_counts.json) are re-counted from the SASS of the
built library: a kernel change that is not followed by `python tools/loop_count.py --write` (or a build through
__graft_entry__.build(), which writes the file) fails here instead of leaving a stale literal in the bench line."""
import json
import os
import shutil
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.mark.skipif(shutil.which("cuobjdump") is None and not os.path.exists("/usr/local/cuda/bin/cuobjdump"),
                    reason="cuobjdump not installed")
def test_committed_counts_match_the_built_library():
    import loop_count
    from rigidmultiblobswall_b200 import _build
    _build.build()
    now = loop_count.flagship_counts(_build.SO)
    with open(os.path.join(ROOT, "profiles", "sass_counts.json")) as f:
        committed = json.load(f)["flagship"]
    for key in ("total", "fp64", "fp64_near_path", "fp64_far_pair", "fp64_per_interaction"):
        assert now[key] == committed[key], (key, now[key], committed[key])
    # sanity of the count itself: an FP64-bound pair loop, two interactions per trip
    assert 80 <= now["fp64_far_pair"] <= 100 and now["opcodes"].get("SHFL", 0) == 6 and now["opcodes"].get("LDS", 0) == 3
    assert now["opcodes"].get("MUFU", 0) == 2          # one reciprocal square root per distance (|r|, |R|)


def test_bench_reads_the_counts_file():
    import bench
    sc = bench.sass_counts()
    assert sc["fp64_per_interaction"] * 2 == sc["fp64_far_pair"]
