"""oracle/_ref (git-ignored, shipped to the GPU box by gpurun) is a byte-identical copy of the reference files the
CPU baseline and the executed drop-in test import: sha256 of every staged file against its manifest, and against
/root/reference where that exists (build container)."""
import os

import pytest

from oracle import make_ref


def test_staged_reference_is_unmodified():
    if not make_ref.available() and not os.path.isdir(make_ref.REF_SRC):
        pytest.skip("neither oracle/_ref nor /root/reference present")
    assert make_ref.stage() is True
    assert make_ref.verify() == []
    with open(os.path.join(make_ref.REF_DST, "MANIFEST.json")) as f:
        import json
        files = json.load(f)["files"]
    for must in ("mobility/mobility.py", "mobility/mobility_numba.py", "multi_bodies/multi_bodies.py", "body/body.py",
                 "stochastic_forcing/stochastic_forcing.py", "general_application_utils.py"):
        assert must in files


def test_oracle_ref_is_git_ignored():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert "oracle/_ref/" in open(os.path.join(root, ".gitignore")).read().split()
    gi = os.path.join(root, ".gpurunignore")
    assert not os.path.exists(gi) or "oracle/_ref" not in open(gi).read()       # it must travel to the GPU box
