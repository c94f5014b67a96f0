"""Golden dense blob mobilities from the REFERENCE's own mobility.py: single_wall_fluid_mobility (:1018-1116, with
the height clamp and the damping B) and rotne_prager_tensor (:967-1013), the functions multi_bodies.set_mobility_blobs
hands out, on a 40-blob cloud drawn like BASELINE config 1 (r = 5 a rand: blobs below z = a, overlapping pairs).

    python tests/golden/make_golden_dense.py     (build container only)
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden_rigid import import_multi_bodies  # noqa: E402


def main():
    mbod = import_multi_bodies()
    rng = np.random.default_rng(2024)
    n, eta, a = 40, 7.0, 0.13
    r = 5 * a * rng.random((n, 3))
    print("blobs below z = a:", int((r[:, 2] < a).sum()))
    wall = mbod.set_mobility_blobs("python")(r, eta, a)
    free = mbod.set_mobility_blobs("python_no_wall")(r, eta, a)
    print("symmetry", np.abs(wall - wall.T).max(), np.abs(free - free.T).max())
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dense_golden.npz")
    np.savez_compressed(path, r=r, eta_a=np.array([eta, a]), single_wall=wall, no_wall=free)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
