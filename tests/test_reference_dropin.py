"""Registration into the UNMODIFIED reference (build container only: /root/reference does not exist on the GPU box).
The reference's driver module is imported in place with the two harness stubs of SURVEY 8c (`imp`, removed from
Python 3.12, and quaternion_integrator/gmres.py, which imports a SciPy private module that no longer exists); no GPU
call is made -- this pins the selection mechanism (multi_bodies.py:233-287, quaternion_integrator_rollers.py:59-91),
the compute parity is in the GPU suite."""
import inspect
import os
import sys

import pytest

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "multi_bodies", "multi_bodies.py")),
                                reason="reference checkout absent")
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def multi_bodies():
    sys.path.insert(0, os.path.join(HERE, "golden"))
    from make_golden_rigid import import_multi_bodies
    return import_multi_bodies()


def test_implementation_strings_resolve_in_the_reference_driver(multi_bodies):
    from rigidmultiblobswall_b200 import integration, mobility as mb
    numba_fn = multi_bodies.set_mobility_vector_prod("numba")
    integration.register(multi_bodies)
    assert multi_bodies.set_mobility_vector_prod("numba") is numba_fn                    # reference strings untouched
    assert multi_bodies.set_mobility_vector_prod("numba_no_wall") is multi_bodies.mb.no_wall_mobility_trans_times_force_numba
    assert multi_bodies.set_mobility_vector_prod("b200") is mb.single_wall_mobility_trans_times_force_b200
    assert multi_bodies.set_mobility_vector_prod("b200_no_wall") is mb.no_wall_mobility_trans_times_force_b200
    assert multi_bodies.set_mobility_vector_prod("b200_fp32") is mb.single_wall_mobility_trans_times_force_b200_fp32
    assert multi_bodies.set_mobility_vector_prod("b200_free_surface") is mb.free_surface_mobility_trans_times_force_b200
    # the per-blob-radii family binds the radii of kwargs['bodies'] like the reference's 'radii_numba*' (:262-287)
    import numpy as np

    class B:
        blobs_radius = np.array([0.1, 0.2])
    for name, fn in (("radii_b200", mb.single_wall_mobility_trans_times_force_source_target_b200),
                     ("radii_b200_no_wall", mb.no_wall_mobility_trans_times_force_source_target_b200),
                     ("radii_b200_free_surface", mb.free_surface_mobility_trans_times_force_source_target_b200)):
        p = multi_bodies.set_mobility_vector_prod(name, bodies=[B(), B()])
        assert p.func is mb.mobility_radii_trans_times_force and p.keywords["function"] is fn
        assert p.keywords["radius_blobs"].tolist() == [0.1, 0.2, 0.1, 0.2]
    # dense blob mobility of one body (mobility_blobs_implementation, multi_bodies.py:205-227)
    assert multi_bodies.set_mobility_blobs("python") is multi_bodies.mb.single_wall_fluid_mobility
    assert multi_bodies.set_mobility_blobs("b200") is mb.single_wall_fluid_mobility_b200
    assert multi_bodies.set_mobility_blobs("b200_no_wall") is mb.rotne_prager_tensor_b200
    # the functions are reachable under the reference's own module, next to their numba twins
    for name in dir(multi_bodies.mb):
        if name.endswith("_numba") and name.startswith(("single_wall_mobility", "no_wall_mobility", "in_plane_mobility")) \
                and "source_target" not in name:
            assert hasattr(multi_bodies.mb, name[:-len("_numba")] + "_b200"), name


def test_signatures_match_the_numba_wrappers(multi_bodies):
    """every *_b200 function takes the positional arguments of its reference twin plus *args, **kwargs"""
    from rigidmultiblobswall_b200 import integration
    integration.register(multi_bodies)
    ref = multi_bodies.mb
    checked = 0
    for name in dir(ref):
        if not name.endswith("_b200"):
            continue
        twin = next((getattr(ref, name[:-len("_b200")] + sfx) for sfx in ("_numba", "_pycuda")
                     if hasattr(ref, name[:-len("_b200")] + sfx)), None)
        if twin is None:
            continue
        want = [p.name for p in inspect.signature(twin).parameters.values()
                if p.kind == inspect.Parameter.POSITIONAL_OR_KEYWORD]
        got = [p.name for p in inspect.signature(getattr(ref, name)).parameters.values()
               if p.kind == inspect.Parameter.POSITIONAL_OR_KEYWORD]
        assert len(got) == len(want), (name, got, want)
        kinds = {p.kind for p in inspect.signature(getattr(ref, name)).parameters.values()}
        assert inspect.Parameter.VAR_POSITIONAL in kinds and inspect.Parameter.VAR_KEYWORD in kinds, name
        checked += 1
    assert checked >= 14


def test_roller_integrator_lookup_by_pycuda_names(multi_bodies):
    """alias_pycuda=True: the substring lookups of quaternion_integrator_rollers.py:59-91 and multi_bodies.calc_slip
    (:99-102) find the B200 functions under the *_pycuda names"""
    from rigidmultiblobswall_b200 import integration, mobility as mb
    integration.register(multi_bodies, alias_pycuda=True)
    ref = multi_bodies.mb
    assert ref.single_wall_mobility_trans_times_torque_pycuda is mb.single_wall_mobility_trans_times_torque_b200
    assert ref.no_wall_mobility_rot_times_torque_pycuda is mb.no_wall_mobility_rot_times_torque_b200
    assert ref.single_wall_mobility_trans_times_force_torque_pycuda is mb.single_wall_mobility_trans_times_force_torque_b200
