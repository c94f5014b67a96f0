"""CPU-side checks of the C-ABI library: it loads and exports every symbol
include/rmb_mobility.h declares.  No compute calls (no GPU here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        txt = open(os.path.join(ROOT, "include", fn)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        names |= set(re.findall(r"\b(rmb_[a-z0-9_]+)\s*\(", txt))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    from rigidmultiblobswall_b200 import _build
    so = _build.build()
    lib = ctypes.CDLL(so)
    decl = _declared()
    assert "rmb_mobility_product" in decl and len(decl) >= 5
    for name in decl:
        assert hasattr(lib, name), name


def test_version_and_argument_errors_without_gpu():
    from rigidmultiblobswall_b200 import _lib
    lib = _lib.load()
    assert lib.rmb_version() == 105
    # argument validation happens before any CUDA call
    rc = lib.rmb_mobility_product(99, 0, 0, 0, 1, 1, 0, None, None, None, 1.0, 1.0, None, None, None, 0, None)
    assert rc == -1 and b"block" in lib.rmb_last_error()
    rc = lib.rmb_mobility_product(0, 0, 0, 0, 5, 4, 0, None, None, None, 1.0, 1.0, None, None, None, 0, None)
    assert rc == -1 and b"target range" in lib.rmb_last_error()
    rc = lib.rmb_mobility_source_target(7, 0, 0, 1, 1, None, None, None, None, None, 1.0, None, None, None, 0, None)
    assert rc == -1 and b"wall" in lib.rmb_last_error()
    rc = lib.rmb_mobility_source_target(1, 0, 0, 3, 2, None, None, None, None, None, 1.0, None, 8, None, 0, None)
    assert rc == -1 and b"null" in lib.rmb_last_error()
    # multi-rank pair-symmetric form: every block with a single wall or none; never a free surface or a periodic z
    Lz = (ctypes.c_double * 3)(0.0, 0.0, 5.0)
    p = 16      # a non-null "device pointer" that is never dereferenced: every call below fails validation first
    rc = lib.rmb_symmetric_partial(1, 2, 0, 100, 0, 1, p, p, None, 1.0, 1.0, None, p, None, 0, None)
    assert rc == -1 and b"RMB_NO_WALL" in lib.rmb_last_error()
    rc = lib.rmb_symmetric_partial(3, 0, 0, 100, 0, 1, p, p, None, 1.0, 1.0, Lz, p, None, 0, None)
    assert rc == -1 and b"periodic in z" in lib.rmb_last_error()
    rc = lib.rmb_symmetric_partial(3, 0, 0, 100, 0, 1, p, p, None, 1.0, 1.0, None, p, None, 0, None)
    assert rc == -1 and b"null" in lib.rmb_last_error()          # no_wall WT is a valid block: the workspace check fires
    rc = lib.rmb_symmetric_partial(4, 0, 0, 100, 0, 1, p, p, None, 1.0, 1.0, None, p, None, 0, None)
    assert rc == -1 and b"vec_b" in lib.rmb_last_error()
    rc = lib.rmb_symmetric_partial(0, 1, 0, 100, 2, 2, p, p, None, 1.0, 1.0, None, p, None, 0, None)
    assert rc == -1 and b"rank" in lib.rmb_last_error()
    assert lib.rmb_symmetric_acc_len(1000) == 1024 * 3


def test_python_api_mirrors_reference_names():
    from rigidmultiblobswall_b200 import mobility as mb
    for stem in ("no_wall_mobility_trans_times_force", "single_wall_mobility_trans_times_force",
                 "in_plane_mobility_trans_times_force", "single_wall_mobility_trans_times_torque",
                 "no_wall_mobility_rot_times_force", "single_wall_mobility_rot_times_torque",
                 "single_wall_mobility_trans_times_force_torque"):
        assert callable(getattr(mb, stem + "_b200"))
        assert callable(getattr(mb, stem + "_b200_fp32"))
    for wall in ("single_wall", "no_wall", "free_surface"):
        assert callable(getattr(mb, wall + "_mobility_trans_times_force_source_target_b200"))
    assert callable(mb.mobility_radii_trans_times_force)


def test_product_path_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "rigidmultiblobswall_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "mobility_oracle" not in txt, f


def test_registration_into_a_reference_like_module():
    """integration.register() against a stand-in with the reference's selection function shape
    (multi_bodies.py:233-287); no GPU call is made."""
    import types
    from rigidmultiblobswall_b200 import integration, mobility as mb
    ref_mb = types.SimpleNamespace(single_wall_mobility_trans_times_force_numba=lambda *a, **k: "numba")
    mod = types.SimpleNamespace(mb=ref_mb)

    def set_mobility_vector_prod(implementation, *args, **kwargs):
        if implementation == "numba":
            return ref_mb.single_wall_mobility_trans_times_force_numba
    mod.set_mobility_vector_prod = set_mobility_vector_prod
    mod.set_mobility_blobs = lambda implementation: "python-dense" if implementation == "python" else None
    integration.register(mod, alias_pycuda=True)
    integration.register(mod)          # idempotent
    assert mod.set_mobility_vector_prod("numba")() == "numba"
    assert mod.set_mobility_vector_prod("b200") is mb.single_wall_mobility_trans_times_force_b200
    assert mod.set_mobility_blobs("python") == "python-dense"
    assert mod.set_mobility_blobs("b200") is mb.single_wall_fluid_mobility_b200
    assert mod.set_mobility_blobs("b200_no_wall") is mb.rotne_prager_tensor_b200
    assert mod.set_mobility_vector_prod("b200_no_wall_fp32") is mb.no_wall_mobility_trans_times_force_b200_fp32
    assert ref_mb.single_wall_mobility_rot_times_torque_b200 is mb.single_wall_mobility_rot_times_torque_b200
    assert ref_mb.single_wall_mobility_trans_times_torque_pycuda is mb.single_wall_mobility_trans_times_torque_b200
    blocks = integration.roller_blocks("single_wall")
    assert set(blocks) == {"mobility_trans_times_force", "mobility_trans_times_torque", "mobility_rot_times_force",
                           "mobility_rot_times_torque", "mobilit_trans_times_force_torque"}
    assert set(integration.roller_blocks("in_plane")) == {"mobility_trans_times_force", "mobility_trans_times_torque"}
