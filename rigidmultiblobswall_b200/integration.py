"""Registration of the B200 path inside an UNMODIFIED reference checkout.

The reference selects its Mf implementation by string:
  (1) multi_bodies.set_mobility_vector_prod(implementation)      multi_bodies/multi_bodies.py:233-287
      (input-file key `mobility_vector_prod_implementation`, read_input/read_input.py:66)
  (2) QuaternionIntegratorRollers.__init__ picks the four blocks by `domain` and by substring match on
      'pycuda' / 'numba'                                          quaternion_integrator_rollers.py:59-91
  (3) multi_bodies.calc_slip hard-codes mb.no_wall/single_wall_mobility_trans_times_torque_pycuda
                                                                  multi_bodies.py:99-102
register() makes all three reach the B200 kernels without editing the reference:
  * every `<name>_b200` / `<name>_b200_fp32` function is attached to the reference's `mobility` module;
  * set_mobility_vector_prod is wrapped to accept 'b200', 'b200_no_wall', 'b200_free_surface' (+ '_fp32' of each)
    and the per-blob-radii family 'radii_b200', 'radii_b200_no_wall', 'radii_b200_free_surface' (which, like the
    reference's 'radii_numba*' strings, multi_bodies.py:262-287, bind the radii of kwargs['bodies']);
  * set_mobility_blobs (the DENSE blob mobility of one body, multi_bodies.py:205-227, input-file key
    `mobility_blobs_implementation`, read_input.py:65) is wrapped to accept 'b200' and 'b200_no_wall';
  * optionally (alias_pycuda=True) the `<name>_pycuda` attributes of the mobility module are pointed at the B200
    functions, so that code paths (2) and (3), which look the functions up under their PyCUDA names, run on the
    B200 path when the input file says `mobility_vector_prod_implementation  pycuda`.
"""
import functools

import numpy as np

from . import mobility as _mb

# string -> function name: the B200 twins of the reference's 'numba' / 'numba_no_wall' / 'numba_free_surface'
# (multi_bodies.py:246-261)
IMPLEMENTATIONS = {
    "b200": "single_wall_mobility_trans_times_force_b200",
    "b200_fp32": "single_wall_mobility_trans_times_force_b200_fp32",
    "b200_no_wall": "no_wall_mobility_trans_times_force_b200",
    "b200_no_wall_fp32": "no_wall_mobility_trans_times_force_b200_fp32",
    "b200_free_surface": "free_surface_mobility_trans_times_force_b200",
    "b200_free_surface_fp32": "free_surface_mobility_trans_times_force_b200_fp32",
}

# the twins of 'radii_numba' / 'radii_numba_no_wall' / 'radii_numba_free_surface' (multi_bodies.py:262-287): a
# source-target product bound to the bodies' own blob radii through mobility_radii_trans_times_force
RADII_IMPLEMENTATIONS = {
    "radii_b200": "single_wall_mobility_trans_times_force_source_target_b200",
    "radii_b200_no_wall": "no_wall_mobility_trans_times_force_source_target_b200",
    "radii_b200_free_surface": "free_surface_mobility_trans_times_force_source_target_b200",
}


DENSE_IMPLEMENTATIONS = {
    "b200": "single_wall_fluid_mobility_b200",
    "b200_no_wall": "rotne_prager_tensor_b200",
}


def attach(mobility_module, alias_pycuda=False):
    """Add the *_b200 functions to the reference's mobility module (mobility/mobility.py)."""
    for name in _mb.__all__:
        if name.endswith(("_b200", "_b200_fp32")):
            setattr(mobility_module, name, getattr(_mb, name))
            if alias_pycuda and name.endswith("_b200"):
                setattr(mobility_module, name[:-len("_b200")] + "_pycuda", getattr(_mb, name))
    return mobility_module


def register(multi_bodies_module, alias_pycuda=False):
    """Wrap multi_bodies.set_mobility_vector_prod so the new implementation strings resolve."""
    attach(multi_bodies_module.mb, alias_pycuda=alias_pycuda)
    original_blobs = getattr(multi_bodies_module, "set_mobility_blobs", None)
    if original_blobs is not None and not getattr(original_blobs, "_b200_wrapped", False):
        @functools.wraps(original_blobs)
        def set_mobility_blobs(implementation, *args, **kwargs):
            if implementation in DENSE_IMPLEMENTATIONS:
                return getattr(_mb, DENSE_IMPLEMENTATIONS[implementation])
            return original_blobs(implementation, *args, **kwargs)

        set_mobility_blobs._b200_wrapped = True
        multi_bodies_module.set_mobility_blobs = set_mobility_blobs
    original = multi_bodies_module.set_mobility_vector_prod
    if getattr(original, "_b200_wrapped", False):
        return multi_bodies_module

    @functools.wraps(original)
    def set_mobility_vector_prod(implementation, *args, **kwargs):
        if implementation in IMPLEMENTATIONS:
            return getattr(_mb, IMPLEMENTATIONS[implementation])
        if implementation in RADII_IMPLEMENTATIONS:
            # multi_bodies.py:279-286: the radii of every blob of kwargs['bodies'], in blob order
            radius_blobs = np.concatenate([np.asarray(b.blobs_radius, dtype=np.float64).reshape(-1)
                                           for b in kwargs.get("bodies")], axis=0)
            return functools.partial(_mb.mobility_radii_trans_times_force, radius_blobs=radius_blobs,
                                     function=getattr(_mb, RADII_IMPLEMENTATIONS[implementation]))
        return original(implementation, *args, **kwargs)

    set_mobility_vector_prod._b200_wrapped = True
    multi_bodies_module.set_mobility_vector_prod = set_mobility_vector_prod
    return multi_bodies_module


def roller_blocks(domain="single_wall", precision="fp64"):
    """The block table QuaternionIntegratorRollers builds at quaternion_integrator_rollers.py:59-91."""
    sfx = "_b200" if precision == "fp64" else "_b200_fp32"
    out = {"mobility_trans_times_force": getattr(_mb, "%s_mobility_trans_times_force%s" % (domain, sfx)),
           "mobility_trans_times_torque": getattr(_mb, "%s_mobility_trans_times_torque%s" % (domain, sfx))}
    if domain != "in_plane":
        out["mobility_rot_times_force"] = getattr(_mb, "%s_mobility_rot_times_force%s" % (domain, sfx))
        out["mobility_rot_times_torque"] = getattr(_mb, "%s_mobility_rot_times_torque%s" % (domain, sfx))
        out["mobilit_trans_times_force_torque"] = getattr(_mb, "%s_mobility_trans_times_force_torque%s" % (domain, sfx))
    return out
