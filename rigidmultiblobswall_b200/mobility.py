"""Host-side mirror of the reference's blob-mobility API for the B200 path.

Every public function keeps the signature of its namesake in
/root/reference/mobility/mobility.py,

    f(r_vectors, vec, eta, a, *args, **kwargs) -> flat array of length 3N

(`periodic_length` is the only kwarg read, mobility.py:1150), so that
multi_bodies.set_mobility_vector_prod (multi_bodies/multi_bodies.py:233-287),
QuaternionIntegratorRollers (quaternion_integrator_rollers.py:59-91) and the
GMRES / Lanczos callers can select them as drop-in implementations:

    reference name (numba / pycuda)                        here
    single_wall_mobility_trans_times_force_{numba,pycuda}  single_wall_mobility_trans_times_force_b200
    no_wall_mobility_rot_times_torque_{numba,pycuda}       no_wall_mobility_rot_times_torque_b200
    single_wall_mobility_trans_times_force_torque_pycuda   single_wall_mobility_trans_times_force_torque_b200
    ...

Inputs may be NumPy arrays (uploaded, result downloaded: what SciPy callers
need) or torch CUDA tensors (everything stays on the device: what the
device-resident GMRES / Lanczos use).  Inputs are never modified.  Like the
reference, nothing is raised for bad geometry (NaN/Inf come back); a missing
CUDA extension or device DOES raise -- there is no CPU fallback.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib
from ._lib import (DETERMINISTIC, FP32, FP64, FREE_SURFACE, IN_PLANE, NO_TMA, NO_WALL, SINGLE_WALL,  # noqa: F401
                   SYMMETRIC, UF,
                   UF_UT, UT, WF, WT)

_workspaces = {}
default_flags = int(os.environ.get("RMB_FLAGS", "0"))  # OR-ed into every call (debug: RMB_FLAGS=2 -> RMB_NO_TMA)


def _device(dev=None):
    if not torch.cuda.is_available():
        raise _lib.RmbError("rigidmultiblobswall_b200 needs a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device()) if dev is None else dev


def _workspace(nbytes, dev):
    """Scratch buffer of the current stream on `dev` (one per (device, stream): products enqueued on different
    streams may run concurrently and must not share records; on one stream they are ordered)."""
    key = (dev, torch.cuda.current_stream(dev).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    return ws


class _HostStage:
    """Pinned staging of one NumPy-boundary call (what SciPy's GMRES makes at every iteration through the drop-in
    functions): the inputs are packed into ONE pinned buffer and go up in one asynchronous copy, the result comes
    down into a pinned buffer, and the stream is synchronised once.  One object per (device, stream, 3N); the
    reference's PyCUDA path allocates and copies three arrays per call (mobility_pycuda.py:2249-2266)."""

    def __init__(self, n3, dev):
        self.h_in = torch.empty(3, n3, dtype=torch.float64).pin_memory()
        self.h_in_np = self.h_in.numpy()
        self.d_in = torch.empty(3, n3, dtype=torch.float64, device=dev)
        self.h_out = torch.empty(n3, dtype=torch.float64).pin_memory()
        self.h_out_np = self.h_out.numpy()
        self.d_out = torch.empty(n3, dtype=torch.float64, device=dev)

    def upload(self, arrays):
        """arrays: up to three array-likes of 3N entries (any shape / stride / dtype); returns their device views"""
        k = len(arrays)
        for i, x in enumerate(arrays):
            x = np.asarray(x)
            if x.size != self.h_in_np.shape[1]:
                raise ValueError("vector has %d entries, expected %d" % (x.size, self.h_in_np.shape[1]))
            np.copyto(self.h_in_np[i], x.reshape(-1), casting="unsafe")
        self.d_in[:k].copy_(self.h_in[:k], non_blocking=True)
        return [self.d_in[i] for i in range(k)]

    def download(self, count, dev):
        self.h_out[:count].copy_(self.d_out[:count], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return self.h_out_np[:count].copy()


_stages = {}


def _stage(n3, dev):
    key = (dev, torch.cuda.current_stream(dev).cuda_stream, n3)
    st = _stages.get(key)
    if st is None:
        if len(_stages) > 8:                 # a handful of system sizes per process is the use case
            _stages.clear()
        st = _stages[key] = _HostStage(n3, dev)
    return st


def _as_device_f64(x, n3, dev):
    """(N,3)/flat/strided NumPy or torch -> contiguous flat float64 CUDA tensor (a copy only if needed)."""
    if isinstance(x, torch.Tensor):
        t = x.to(device=dev, dtype=torch.float64).reshape(-1)
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1))).to(dev)
    if t.numel() != n3:
        raise ValueError("vector has %d entries, expected %d" % (t.numel(), n3))
    return t.contiguous()


def mobility_product(block, wall, r_vectors, vec, eta, a, periodic_length=None, vec_b=None,
                     precision=FP64, flags=0, targets=None, out=None):
    """out = rows `targets` of the (block, wall) blob mobility applied to vec.

    targets = (offset, count) restricts the computed rows to a shard of the
    blobs (multi-GPU: each rank owns a target range, sources are replicated).
    """
    lib = _lib.load()
    host_in = not isinstance(r_vectors, torch.Tensor)
    dev = _device(None if host_in else r_vectors.device)
    n = (r_vectors.numel() if isinstance(r_vectors, torch.Tensor) else np.asarray(r_vectors).size) // 3
    off, cnt = (0, n) if targets is None else targets
    with torch.cuda.device(dev):
        stage = None
        if host_in and out is None and n > 0 and not isinstance(vec, torch.Tensor):
            stage = _stage(3 * n, dev)
            ups = stage.upload([r_vectors, vec] + ([vec_b] if block == UF_UT else []))
            r, va, vb = ups[0], ups[1], (ups[2] if block == UF_UT else None)
            out = stage.d_out[:3 * cnt]
        else:
            r = _as_device_f64(r_vectors, 3 * n, dev)
            va = _as_device_f64(vec, 3 * n, dev)
            vb = _as_device_f64(vec_b, 3 * n, dev) if block == UF_UT else None
        Lc = _length3(periodic_length)
        if out is None:
            out = torch.empty(3 * cnt, dtype=torch.float64, device=dev)
        if n > 0 and cnt > 0:
            nbytes = lib.rmb_workspace_bytes(block, precision, cnt, n, Lc)
            ws = _workspace(nbytes, dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.rmb_mobility_product(
                block, wall, precision, flags | default_flags, cnt, n, off,
                r.data_ptr(), va.data_ptr(), vb.data_ptr() if vb is not None else None,
                float(eta), float(a), Lc, out.data_ptr(), ws.data_ptr(), ws.numel(), stream))
        if stage is not None:
            return stage.download(3 * cnt, dev)
        return out.cpu().numpy() if host_in else out


def mobility_source_target(wall, source, target, force, radius_source, radius_target, eta, periodic_length=None,
                           precision=FP64, flags=0, out=None):
    """Flow at `target` created by `force` on `source`, blobs with individual radii (argument order of
    mobility.py:551: source first).  The wall form clamps heights to each blob's radius and damps (mobility.py:87-119)."""
    lib = _lib.load()
    host_in = not isinstance(source, torch.Tensor)
    dev = _device(None if host_in else source.device)
    ns = (source.numel() if isinstance(source, torch.Tensor) else np.asarray(source).size) // 3
    nt = (target.numel() if isinstance(target, torch.Tensor) else np.asarray(target).size) // 3
    with torch.cuda.device(dev):
        src = _as_device_f64(source, 3 * ns, dev)
        trg = _as_device_f64(target, 3 * nt, dev)
        f = _as_device_f64(force, 3 * ns, dev)
        rs = _as_device_f64(radius_source, ns, dev)
        rt = _as_device_f64(radius_target, nt, dev)
        L = np.zeros(3) if periodic_length is None else np.asarray(
            periodic_length.cpu() if isinstance(periodic_length, torch.Tensor) else periodic_length, dtype=np.float64)
        Lc = (ctypes.c_double * 3)(*[float(x) for x in L.reshape(-1)[:3]])
        if out is None:
            out = torch.empty(3 * nt, dtype=torch.float64, device=dev)
        if nt > 0:
            nbytes = lib.rmb_source_target_workspace_bytes(precision, nt, ns, Lc)
            ws = _workspace(nbytes, dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.rmb_mobility_source_target(
                wall, precision, flags | default_flags, nt, ns, trg.data_ptr(), src.data_ptr(), f.data_ptr(),
                rt.data_ptr(), rs.data_ptr(), float(eta), Lc, out.data_ptr(), ws.data_ptr(), ws.numel(), stream))
        return out.cpu().numpy() if host_in else out


def _make_source_target(wall, precision=FP64, doc=""):
    def f(source, target, force, radius_source, radius_target, eta, *args, **kwargs):
        return mobility_source_target(wall, source, target, force, radius_source, radius_target, eta,
                                      periodic_length=kwargs.get("periodic_length"), precision=precision)
    f.__doc__ = doc
    return f


def mobility_radii_trans_times_force(r_vectors, force, eta, a, radius_blobs, function, *args, **kwargs):
    """mobility.py:1369: M f for blobs with different radii; `function` is one of the *_source_target_* products
    (the argument `a` is ignored, as in the reference)."""
    return function(r_vectors, r_vectors, force, radius_blobs, radius_blobs, eta, *args, **kwargs)


def dense_mobility(wall, r_vectors, eta, a):
    """Dense (3N, 3N) translational blob mobility, the matrix whose product with a force vector mobility_product(UF,
    wall, ...) computes (open boundaries): what the reference assembles in Python loops / C++ for the bodies'
    preconditioner blocks and the dense schemes (mobility.py:967-1116 -> multi_bodies.set_mobility_blobs,
    multi_bodies.py:205-227).  One CUDA launch (rmb_body_mobility_blocks with the whole cloud as a single body)."""
    from . import rigid
    if wall not in (SINGLE_WALL, NO_WALL):
        raise ValueError("dense_mobility: wall must be SINGLE_WALL or NO_WALL")
    host_in = not isinstance(r_vectors, torch.Tensor)
    dev = _device(None if host_in else r_vectors.device)
    n = (r_vectors.numel() if isinstance(r_vectors, torch.Tensor) else np.asarray(r_vectors).size) // 3
    if n > 30000:
        raise ValueError("dense_mobility: %d blobs would need a %.0f GB matrix; use the matrix-free product" %
                         (n, 9e-9 * 8 * n * n))
    with torch.cuda.device(dev):
        r = _as_device_f64(r_vectors, 3 * n, dev)
        M = torch.empty(3 * n, 3 * n, dtype=torch.float64, device=dev)
        if n > 0:
            off = torch.tensor([0, n], dtype=torch.int64, device=dev)
            rigid._check(rigid._rlib().rmb_body_mobility_blocks(
                1, off.data_ptr(), n, r.data_ptr(), float(eta), float(a), int(wall == SINGLE_WALL), M.data_ptr(),
                torch.cuda.current_stream(dev).cuda_stream))
    return M.cpu().numpy() if host_in else M


def single_wall_fluid_mobility_b200(r_vectors, eta, a, *args, **kwargs):
    """B200 drop-in for mobility.single_wall_fluid_mobility / _cpp (mobility/mobility.py:1015-1116): dense mobility
    above a wall, heights clamped to z >= a and damped by B on both sides."""
    return dense_mobility(SINGLE_WALL, r_vectors, eta, a)


def rotne_prager_tensor_b200(r_vectors, eta, a, *args, **kwargs):
    """B200 drop-in for mobility.rotne_prager_tensor / _cpp (mobility/mobility.py:960-1013): dense free-space RPY."""
    return dense_mobility(NO_WALL, r_vectors, eta, a)


_ZERO_L = (ctypes.c_double * 3)(0.0, 0.0, 0.0)


def _length3(periodic_length):
    if periodic_length is None:
        return _ZERO_L
    L = np.asarray(periodic_length.cpu() if isinstance(periodic_length, torch.Tensor) else periodic_length,
                   dtype=np.float64)
    return (ctypes.c_double * 3)(*[float(x) for x in L.reshape(-1)[:3]])


def symmetric_partial(r_vectors, force, eta, a, rank=0, world=1, acc=None, flags=0, wall=SINGLE_WALL,
                      periodic_length=None, block=UF, vec_b=None):
    """Multi-rank pair-symmetric product, step 1 (rmb_symmetric_partial): this rank's share of the unordered
    pairs, raw sums for ALL blobs -> acc (device tensor, summed over the ranks by the caller).  block = UF / UT / WF /
    WT / UF_UT (vec_b = torque), wall = SINGLE_WALL or NO_WALL; periodic_length may be periodic in x and/or y only."""
    lib = _lib.load()
    dev = _device(r_vectors.device if isinstance(r_vectors, torch.Tensor) else None)
    n = (r_vectors.numel() if isinstance(r_vectors, torch.Tensor) else np.asarray(r_vectors).size) // 3
    with torch.cuda.device(dev):
        r = _as_device_f64(r_vectors, 3 * n, dev)
        f = _as_device_f64(force, 3 * n, dev)
        vb = _as_device_f64(vec_b, 3 * n, dev) if block == UF_UT else None
        Lc = _length3(periodic_length)
        if acc is None:
            acc = torch.empty(lib.rmb_symmetric_acc_len(n), dtype=torch.float64, device=dev)
        ws = _workspace(lib.rmb_workspace_bytes(block, FP64, n, n, Lc), dev)
        _lib.check(lib.rmb_symmetric_partial(block, wall, flags | default_flags, n, rank, world, r.data_ptr(), f.data_ptr(),
                                             vb.data_ptr() if vb is not None else None,
                                             float(eta), float(a), Lc, acc.data_ptr(), ws.data_ptr(), ws.numel(),
                                             torch.cuda.current_stream(dev).cuda_stream))
    return acc


def symmetric_finish(r_vectors, vec, acc, eta, a, targets=None, out=None, flags=0, wall=SINGLE_WALL, block=UF,
                     vec_b=None):
    """Step 2 (rmb_symmetric_finish) after the ranks' acc have been summed: self term, normalisation, damping for
    the target range, from the same r_vectors / vec (/ vec_b) the partial step was given.  Stateless: it reads its
    arguments only, so other products may run between the two steps."""
    lib = _lib.load()
    dev = acc.device
    n = (r_vectors.numel() if isinstance(r_vectors, torch.Tensor) else np.asarray(r_vectors).size) // 3
    off, cnt = (0, n) if targets is None else targets
    with torch.cuda.device(dev):
        r = _as_device_f64(r_vectors, 3 * n, dev)
        f = _as_device_f64(vec, 3 * n, dev)
        vb = _as_device_f64(vec_b, 3 * n, dev) if block == UF_UT else None
        if out is None:
            out = torch.empty(3 * cnt, dtype=torch.float64, device=dev)
        _lib.check(lib.rmb_symmetric_finish(block, wall, flags | default_flags, n, off, cnt, r.data_ptr(), f.data_ptr(),
                                            vb.data_ptr() if vb is not None else None, acc.data_ptr(), float(eta),
                                            float(a), out.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
    return out


_DETERMINISTIC_DOC = ("\n    Extra keyword (not in the reference): deterministic=True forces the one-thread-per-target row kernel "
                      "(fixed summation order, bit-reproducible, like the reference's per-row sum).  By default products "
                      "of >= %d blobs use the pair-symmetric kernel, whose atomics make the last bits run-dependent."
                      % _lib.SYMMETRIC_MIN_N)


def _kw_flags(kwargs, flags):
    return flags | (DETERMINISTIC if kwargs.get("deterministic") else 0)


def _make(block, wall, flags=0, precision=FP64, doc=""):
    def f(r_vectors, vec, eta, a, *args, **kwargs):
        return mobility_product(block, wall, r_vectors, vec, eta, a, periodic_length=kwargs.get("periodic_length"),
                                precision=precision, flags=_kw_flags(kwargs, flags))
    f.__doc__ = doc + _DETERMINISTIC_DOC
    return f


def _make_fused(wall, precision=FP64, doc=""):
    def f(r_vectors, force, torque, eta, a, *args, **kwargs):
        return mobility_product(UF_UT, wall, r_vectors, force, eta, a, vec_b=torque,
                                periodic_length=kwargs.get("periodic_length"), precision=precision,
                                flags=_kw_flags(kwargs, 0))
    f.__doc__ = doc + _DETERMINISTIC_DOC
    return f


_TABLE = [
    # name stem, block, wall, flags, reference lines in mobility/mobility.py (numba; pycuda)
    ("no_wall_mobility_trans_times_force", UF, NO_WALL, 0, ":1119; :288"),
    ("single_wall_mobility_trans_times_force", UF, SINGLE_WALL, 0, ":1132; :222"),
    ("in_plane_mobility_trans_times_force", UF, SINGLE_WALL, IN_PLANE, ":1166; :255"),
    ("no_wall_mobility_trans_times_torque", UT, NO_WALL, 0, ":1200; :482"),
    ("single_wall_mobility_trans_times_torque", UT, SINGLE_WALL, 0, ":1213; :427"),
    ("in_plane_mobility_trans_times_torque", UT, SINGLE_WALL, IN_PLANE, ":1235; :454"),
    ("no_wall_mobility_rot_times_force", WF, NO_WALL, 0, ":1258; :328"),
    ("single_wall_mobility_rot_times_force", WF, SINGLE_WALL, 0, ":1273; :300"),
    ("no_wall_mobility_rot_times_torque", WT, NO_WALL, 0, ":1301; :370"),
    ("single_wall_mobility_rot_times_torque", WT, SINGLE_WALL, 0, ":1316; :342"),
    ("free_surface_mobility_trans_times_force", UF, FREE_SURFACE, 0, ":1390; :533"),
]

__all__ = ["mobility_product", "mobility_source_target", "mobility_radii_trans_times_force", "dense_mobility",
           "single_wall_fluid_mobility_b200", "rotne_prager_tensor_b200"]
for _stem, _b, _w, _fl, _ref in _TABLE:
    for _suffix, _prec in (("_b200", FP64), ("_b200_fp32", FP32)):
        _name = _stem + _suffix
        globals()[_name] = _make(_b, _w, _fl, _prec,
                                 "B200 drop-in for mobility.%s_numba / _pycuda (mobility/mobility.py%s)." % (_stem, _ref))
        globals()[_name].__name__ = _name
        __all__.append(_name)
for _wname, _w, _ref in (("single_wall", SINGLE_WALL, ":384"), ("no_wall", NO_WALL, ":413")):
    for _suffix, _prec in (("_b200", FP64), ("_b200_fp32", FP32)):
        _name = "%s_mobility_trans_times_force_torque%s" % (_wname, _suffix)
        globals()[_name] = _make_fused(_w, _prec, "B200 drop-in for mobility.%s_mobility_trans_times_force_torque_pycuda "
                                       "(mobility/mobility.py%s): UF f + UT t in one pass." % (_wname, _ref))
        globals()[_name].__name__ = _name
        __all__.append(_name)
for _wname, _w, _ref in (("single_wall", SINGLE_WALL, ":494; :551"), ("no_wall", NO_WALL, ":588"),
                          ("free_surface", FREE_SURFACE, ":1407")):
    for _suffix, _prec in (("_b200", FP64), ("_b200_fp32", FP32)):
        _name = "%s_mobility_trans_times_force_source_target%s" % (_wname, _suffix)
        globals()[_name] = _make_source_target(
            _w, _prec, "B200 drop-in for mobility.%s_mobility_trans_times_force_source_target_numba / _pycuda "
            "(mobility/mobility.py%s): separate source / target sets, per-blob radii." % (_wname, _ref))
        globals()[_name].__name__ = _name
        __all__.append(_name)
