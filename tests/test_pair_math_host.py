"""Checks the product's restructured per-pair algebra (csrc/pair_math.cuh +
mf_core.cuh) on the CPU: the headers are compiled with g++ into a test-only
harness and compared with the reference's golden outputs.  No GPU needed."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BLOCKS = {"UF": 0, "UT": 1, "WF": 2, "WT": 3, "UF_UT": 4, "UF_FS": 5}


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("h") / "host_pair_harness.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-DRMB_HOST_ONLY", "-shared", "-fPIC",
                           os.path.join(HERE, "host_pair_harness.cpp"), "-o", so])
    lib = ctypes.CDLL(so)
    dp = ctypes.POINTER(ctypes.c_double)
    lib.host_mf.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long, dp, dp, dp,
                            ctypes.c_double, ctypes.c_double, dp, ctypes.c_int, dp]

    lib.host_st.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_long, dp, dp, dp, dp, dp,
                            ctypes.c_double, dp, dp]

    def run_st(wallmode, src, trg, f, rs, rt, eta, L, fp32=False):
        arrs = [np.ascontiguousarray(x, dtype=np.float64) for x in (trg, src, f, rt, rs, L)]
        out = np.empty(arrs[0].size)
        p = lambda x: x.ctypes.data_as(dp)
        lib.host_st(wallmode, int(fp32), arrs[0].size // 3, arrs[1].size // 3, p(arrs[0]), p(arrs[1]), p(arrs[2]),
                    p(arrs[3]), p(arrs[4]), eta, p(arrs[5]), p(out))
        return out

    lib.host_uf_sym.argtypes = [ctypes.c_int, ctypes.c_long, dp, dp, ctypes.c_double, ctypes.c_double, dp]

    def run_sym(r, f, eta, a, wall=True):
        r = np.ascontiguousarray(r, dtype=np.float64); f = np.ascontiguousarray(f, dtype=np.float64)
        out = np.empty(r.size)
        lib.host_uf_sym(int(wall), r.shape[0], r.ctypes.data_as(dp), f.ctypes.data_as(dp), eta, a, out.ctypes.data_as(dp))
        return out

    lib.host_sym_periodic.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, dp, dp, ctypes.c_double,
                                      ctypes.c_double, dp, dp]

    def run_sym_periodic(r, f, eta, a, L, wall=True, block="UF"):
        r, f, L = (np.ascontiguousarray(x, dtype=np.float64) for x in (r, f, L))
        out = np.empty(r.size)
        lib.host_sym_periodic(BLOCKS[block], int(wall), r.shape[0], r.ctypes.data_as(dp), f.ctypes.data_as(dp), eta, a,
                              L.ctypes.data_as(dp), out.ctypes.data_as(dp))
        return out

    def run(block, wall, r, va, eta, a, L, fp32=False, in_plane=False, vb=None):
        r = np.ascontiguousarray(r, dtype=np.float64)
        va = np.ascontiguousarray(va, dtype=np.float64)
        vb = va if vb is None else np.ascontiguousarray(vb, dtype=np.float64)
        L = np.ascontiguousarray(L, dtype=np.float64)
        out = np.empty(r.size)
        p = lambda x: x.ctypes.data_as(dp)
        lib.host_mf(BLOCKS[block], int(wall), int(fp32), r.shape[0], p(r), p(va), p(vb), eta, a, p(L), int(in_plane), p(out))
        return out
    run.st = run_st
    run.sym = run_sym
    run.sym_periodic = run_sym_periodic
    return run


def _cases():
    g = np.load(os.path.join(HERE, "golden", "mf_golden.npz"))
    return [str(c) for c in g["cases"]]


def relerr(x, ref):
    return np.linalg.norm(x - ref) / np.linalg.norm(ref)


@pytest.mark.parametrize("case", _cases())
def test_algebra_fp64(golden, harness, case):
    cloud, lname, wall, blk = case.split("/")
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    out = harness("UF_FS" if wall == "free_surface" else blk, wall != "no_wall", r, v, eta, a, L,
                  in_plane=(wall == "in_plane"))
    assert relerr(out, golden[case]) < 1e-12   # BASELINE.json: fp64 parity bar


@pytest.mark.parametrize("case", [c for c in _cases() if c.startswith("dilute/") or c.startswith("c1/")])
def test_algebra_fp32(golden, harness, case):
    cloud, lname, wall, blk = case.split("/")
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    out = harness("UF_FS" if wall == "free_surface" else blk, wall != "no_wall", r, v, eta, a, L, fp32=True,
                  in_plane=(wall == "in_plane"))
    assert relerr(out, golden[case]) < 1e-5    # BASELINE.json: fp32 parity bar


@pytest.mark.parametrize("lname", ["open", "xy"])
@pytest.mark.parametrize("wall", ["single_wall", "no_wall"])
def test_fused_force_torque(golden, harness, lname, wall):
    """Fused UF f + UT t (mobility_pycuda.py:1266; no numba twin): oracle = sum of the two golden products."""
    r, f = golden["dense/r"], golden["dense/v"]
    a, eta = golden["dense/a_eta"]
    L = golden[f"dense/L_{lname}"]
    t = f[::-1].copy()
    ut = harness("UT", wall != "no_wall", r, t, eta, a, L)
    ref = golden[f"dense/{lname}/{wall}/UF"] + ut
    out = harness("UF_UT", wall != "no_wall", r, f, eta, a, L, vb=t)
    assert relerr(out, ref) < 1e-12


# ---- source-target kernels with per-blob radii (csrc/st_math.cuh) ----
ST_WALLMODE = {"no_wall": 0, "single_wall": 1, "radii_single_wall": 1, "free_surface": 2}


def _st_golden():
    return np.load(os.path.join(HERE, "golden", "st_golden.npz"))


@pytest.mark.parametrize("fp32", [False, True])
@pytest.mark.parametrize("case", [str(c) for c in _st_golden()["cases"]])
def test_source_target_algebra(harness, case, fp32):
    g = _st_golden()
    name, lname, wall = case.split("/")
    src, trg, f = g[f"{name}/source"], g[f"{name}/target"], g[f"{name}/force"]
    rs, rt, eta, L = g[f"{name}/radius_source"], g[f"{name}/radius_target"], float(g[f"{name}/eta"]), g[f"{name}/L_{lname}"]
    if wall == "radii_single_wall":
        trg, rt = src, rs
    out = harness.st(ST_WALLMODE[wall], src, trg, f, rs, rt, eta, L, fp32=fp32)
    # the "mixed" cloud has partially overlapping pairs whose fp32 C1 cancels ((a+b)/2 - ...): only the dilute
    # cloud is held to the 1e-5 fp32 bar
    if fp32 and name == "mixed":
        assert relerr(out, g[case]) < 1e-3
    else:
        assert relerr(out, g[case]) < (1e-5 if fp32 else 1e-12)


@pytest.mark.parametrize("cloud", ["dense", "dilute", "c1"])
def test_pair_symmetric_evaluation(golden, harness, cloud):
    """pair_uf_wall_sym: each unordered pair once, u_i += M_ij f_j and u_j += M_ji f_i from shared coefficients"""
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    assert relerr(harness.sym(r, v, eta, a), golden[f"{cloud}/open/single_wall/UF"]) < 1e-12
    if cloud != "c1":
        assert relerr(harness.sym(r, v, eta, a, wall=False), golden[f"{cloud}/open/no_wall/UF"]) < 1e-12


@pytest.mark.parametrize("lname", ["xy", "x"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_pair_symmetric_periodic_evaluation(golden, harness, cloud, lname):
    """pseudo-periodic images in the pair-symmetric scheme: one wrap per unordered pair, then every box with (r, -r)
    (box b seen from i is box -b seen from j); the blob's own images come from the i == j pair"""
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    assert relerr(harness.sym_periodic(r, v, eta, a, L), golden[f"{cloud}/{lname}/single_wall/UF"]) < 1e-12
    assert relerr(harness.sym_periodic(r, v, eta, a, L, wall=False), golden[f"{cloud}/{lname}/no_wall/UF"]) < 1e-12


def test_pair_symmetric_periodic_outside_cell(harness):
    """blobs far outside the primary cell: the truncating wrap is odd, so the two directions still agree"""
    import oracle
    rng = np.random.default_rng(8)
    n, a, eta = 150, 0.3, 1.1
    r = 4.0 * rng.random((n, 3)) * np.array([3.0, 3.0, 1.0]) - np.array([4.0, 4.0, 0.0])
    f = rng.standard_normal((n, 3))
    L = np.array([4.0, 3.5, 0.0])
    for wall, name in ((True, "single_wall"), (False, "no_wall")):
        ref = oracle.mobility_product("UF", name, r, f, eta, a, periodic_length=L)
        assert relerr(harness.sym_periodic(r, f, eta, a, L, wall=wall), ref) < 1e-12


@pytest.mark.parametrize("block", ["UT", "WF", "WT"])
@pytest.mark.parametrize("lname", ["open", "xy", "x"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_pair_symmetric_rotation_blocks(golden, harness, cloud, lname, block):
    """pair_{ut,wf,wt}_wall_sym: the (j <- i) interaction from the (i <- j) one's shared factors, r -> -r and the
    wall term's height exchanged; open and pseudo-periodic"""
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    assert relerr(harness.sym_periodic(r, v, eta, a, L, block=block), golden[f"{cloud}/{lname}/single_wall/{block}"]) < 1e-12
    # free-space twins (pair_coupling_free_sym, pair_wt_free_sym)
    assert relerr(harness.sym_periodic(r, v, eta, a, L, wall=False, block=block),
                  golden[f"{cloud}/{lname}/no_wall/{block}"]) < 1e-12
