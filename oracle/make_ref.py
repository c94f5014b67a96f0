"""TEST / BASELINE INFRASTRUCTURE ONLY -- stages the UNMODIFIED reference next to the oracle.

    python oracle/make_ref.py            (build container; also called by __graft_entry__.build())

The reference is Python with no installer and no package root, so "building" it means placing the
files its own import statements need where the GPU box can see them: /root/reference does not exist there, but
git-ignored, gpurun-shipped oracle/_ref/ does.  Files are copied byte for byte from /root/reference (sha256 of
every file recorded in oracle/_ref/MANIFEST.json and re-checked by tests/test_reference_stage.py); nothing under
oracle/_ref/ is ever committed and the product (rigidmultiblobswall_b200/) never imports it.

What it is used for (tests/, bench.py's cpu_baseline / --impl reference legs only):
  * mobility/mobility_numba.py + mobility/mobility.py   the reference's own CPU implementation of the hot path, timed on
                                                        the GPU box's host cores;
  * multi_bodies/multi_bodies.py and what it imports    the reference's own driver: set_mobility_vector_prod,
                                                        linear_operator_rigid, build_block_diagonal_preconditioner
                                                        (multi_bodies.py:233-287, 424-471, 752-903) under SciPy GMRES with
                                                        the B200 product registered as an implementation string.
Harness-side stubs, never edits: `imp` (removed from Python 3.12; mobility.py:7,
multi_bodies_functions.py:10) and quaternion_integrator/gmres.py (imports a private SciPy module that is gone).
"""
import hashlib
import importlib.util
import json
import os
import shutil
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_DST = os.path.join(HERE, "_ref")

# every *.py of these directories (the closure of `import multi_bodies` + the Lanczos module), plus single files
DIRS = ["mobility", "body", "quaternion_integrator", "read_input", "stochastic_forcing", "constraint", "articulated",
        "Laplace_kernels", "plot"]
SKIP = {"mobility/mobility_pycuda.py",      # needs pycuda (absent); mobility.py skips it when pycuda is not importable
        "quaternion_integrator/gmres.py",   # imports scipy.sparse.linalg._isolve._iterative (gone); stubbed
        "Laplace_kernels/Laplace_kernels_stkfmm.py"}
FILES = ["general_application_utils.py",
         "multi_bodies/multi_bodies.py", "multi_bodies/multi_bodies_functions.py", "multi_bodies/forces_numba.py",
         "multi_bodies/multi_bodies_utilities.py", "multi_bodies/inputfile.dat",
         "multi_bodies/Structures/boomerang_N_15.vertex", "multi_bodies/Structures/blob.vertex",
         "multi_bodies/Structures/shell_N_12_Rg_0_7921_Rh_1.vertex",
         "multi_bodies/Structures/shell_N_162_Rg_0_9497_Rh_1.vertex",
         "multi_bodies/examples/Mobility_Prod_Timing/Multiblob.inputfile"]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def file_list():
    out = list(FILES)
    for d in DIRS:
        for name in sorted(os.listdir(os.path.join(REF_SRC, d))):
            rel = d + "/" + name
            if name.endswith(".py") and rel not in SKIP:
                out.append(rel)
    return out


def available():
    return os.path.exists(os.path.join(REF_DST, "MANIFEST.json"))


def stage(force=False):
    """Copy the reference files into oracle/_ref/ (no-op when /root/reference is absent, e.g. on the GPU box)."""
    if not os.path.isdir(REF_SRC):
        return available()
    manifest = {}
    for rel in file_list():
        src, dst = os.path.join(REF_SRC, rel), os.path.join(REF_DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if force or not os.path.exists(dst) or _sha(dst) != _sha(src):
            shutil.copyfile(src, dst)
        manifest[rel] = _sha(dst)
    with open(os.path.join(REF_DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": REF_SRC, "files": manifest}, f, indent=1, sort_keys=True)
    return True


def verify():
    """sha256 of every staged file against the manifest (and against /root/reference where that exists)."""
    with open(os.path.join(REF_DST, "MANIFEST.json")) as f:
        manifest = json.load(f)["files"]
    bad = []
    for rel, sha in manifest.items():
        if _sha(os.path.join(REF_DST, rel)) != sha:
            bad.append(rel)
        elif os.path.isdir(REF_SRC) and _sha(os.path.join(REF_SRC, rel)) != sha:
            bad.append(rel)
    return bad


def _install_stubs():
    if "imp" not in sys.modules:
        imp = types.ModuleType("imp")

        def find_module(name, path=None):
            if importlib.util.find_spec(name) is None:
                raise ImportError(name)
            return (None, None, None)

        imp.find_module = find_module
        sys.modules["imp"] = imp
    for name in ("quaternion_integrator.gmres", "gmres"):
        sys.modules.setdefault(name, types.ModuleType(name))


def root():
    """Directory to import the reference from: the staged copy (works on both machines)."""
    if not available() and not stage():
        raise RuntimeError("oracle/_ref is not staged and /root/reference is absent: run oracle/make_ref.py in the "
                           "build container (it is called by __graft_entry__.build())")
    return REF_DST


def import_mobility():
    """The reference's mobility/mobility.py (wrappers) with its numba kernels; returns the module."""
    _install_stubs()
    r = root()
    if r not in sys.path:
        sys.path.insert(0, r)
    from mobility import mobility as mb
    return mb


def import_multi_bodies():
    """The reference's driver module multi_bodies/multi_bodies.py, unmodified; returns the module."""
    _install_stubs()
    r = root()
    for p in (os.path.join(r, "multi_bodies"), r):
        if p not in sys.path:
            sys.path.insert(0, p)
    import multi_bodies
    return multi_bodies


if __name__ == "__main__":
    ok = stage(force="--force" in sys.argv)
    print("oracle/_ref staged:", ok, "files:", len(json.load(open(os.path.join(REF_DST, "MANIFEST.json")))["files"]) if ok else 0)
    print("modified files:", verify() if ok else "n/a")
