"""Algorithmic flop count per pair interaction of the reference's numba kernels: the numerator of
bench.py's roofline.  Parses /root/reference/mobility/mobility_numba.py with `ast` (nothing is imported or run) and
counts, in the body of each kernel's innermost `for j in range(...)` loop, every + - * / and sqrt as ONE flop, the way
the source is written (no CSE, no strength reduction; `x**n` = n-1 multiplications; unary minus, comparisons, int()
casts and index arithmetic are free).  Branches are resolved per path:

    i == j / j_image tests     -> `pair` path (else branch); the self path is reported separately
    r > 2 / r >= 2 tests       -> `far` (if branch) and `near` (else branch) reported separately
    L? > 0 tests               -> the image wrap, reported as `wrap_per_periodic_direction`, not part of `pair_*`
    box? != 0 tests            -> integer bookkeeping, free

    python tools/count_flops.py            (build container only; writes profiles/flops_per_pair.json)
"""
import ast
import json
import os
import sys

REF = "/root/reference/mobility/mobility_numba.py"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "flops_per_pair.json")

KERNELS = [
    "no_wall_mobility_trans_times_force_numba",
    "single_wall_mobility_trans_times_force_numba",
    "in_plane_mobility_trans_times_force_numba",
    "no_wall_mobility_trans_times_torque_numba",
    "single_wall_mobility_trans_times_torque_numba",
    "in_plane_mobility_trans_times_torque_numba",
    "no_wall_mobility_rot_times_force_numba",
    "single_wall_mobility_rot_times_force_numba",
    "no_wall_mobility_rot_times_torque_numba",
    "single_wall_mobility_rot_times_torque_numba",
    "free_surface_mobility_trans_times_force_numba",
]


class Count:
    def __init__(self):
        self.flops = self.sqrt = self.div = 0

    def __iadd__(self, o):
        self.flops += o.flops; self.sqrt += o.sqrt; self.div += o.div
        return self

    def as_dict(self):
        return {"flops": self.flops, "sqrt": self.sqrt, "div": self.div}


def expr_count(node):
    """flops of one expression tree"""
    c = Count()
    for n in ast.walk(node):
        if isinstance(n, ast.BinOp):
            if isinstance(n.op, (ast.Add, ast.Sub, ast.Mult)):
                c.flops += 1
            elif isinstance(n.op, ast.Div):
                c.flops += 1; c.div += 1
            elif isinstance(n.op, ast.Pow):
                e = n.right.value if isinstance(n.right, ast.Constant) else 2
                c.flops += max(int(e) - 1, 1)
        elif isinstance(n, ast.Call):
            f = n.func
            name = f.attr if isinstance(f, ast.Attribute) else getattr(f, "id", "")
            if name == "sqrt":
                c.flops += 1; c.sqrt += 1
    return c


def names(node):
    return {n.id for n in ast.walk(node) if isinstance(n, ast.Name)}


def classify(test):
    ids = names(test)
    if ids & {"boxX", "boxY", "boxZ"}:
        return "box"
    if ids & {"Lx", "Ly", "Lz", "L"}:
        return "periodic"
    if "j_image" in ids or {"i", "j"} <= ids:
        return "self"
    if ids & {"r", "r2", "r_norm"}:
        return "near_far"
    return "other"


def far_is_body(test):
    """True when the `if` branch is the far field (r > 2, r >= 2, r2 > 4); False for r < 2 style tests."""
    if isinstance(test, ast.Compare) and isinstance(test.ops[0], (ast.Gt, ast.GtE)):
        return True
    return False


def walk(stmts, path, acc, wrap, notes):
    """path: dict(self=bool, far=bool).  Adds the flops of the executed statements to acc."""
    for s in stmts:
        if isinstance(s, ast.If):
            kind = classify(s.test)
            if kind == "box":
                continue
            if kind == "periodic":
                w = Count()
                walk(s.body, path, w, wrap, notes)
                w += expr_count(s.test)
                wrap.append(w.flops)
                continue
            if kind == "self":
                walk(s.body if path["self"] else s.orelse, path, acc, wrap, notes)
                continue
            if kind == "near_far":
                far_body = far_is_body(s.test)
                take_body = (path["far"] == far_body)
                walk(s.body if take_body else s.orelse, path, acc, wrap, notes)
                continue
            notes.append("unclassified if at line %d: both branches counted" % s.lineno)
            walk(s.body, path, acc, wrap, notes)
            walk(s.orelse, path, acc, wrap, notes)
        elif isinstance(s, ast.AugAssign):
            acc.flops += 1
            acc += expr_count(s.value)
        elif isinstance(s, (ast.Assign, ast.Expr)):
            acc += expr_count(s.value)
        elif isinstance(s, ast.For):
            notes.append("nested loop at line %d counted once" % s.lineno)
            walk(s.body, path, acc, wrap, notes)


def inner_j_loop(fn):
    best = None
    for n in ast.walk(fn):
        if isinstance(n, ast.For) and isinstance(n.target, ast.Name) and n.target.id == "j":
            best = n
    return best


def main():
    src = open(REF).read()
    tree = ast.parse(src)
    fns = {n.name: n for n in tree.body if isinstance(n, ast.FunctionDef)}
    out = {"source": "mobility/mobility_numba.py", "rule": "one flop per + - * / sqrt as written; x**n = n-1",
           "kernels": {}}
    for name in KERNELS:
        fn = fns[name]
        loop = inner_j_loop(fn)
        entry = {"lines": [loop.lineno, loop.end_lineno]}
        for label, path in (("pair_far", {"self": False, "far": True}), ("pair_near", {"self": False, "far": False}),
                            ("self", {"self": True, "far": True})):
            acc, wrap, notes = Count(), [], []
            walk(loop.body, path, acc, wrap, notes)
            entry[label] = acc.as_dict()
            entry["wrap_per_periodic_direction"] = wrap[0] if wrap else 0
            if notes:
                entry["notes"] = sorted(set(notes))
        out["kernels"][name] = entry
        print("%-52s far %3d (sqrt %d, div %d)  near %3d  self %3d  wrap/dir %d" % (
            name, entry["pair_far"]["flops"], entry["pair_far"]["sqrt"], entry["pair_far"]["div"],
            entry["pair_near"]["flops"], entry["self"]["flops"], entry["wrap_per_periodic_direction"]))
    uf = out["kernels"]["single_wall_mobility_trans_times_force_numba"]["pair_far"]["flops"]
    ut = out["kernels"]["single_wall_mobility_trans_times_torque_numba"]["pair_far"]["flops"]
    out["headline"] = {"single_wall_UF_counted": uf, "single_wall_UF_used_by_bench": 200,
                       "why": "SURVEY 8d fixes 200 flop/pair for the headline so that numbers stay comparable across "
                              "rounds; the counted value is reported next to it",
                       "fused_UF_UT_counted": uf + ut}
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
