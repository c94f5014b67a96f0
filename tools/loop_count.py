"""SASS instruction counts of the pair loops of the built library: the figures bench.py's roofline uses.

  python tools/loop_count.py                       print the flagship's counts
  python tools/loop_count.py --write               (re)write profiles/sass_counts.json   (__graft_entry__.build() does)
  python tools/loop_count.py --hist <substring>    opcode histogram of every inner loop of the matching kernels

The flagship is sym_kernel<UF, TMA, NT = 1, WALL, !PER>; its hot loop is the 32-step rotation of sym_block (the
first innermost backward branch of the function).  One trip = one unordered pair per lane = TWO interactions.  The
near-field coefficients sit behind a warp-uniform forward branch (`@!P0 BRA`) inside the loop: instructions between
that branch and its target are the near path, executed only when some lane of the warp has r < 2a."""
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "rigidmultiblobswall_b200", "librmb_mobility.so")
OUT = os.path.join(ROOT, "profiles", "sass_counts.json")
FLAGSHIP = "sym_kernelILi0ELb1ELi1ELb1ELb0"      # <UF, TMA, NT = 1, WALL, !PER>
FP64 = ("DFMA", "DMUL", "DADD", "DSETP")


def _cuobjdump():
    for cand in ("cuobjdump", "/usr/local/cuda/bin/cuobjdump"):
        try:
            subprocess.run([cand, "--version"], capture_output=True, check=True)
            return cand
        except (OSError, subprocess.CalledProcessError):
            pass
    raise RuntimeError("cuobjdump not found")


def functions(so):
    txt = subprocess.run([_cuobjdump(), "-sass", so], capture_output=True, text=True, check=True).stdout
    for f in re.split(r"\n\s*Function : ", txt)[1:]:
        name = f.split("\n")[0]
        ins = []
        for line in f.split("\n"):
            m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
            if m:
                ins.append((int(m.group(1), 16), m.group(2).strip()))
        yield name, ins


def _op(text):
    return re.sub(r"^@!?U?P\d\s+", "", text).split()[0].split(".")[0]


def inner_loops(ins, min_bytes=0x400, max_bytes=0x3000):
    """[(first address, last address, body)] of the backward branches spanning min..max bytes, in address order"""
    out = []
    for a, t in ins:
        m = re.search(r"BRA\S*\s+(?:\S+,\s*)?`?\(?0x([0-9a-f]+)", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < a and min_bytes < a - tgt < max_bytes:
                out.append((tgt, a, [x for x in ins if tgt <= x[0] <= a]))
    return out


def count_loop(body):
    ops = collections.Counter(_op(t) for _, t in body)
    fp64 = sum(ops[k] for k in FP64)
    near = 0
    for a, t in body:                                   # forward, predicated, warp-uniform branch around the near path
        m = re.match(r"@!?U?P\d\s+BRA\s+(?:\S+,\s*)?`?\(?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) > a and int(m.group(1), 16) <= body[-1][0]:
            tgt = int(m.group(1), 16)
            near = sum(1 for x in body if a < x[0] < tgt and _op(x[1]) in FP64)
            break
    return {"total": len(body), "fp64": fp64, "other": len(body) - fp64, "fp64_near_path": near,
            "fp64_far_pair": fp64 - near, "opcodes": dict(sorted(ops.items(), key=lambda kv: -kv[1]))}


def flagship_counts(so=SO):
    for name, ins in functions(so):
        if FLAGSHIP in name:
            loops = inner_loops(ins)
            if not loops:
                raise RuntimeError("no inner loop found in " + name)
            c = count_loop(loops[0][2])
            c["kernel"] = "sym_kernel<UF, TMA, NT=1, WALL, !PER>"
            c["mangled"] = name.strip()[-64:]
            c["loop"] = "rotation step of sym_block: one unordered pair per lane = two interactions"
            c["fp64_per_interaction"] = c["fp64_far_pair"] / 2.0
            return c
    raise RuntimeError("flagship kernel not found in " + so)


def write(so=SO, out=OUT):
    c = flagship_counts(so)
    with open(out, "w") as f:
        json.dump({"flagship": c}, f, indent=1)
        f.write("\n")
    return c


if __name__ == "__main__":
    if "--hist" in sys.argv:
        key = sys.argv[sys.argv.index("--hist") + 1]
        for name, ins in functions(SO if len(sys.argv) < 5 else sys.argv[4]):
            if key in name:
                print(name[-90:])
                for a0, a1, body in inner_loops(ins):
                    c = count_loop(body)
                    print(hex(a0), hex(a1), {k: c[k] for k in ("total", "fp64", "other", "fp64_near_path")},
                          list(c["opcodes"].items())[:16])
    else:
        c = write() if "--write" in sys.argv else flagship_counts()
        print(json.dumps({k: v for k, v in c.items() if k != "opcodes"}))
