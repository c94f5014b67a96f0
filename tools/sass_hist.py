"""Dev tool: per-kernel SASS opcode histogram, split into (whole function).
usage: python tools_sass_hist.py <lib.so> <substring of mangled kernel name>"""
import collections, re, subprocess, sys
txt = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n')[0]
    if sys.argv[2] not in name: continue
    ops = collections.Counter()
    for line in f.split('\n'):
        m = re.search(r'/\*[0-9a-f]{4}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)', line)
        if m: ops[m.group(2).split('.')[0]] += 1
    print(name[:100]); print(sorted(ops.items(), key=lambda x: -x[1]))
