"""Dev tool: opcode histogram of the inner loops of one kernel.  usage: python tools/loop_count.py <mangled substring> [next function substring]"""
import collections, re, subprocess, sys
txt = subprocess.run(["cuobjdump", "-sass", "rigidmultiblobswall_b200/librmb_mobility.so"], capture_output=True, text=True).stdout
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n')[0]
    if sys.argv[1] not in name: continue
    ins = []
    for l in f.split('\n'):
        m = re.search(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', l)
        if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
    print(name[-90:])
    for a, t in ins:
        m = re.search(r'BRA\S*\s+(?:\S+,\s*)?`?\(?0x([0-9a-f]+)', t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < a and 0x400 < a - tgt < 0x3000:
                body = [x for x in ins if tgt <= x[0] <= a]
                c = collections.Counter(re.sub(r'^@!?U?P\d\s+', '', x[1]).split()[0].split('.')[0] for x in body)
                n64 = c['DFMA'] + c['DMUL'] + c['DADD'] + c['DSETP']
                print(hex(tgt), hex(a), 'total', len(body), 'fp64', n64, 'other', len(body) - n64, sorted(c.items(), key=lambda x: -x[1]))
