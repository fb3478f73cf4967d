#!/usr/bin/env python
"""Print the fp64-heaviest innermost loop of a kernel's SASS: tools/hotloop.py <lib.so> <mangled-name-substring>"""
import re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = out.split("Function : ")
blk = [b for b in blocks if b.split("\n", 1)[0].find(pat) >= 0][0]
ins = []
for l in blk.splitlines():
    m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
best = None
for a, t in ins:
    m = re.search(r"BRA.* 0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a:
        body = [x for x in ins if int(m.group(1), 16) <= x[0] <= a]
        nd = sum(1 for x in body if re.match(r"(@!?U?P\d\s+)?D(FMA|ADD|MUL|SETP)", x[1]))
        if nd >= int(sys.argv[3]) if len(sys.argv) > 3 else 40:
            if best is None or len(body) < len(best[0]): best = (body, nd)
body, nd = best
ops = {}
for a, t in body:
    op = re.sub(r"^@!?U?P\d\s+", "", t).split()[0].split(".")[0]
    ops[op] = ops.get(op, 0) + 1
print(len(body), "instructions, fp64", nd, ops)
if "-v" in sys.argv:
    for a, t in body: print(hex(a), t)
