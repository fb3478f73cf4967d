#!/usr/bin/env python
"""SASS evidence for profiles/: the TMA bulk-copy + mbarrier instructions and one unrolled pair of the k_kde_groups<2,false>
inner loop (fp64 instruction mix per pair), cut out of the shipped libmemsurfer_b200.so with cuobjdump."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "memsurfer_b200", "libmemsurfer_b200.so")
names = subprocess.run(["cuobjdump", "-elf", so], capture_output=True, text=True).stdout     # noqa: F841 (presence check)
fun = "_Z12k_kde_groupsILi2ELb0EEvPKN4SrcTIXT_EE4typeEPKiS6_PK8KdeGroupS6_S4_S6_il9KdeTilingPjPd"
sass = subprocess.run(["cuobjdump", "-sass", "-fun", fun, so], capture_output=True, text=True).stdout.splitlines()
def instr(l):
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(.*?);", l)
    return m.group(1).strip() if m else None
ins = [(i, instr(l)) for i, l in enumerate(sass)]
ins = [(i, t) for i, t in ins if t]
print(f"# {os.path.basename(so)}: k_kde_groups<2,false> ({len(ins)} SASS instructions), cuobjdump -sass")
print("\n## TMA bulk copies and mbarrier operations (cp.async.bulk / mbarrier.* in the source)")
for i, t in ins:
    if re.search(r"UBLKCP|SYNCS|FENCE", t):
        print(f"  {t}")
fp = [k for k, (i, t) in enumerate(ins) if re.search(r"\b(DFMA|DADD|DMUL|DSETP)\b", t)]
# the unrolled pair loop = the longest run of fp64 instructions at most 6 instructions apart
best = (0, 0); a = 0
for b in range(1, len(fp) + 1):
    if b == len(fp) or fp[b] - fp[b - 1] > 6:
        if b - 1 - a > best[1] - best[0]:
            best = (a, b - 1)
        a = b
lo, hi = fp[best[0]], fp[best[1]]
body = [t for _, t in ins[lo:hi + 1]]
cnt = {}
for t in body:
    op = t.split()[0] if not t.startswith("@") else t.split()[1]
    op = op.split(".")[0]
    cnt[op] = cnt.get(op, 0) + 1
npairs = cnt.get("DMUL", 1)
print(f"\n## unrolled pair loop: {len(body)} instructions for {npairs} pairs (one DMUL per pair)")
print("  per pair: " + ", ".join(f"{k} {v / npairs:.2f}" for k, v in sorted(cnt.items(), key=lambda kv: -kv[1])))
print("\n## first two pairs of the loop body")
seen = 0
for t in body:
    print(f"  {t}")
    if t.split()[0].startswith("DMUL") or (t.startswith("@") and "DMUL" in t):
        seen += 1
    if seen == 3:
        break
