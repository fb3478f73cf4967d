#!/usr/bin/env python
"""Print the key numbers of one or more bench.py JSON lines."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "ERR", e); continue
    print("==", f, " value", round(d["value"], 3), " ms/step", round(d["ms_per_step"], 3), " launches", d.get("gpu_launches"))
    e = d.get("e2e")
    if e: print("   e2e", round(e["value"], 3), "ms", round(e.get("ms_per_step", 0), 2), "h2d", e["h2d_bytes_per_step"], "d2h", e["d2h_bytes_per_step"])
    r = d.get("roofline")
    if r: print("   roofline", r["kernel"], "frac", round(r["frac"], 4), "ach", round(r["achieved"], 2), "avg_ms", round(r["avg_launch_ms"], 3),
                "pairs/launch %.3g" % r["pairs_per_launch"], "share", round(r["share_of_step"], 3))
    g = d.get("roofline_geometry")
    if g: print("   geom frac", round(g["frac"], 4), "ms/mesh", round(g["avg_ms_per_mesh"], 3))
    k = d.get("kernel_ms_per_step")
    if k: print("   " + ", ".join(f"{a} {b}" for a, b in list(k.items())[:14]))
    c = d.get("cpu_baseline")
    if c: print("   cpu", c["value"], c["cores"], c["kind"])
