#!/usr/bin/env python
"""GPU timeline of one end-to-end frame (torch.profiler / CUPTI): when every kernel and copy ran, how long the GPU
sat idle between them, and what the host was doing at the time.  Diagnostic for the e2e arm."""
import json, os, sys, tempfile
import torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from memsurfer_b200 import Membrane
dev = torch.device("cuda:0")
leaflets = [bench.pin_inputs(lf) for lf in bench.make_frame_inputs(1000000, 0, device=dev)]


def step():
    ms = []
    for lf in leaflets:
        m = Membrane(lf["points"], labels=lf["labels"], periodic=True, bbox=lf["bbox"], device=dev)
        m.compute_pnormals(18, wait=False)
        m.fit_points_to_box_xy()
        m.memb_smooth = bench.new_mesh(lf, dev)
        ms.append(m)
    for m in ms:
        m.compute_properties('smooth')
    for l in ['all', 0, 1, 2]:
        Membrane.compute_densities(ms, [2], bench.SIGMAS, True, l)
    for m in ms:
        m.pnormals
        for k in ('pnormal', 'parea', 'mean_curv', 'gaus_curv'):
            m.memb_smooth.attributes[k]
        list(m.attributes.values())
    torch.cuda.synchronize()


for _ in range(3):
    step()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step()
path = os.path.join(tempfile.gettempdir(), "e2e_trace.json")
prof.export_chrome_trace(path)
ev = json.load(open(path))["traceEvents"]
gpu = sorted([e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e], key=lambda e: e["ts"])
t0 = min(e["ts"] for e in ev if e.get("cat") in ("cpu_op", "cuda_runtime", "kernel", "gpu_memcpy") and "ts" in e)
print(f"{len(gpu)} GPU activities; first at {gpu[0]['ts'] - t0:.0f} us, last ends at {max(e['ts'] + e['dur'] for e in gpu) - t0:.0f} us")
# union of busy intervals (any stream)
busy, cur_s, cur_e = [], None, None
for e in gpu:
    s, t = e["ts"], e["ts"] + e["dur"]
    if cur_s is None:
        cur_s, cur_e = s, t
    elif s <= cur_e:
        cur_e = max(cur_e, t)
    else:
        busy.append((cur_s, cur_e)); cur_s, cur_e = s, t
busy.append((cur_s, cur_e))
print(f"GPU busy {sum(b - a for a, b in busy):.0f} us in {len(busy)} intervals")
gaps = [(busy[i][1], busy[i + 1][0]) for i in range(len(busy) - 1)]
for a, b in sorted(gaps, key=lambda g: g[0] - g[1])[:12]:
    nxt = next(e for e in gpu if e["ts"] >= b - 1e-3)
    print(f"  idle {b - a:7.0f} us at t = {a - t0:7.0f} us, then {nxt['name'][:60]}")
# kernels only (no copies): busy time
kb = sorted([(e["ts"], e["ts"] + e["dur"]) for e in gpu if e["cat"] == "kernel"])
tot, cs, ce = 0.0, None, None
for s, t in kb:
    if cs is None: cs, ce = s, t
    elif s <= ce: ce = max(ce, t)
    else: tot += ce - cs; cs, ce = s, t
tot += ce - cs
print(f"kernels alone cover {tot:.0f} us; first kernel at {kb[0][0] - t0:.0f} us, last kernel ends at {max(t for _, t in kb) - t0:.0f} us")
# intervals without any kernel: which copies ran meanwhile
merged, cs, ce = [], None, None
for s, t in kb:
    if cs is None: cs, ce = s, t
    elif s <= ce: ce = max(ce, t)
    else: merged.append((cs, ce)); cs, ce = s, t
merged.append((cs, ce))
print("kernel-free intervals > 40 us:")
for i in range(len(merged) - 1):
    a, b = merged[i][1], merged[i + 1][0]
    if b - a < 40: continue
    cps = [e for e in gpu if e["cat"] != "kernel" and e["ts"] < b and e["ts"] + e["dur"] > a]
    nxt = next(e for e in gpu if e["cat"] == "kernel" and e["ts"] >= b - 1e-3)
    prev = max((e for e in gpu if e["cat"] == "kernel" and e["ts"] + e["dur"] <= a + 1e-3), key=lambda e: e["ts"] + e["dur"])
    print(f"  {b - a:6.0f} us at t = {a - t0:6.0f}: after {prev['name'][:28]:28s} before {nxt['name'][:28]:28s} copies: "
          + ", ".join(f"{e['name'][7:11]} {e['args'].get('bytes', 0) / 1e6:.1f}MB {e['dur']:.0f}us" for e in cps[:5]))
agg = {}
for e in gpu:
    if e["cat"] == "kernel":
        k = e["name"].split("(")[0][:40]
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += e["dur"]
print("kernel time by name (us):")
for k, (n, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
    print(f"  {k:42s} {n:4d} {d:9.0f}")
