#!/usr/bin/env python
"""Per-kernel times of one BASELINE configs[3] frame (100k lipids, 8 types: 2 leaflets of ~50k points), from the
library's own per-launch events (ms_profile_enable) -- where the 2.3 ms of a trajectory frame go."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from memsurfer_b200 import _lib
from memsurfer_b200.frame import LeafletFrame
dev = torch.device("cuda:0")
L = _lib.lib()
lfs = bench.make_frame_inputs(100000, 0, device=dev, ntypes=8, fractions=None)
frs = [LeafletFrame(lf["points"], lf["labels"], lf["bbox"], bench.new_mesh(lf, dev), device=dev) for lf in lfs]
for _ in range(3):
    for fr in frs:
        fr.run(bench.SIGMAS, 18, 7.5); fr.summary()
torch.cuda.synchronize()
L.ms_profile_enable(1)
REP = 10
for _ in range(REP):
    for fr in frs:
        fr.run(bench.SIGMAS, 18, 7.5); fr.summary()
torch.cuda.synchronize()
prof = _lib.profile_report()
L.ms_profile_enable(0)
tot = sum(v["ms"] for v in prof.values())
print(f"sum of kernel times per frame: {tot / REP:.3f} ms, {sum(v['launches'] for v in prof.values()) / REP:.0f} launches")
for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
    print(f"  {k:22s} {v['ms'] / REP:7.4f} ms  {v['launches'] / REP:5.1f} launches  {v['ms'] / v['launches'] * 1e3:8.1f} us each")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(REP):
    for fr in frs:
        fr.run(bench.SIGMAS, 18, 7.5); fr.summary()
e1.record(); torch.cuda.synchronize()
print(f"one stream, back to back: {e0.elapsed_time(e1) / REP:.3f} ms per frame")
