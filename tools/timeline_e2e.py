#!/usr/bin/env python
"""Wall-clock phases of one end-to-end frame through the drop-in API (host numpy in, numpy out)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from memsurfer_b200 import Membrane
leaflets = bench.make_frame_inputs(1000000, 0)
dev = torch.device("cuda:0")

def step(log):
    T = [time.perf_counter()]
    def mark(name):
        T.append(time.perf_counter()); log.append((name, (T[-1] - T[-2]) * 1e3))
    ms = []
    for i, lf in enumerate(leaflets):
        m = Membrane(lf["points"], labels=lf["labels"], periodic=True, bbox=lf["bbox"], device=dev); mark(f"L{i} init")
        m.compute_pnormals(18); mark(f"L{i} pnormals")
        m.memb_smooth = bench.new_mesh(lf, dev); mark(f"L{i} new_mesh")
        m.compute_properties('smooth'); mark(f"L{i} properties(queue)")
        m._label_table(); mark(f"L{i}   label table")
        m._prefetch_densities([2], bench.SIGMAS); mark(f"L{i}   prefetch (labels upload + ms_kde_labels_multi)")
        for l in ['all', 0, 1, 2]:
            Membrane.compute_densities([m], [2], bench.SIGMAS, True, l); mark(f"L{i}   densities {l}")
        ms.append(m)
    for i, m in enumerate(ms):
        m.memb_smooth.attributes.resolve(); m.attributes.resolve(); mark(f"L{i} resolve")
    torch.cuda.synchronize(); mark("sync")

for _ in range(3): step([])
log = []; t0 = time.perf_counter(); step(log); tot = (time.perf_counter() - t0) * 1e3
for n, v in log: print(f"{n:26s} {v:7.2f} ms")
print(f"{'total':26s} {tot:7.2f} ms")
