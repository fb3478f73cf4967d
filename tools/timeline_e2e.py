#!/usr/bin/env python
"""Wall-clock phases of one end-to-end frame through the drop-in API (page-locked host numpy in, numpy out),
then a cProfile of the same step (host time per function)."""
import cProfile, os, pstats, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from memsurfer_b200 import Membrane
dev = torch.device("cuda:0")
leaflets = [bench.pin_inputs(lf) for lf in bench.make_frame_inputs(1000000, 0, device=dev)]

def step(log):
    T = [time.perf_counter()]
    def mark(name):
        T.append(time.perf_counter()); log.append((name, (T[-1] - T[-2]) * 1e3))
    ms = []
    for i, lf in enumerate(leaflets):
        m = Membrane(lf["points"], labels=lf["labels"], periodic=True, bbox=lf["bbox"], device=dev); mark(f"L{i} init")
        m.compute_pnormals(18, wait=False); mark(f"L{i} pnormals(queue)")
        m.fit_points_to_box_xy(); mark(f"L{i} fit")
        m.memb_smooth = bench.new_mesh(lf, dev); mark(f"L{i} new_mesh")
        ms.append(m)
    for i, m in enumerate(ms):
        m.compute_properties('smooth'); mark(f"L{i} properties(queue)")
    for l in ['all', 0, 1, 2]:
        Membrane.compute_densities(ms, [2], bench.SIGMAS, True, l); mark(f"densities {l}")
    for i, m in enumerate(ms):
        mesh = m.memb_smooth
        m.pnormals; mark(f"L{i} read pnormals")
        for k in ('pnormal', 'parea', 'mean_curv', 'gaus_curv'):
            mesh.attributes[k]
        mark(f"L{i} read mesh fields")
        list(m.attributes.values()); mark(f"L{i} read densities")
    torch.cuda.synchronize(); mark("sync")

for _ in range(3): step([])
log = []; t0 = time.perf_counter(); step(log); tot = (time.perf_counter() - t0) * 1e3
for n, v in log: print(f"{n:26s} {v:7.2f} ms")
print(f"{'total':26s} {tot:7.2f} ms")
pr = cProfile.Profile()
pr.enable(); step([]); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
