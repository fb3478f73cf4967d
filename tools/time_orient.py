#!/usr/bin/env python
"""Phase times of the cooperative MST-orientation kernel on one BASELINE configs[2] leaflet (diagnostic)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from memsurfer_b200 import synth, point_normals_device
from memsurfer_b200.membrane import orient_normals_device

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
pts, labels, bbox = synth.make_leaflet(n, seed=1234, ntypes=3, fractions=[0.4, 0.3, 0.3])
keep = {}
o32, o64, info = point_normals_device(torch.from_numpy(pts).cuda(), bbox, True, 0.2, 18, +1, keep=keep)
print("m", info['m'], "rounds", info['mst_rounds'], "ties", info['knn_ties'], "unoriented", info['n_unoriented'])
pa, nbr, pca = keep['points_all'].contiguous(), keep['nbr'].contiguous(), keep['normals_pca'].contiguous()
for it in range(3):
    nrm = pca.clone()
    times = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    orient_normals_device(pa, nbr, nrm, phase_times=times)
    e1.record(); torch.cuda.synchronize()
    print(f"run {it}: total {e0.elapsed_time(e1)*1e3:.0f} us (incl. host read-backs); kernel phases sum {sum(times):.0f} us")
    print("   " + " ".join(f"{t:.0f}" for t in times))
assert torch.equal(nrm, keep['normals_all'])
