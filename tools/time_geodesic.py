#!/usr/bin/env python
"""Type-1 (geodesic) density: GPU (ms_geodesics + ms_kde_geodesic) against the compiled reference's
TriMesh::kde(type 1) (oracle/_ref, 1 core, float32) on the same mesh."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from memsurfer_b200 import synth, TriMesh
from oracle import oracle as O

for n, with_ref in ((2025, True), (8100, False), (20164, False)):
    pts, labels, bbox = synth.make_leaflet(n, seed=3)
    faces = synth.planar_delaunay(pts[:, :2]).astype(np.uint32)
    TriMesh(pts, faces=faces).compute_density(1, 12.0, "w", True, np.empty([0]))      # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mesh = TriMesh(pts, faces=faces)
    d = mesh.compute_density(1, 12.0, "g", True, np.empty([0]))
    torch.cuda.synchronize()
    tg = time.perf_counter() - t0
    line = f"V {pts.shape[0]:6d}  GPU {tg * 1e3:9.1f} ms (host arrays in, density out)"
    if with_ref:
        t0 = time.perf_counter()
        r = O.ref_kde_geodesic(pts, faces, [], 12.0, True)
        tr = time.perf_counter() - t0
        line += f"   reference C++ {tr:7.2f} s   x{tr / tg:.0f}   max rel diff {np.abs(d - r).max() / np.abs(r).max():.1e}"
    print(line, flush=True)
