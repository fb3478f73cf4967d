#!/usr/bin/env python
"""Small driver for ncu captures: one leaflet of BASELINE configs[2] (499 849 vertices), KDE type 2,
`--mode pairs` (all pairs, sigma 20, 'all') or `--mode cells` (cut-off 8 sigma; sigma 40 and 10),
plus `--mode mesh` (topology + fused properties) and `--mode normals` (point normals)."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from memsurfer_b200 import synth, TriMesh                      # noqa: E402
from memsurfer_b200.frame import LeafletFrame                  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="cells")
ap.add_argument("--n", type=int, default=500000)
a = ap.parse_args()
pts, labels, bbox = synth.make_leaflet(a.n, seed=1234, ntypes=3, fractions=[0.4, 0.3, 0.3])
mesh = TriMesh(pts, periodic=True)
mesh.set_bbox(pts.min(axis=0), pts.max(axis=0))
if a.mode in ("mesh", "all"):
    df, do = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
    t = synth.trim_periodic(pts, df, do, pts.min(axis=0), pts.max(axis=0))
    mesh.set_periodic_triangulation(t["faces"], t["periodic_faces"], t["trimmed_faces"], t["duplicated_vidxs"], t["duplicated_verts"])
else:
    mesh.set_periodic_triangulation(np.zeros((1, 3), np.uint32), np.zeros((0, 3), np.uint32), np.zeros((0, 3), np.uint32),
                                    np.zeros(0, np.uint32), np.zeros((0, 3), np.float32))
fr = LeafletFrame(pts, labels, bbox, mesh)
for rep in range(2):
    if a.mode in ("pairs", "all"):
        fr.run_density("p", 2, 20.0, True, None, 0.0)
    if a.mode in ("cells", "all"):
        fr.run_density("c40", 2, 40.0, True, None, 8.0)
        fr.run_density("c10", 2, 10.0, True, fr.label_ids[1], 8.0)
    if a.mode == "labels":
        fr.run_densities([10.0, 20.0, 30.0, 40.0], cutoff_sigmas=7.5)      # the bench.py call
    if a.mode in ("mesh", "all"):
        fr.run_properties()
    if a.mode in ("normals", "all"):
        fr.run_pnormals(18)
torch.cuda.synchronize()
print("ok", a.mode)
