#!/usr/bin/env python
"""Wall time of the periodic Delaunay producer: GPU (ms_ghosts + ms_knn + ms_delaunay_periodic, host arrays in and
out) against the host stand-in (scipy/Qhull over the tiled points) on one leaflet of the bench frame."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from memsurfer_b200 import synth, _lib
from memsurfer_b200.delaunay import periodic_delaunay_device

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
pts, _, bbox = synth.make_leaflet(n, seed=1234)
xy = pts[:, :2]
for _ in range(2):
    periodic_delaunay_device(xy, bbox[0], bbox[1])
torch.cuda.synchronize()
_lib.lib().ms_profile_enable(1)
t = time.perf_counter()
for _ in range(5):
    gf, go = periodic_delaunay_device(xy, bbox[0], bbox[1])
torch.cuda.synchronize()
tg = (time.perf_counter() - t) / 5
prof = _lib.profile_report()
_lib.lib().ms_profile_enable(0)
t = time.perf_counter()
hf, ho = synth.periodic_delaunay(xy, bbox[0], bbox[1])
th = time.perf_counter() - t
print(f"points {xy.shape[0]}  faces {gf.shape[0]}  GPU {tg * 1e3:.2f} ms (host arrays in/out)  host scipy {th * 1e3:.0f} ms  x{th / tg:.0f}")
print({k: round(v["ms"] / 5, 3) for k, v in prof.items() if "delaunay" in k or k == "knn"})
