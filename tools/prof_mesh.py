#!/usr/bin/env python
"""One BASELINE configs[2] leaflet through topology + mesh properties (for ncu)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from memsurfer_b200.frame import LeafletFrame
dev = torch.device("cuda:0")
lf = bench.make_frame_inputs(1000000, 0, device=dev)[0]
fr = LeafletFrame(lf["points"], lf["labels"], lf["bbox"], bench.new_mesh(lf, dev), device=dev)
for _ in range(3):
    fr.run_properties()
torch.cuda.synchronize()
print("ok", fr.nv, fr.nf)
