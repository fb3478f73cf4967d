#!/usr/bin/env python
"""cProfile of one end-to-end frame through the drop-in API (host numpy in, numpy out)."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                    # noqa: E402
from memsurfer_b200 import Membrane                             # noqa: E402

lipids = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
leaflets = bench.make_frame_inputs(lipids, 0)
dev = torch.device("cuda:0")


def step():
    ms = []
    for lf in leaflets:
        m = Membrane(lf["points"], labels=lf["labels"], periodic=True, bbox=lf["bbox"], device=dev)
        m.compute_pnormals(18)
        m.memb_smooth = bench.new_mesh(lf, dev)
        m.compute_properties('smooth')
        for l in ['all', 0, 1, 2]:
            Membrane.compute_densities([m], [2], bench.SIGMAS, True, l)
        ms.append(m)
    for m in ms:
        m.memb_smooth.attributes.resolve(); m.attributes.resolve()
    torch.cuda.synchronize()


step(); step()
pr = cProfile.Profile()
pr.enable(); step(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(28)
