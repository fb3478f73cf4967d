"""The reference's examples/ex_simple.py flow on memsurfer_b200 (BASELINE.json configs[0]).

    python examples/ex_simple.py [mesh.off] [--out PREFIX]

Without an OFF file the vertices of the reference's own input (examples/data/noisy.off, 10 201 points; kept
as tests/golden/noisy_off_golden.npz) are used.  Same calls in the same order as examples/ex_simple.py:66-125;
the labels are seeded here (the script's np.random.randint is not).  Needs a B200: there is no CPU fallback.
"""
import argparse
import logging
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import memsurfer_b200 as memsurfer                                     # noqa: E402
from memsurfer_b200 import synth                                       # noqa: E402
from memsurfer_b200.utils import Timer                                 # noqa: E402

LOGGER = logging.getLogger('ex_simple')


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('filename', nargs='?', default=None)
    ap.add_argument('--out', default=None, help='prefix of the .vtp files (default: next to the input)')
    args = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format='%(asctime)s %(name)s %(levelname)s: %(message)s')

    if args.filename is not None:
        verts, _ = synth.read_off(args.filename)                       # faces are discarded (ex_simple.py:66)
        outprefix = args.out or args.filename[:args.filename.rfind('.')]
    else:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        verts = np.load(os.path.join(root, 'tests', 'golden', 'noisy_off_golden.npz'))['verts']
        outprefix = args.out or 'noisy'
    verts = np.array(verts)
    LOGGER.info(f'input has {verts.shape} vertices')

    periodic = True
    box = np.array([-1.01, -1.01, -1.01, 1.01, 1.01, 1.01]).reshape(2, 3)
    labels = np.random.default_rng(0).integers(0, 5, size=verts.shape[0])
    labels = np.array([str(l) for l in labels])                        # labels can be any datatype

    timer = Timer()
    knbrs = 18
    m = memsurfer.Membrane(verts, periodic=periodic, bbox=box, labels=labels)
    m.compute_pnormals(knbrs)
    m.compute_approx_surface()
    m.compute_membrane_surface()
    m.compute_properties('exact')
    m.compute_properties('smooth')

    sigmas = [0.2]
    get_nlipdis = True
    memsurfer.Membrane.compute_densities([m], [2], sigmas, get_nlipdis, 'all')
    memsurfer.Membrane.compute_densities([m], [2], sigmas, get_nlipdis, '1')

    m.write_all(outprefix, {'frame': 0, 'time': 0.0})
    timer.end()
    LOGGER.info(f'done in {timer}; wrote {outprefix}_*.vtp')
    return m, outprefix


if __name__ == '__main__':
    main()
