"""The per-frame part of the reference's examples/ex_3lipid.py (lines 142-182) on memsurfer_b200.

    python examples/ex_3lipid.py [--lipids 20000] [--frames 2] [--out PREFIX] [--no-write]

The reference script reads a Martini DPPC/DIPC/CHOL trajectory with MDAnalysis and finds the leaflets with
LeafletFinder; that input is not part of the reference tree (BASELINE.json configs[1]), so the frames here are
the synthetic 40/30/30 bilayer of memsurfer_b200.synth with the undulation phases advanced per frame.  Everything
after the selection is the script's own sequence: Membrane.compute for both leaflets, the total and the per-type
densities at sigma 10..40, (additionally) properties and the leaflet thickness, write_all.  Needs a B200.
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

LOGGER = logging.getLogger('ex_3lipid')
LIPIDS = np.array(['DPPC', 'DIPC', 'CHOL'])


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('--lipids', type=int, default=20000, help='lipids per frame (both leaflets)')
    ap.add_argument('--frames', type=int, default=2)
    ap.add_argument('--out', default='3lipid')
    ap.add_argument('--no-write', action='store_true')
    args = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format='%(asctime)s %(name)s %(levelname)s: %(message)s')
    periodic = True
    results = []
    for frame in range(args.frames):
        timer = Timer()
        leaflets = []
        for li, z0 in enumerate((20.0, -20.0)):
            pts, lab, bbox = synth.make_leaflet(args.lipids // 2, seed=1234 + 2 * frame + li, z0=z0, ntypes=3,
                                                fractions=[0.4, 0.3, 0.3], phase_shift=0.05 * frame)
            leaflets.append((pts, LIPIDS[lab], bbox))
        LOGGER.info(f'Frame: {frame:5d}: {len(leaflets[0][0])} lipids in upper and {len(leaflets[1][0])} in lower leaflets')

        mt = memsurfer.Membrane.compute(leaflets[0][0], leaflets[0][1], leaflets[0][2], periodic)
        mb = memsurfer.Membrane.compute(leaflets[1][0], leaflets[1][1], leaflets[1][2], periodic)
        for m in (mt, mb):
            m.compute_properties('smooth')
        memsurfer.Membrane.compute_thickness(mt, mb, 'smooth')

        sigmas = [10, 20, 30, 40]
        get_nlipdis = True
        memsurfer.Membrane.compute_densities([mt, mb], [2], sigmas, get_nlipdis, 'all')
        for l in LIPIDS:
            memsurfer.Membrane.compute_densities([mt, mb], [2], sigmas, get_nlipdis, l)

        if not args.no_write:
            mt.write_all(f'{args.out}_f{frame}-top', {'frame': frame, 'time': float(frame)})
            mb.write_all(f'{args.out}_f{frame}-bot', {'frame': frame, 'time': float(frame)})
        timer.end()
        LOGGER.info(f'frame {frame} done in {timer}')
        results.append((mt, mb))
    return results


if __name__ == '__main__':
    main()
