"""GPU, >= 2 devices (run with `gpurun --gpus 2`): the C ABI keeps no per-process state -- two Membranes on cuda:0 and
cuda:1 driven from ONE process, first one after the other and then from two host threads at once (ctypes drops the
GIL), must both match the oracle (re-entrant per stream, no global mutable state)."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(dev, pts, labels, bbox, out):
    import torch
    from memsurfer_b200 import Membrane
    try:
        with torch.cuda.device(dev):
            m = Membrane(pts, labels=labels, periodic=True, bbox=bbox, device=dev)
            m.compute_pnormals(18, wait=False)
            m.compute_membrane_surface()
            m.compute_properties('smooth')
            for l in ['all', 0, 1, 2]:
                Membrane.compute_densities([m], [2], [10.0, 25.0], True, l)
            out[dev] = dict(pn=m.pnormals_f64.copy(), H=m.memb_smooth.attributes['mean_curv'].copy(),
                            d={k: m.memb_smooth.attributes_f64[k].copy() for k in list(m.attributes.keys())}, mesh=m.memb_smooth)
    except Exception as e:          # surfaced by the caller
        out[dev] = e


def test_two_devices_one_process_and_two_threads(oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(6400, seed=21)
    want_pn = oracle.point_normals(pts, bbox, 0.2, 18, +1)
    for mode in ("sequential", "threads"):
        out = {}
        if mode == "sequential":
            for dev in ("cuda:1", "cuda:0"):          # the second device first: nothing may have been set up by cuda:0
                _run(dev, pts, labels, bbox, out)
        else:
            th = [threading.Thread(target=_run, args=(dev, pts, labels, bbox, out)) for dev in ("cuda:0", "cuda:1")]
            [t.start() for t in th]; [t.join() for t in th]
        for dev in ("cuda:0", "cuda:1"):
            r = out[dev]
            assert not isinstance(r, Exception), (mode, dev, r)
            assert np.array_equal(r['pn'], want_pn), (mode, dev)
            bb = r['mesh'].bboxl
            for name, got in r['d'].items():
                _, _, l, s = name.split('_')
                ids = np.zeros(0, np.int32) if l == 'all' else np.where(labels == int(l))[0].astype(np.int32)
                want = oracle.kde(pts, ids, 2, float(s[1:]), True, bb)
                assert (np.abs(got - want) <= 1e-12 + 1e-9 * np.abs(want)).all(), (mode, dev, name)
        assert np.array_equal(out["cuda:0"]['H'], out["cuda:1"]['H'])
