"""GPU: fields written ahead of their request (Membrane.EMIT_LABEL_FIELDS -> ms_kde_labels_multi_fields) are the same
numbers as the fields computed on request, and a request that does not match the emitted field (other normalisation,
other selection under the same name) is computed afresh."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIGMAS = [10.0, 25.0, 40.0]


def _densities(pts, labels, bbox, emit, order, get_nlipdis=True):
    from memsurfer_b200 import Membrane, TriMesh
    old = Membrane.EMIT_LABEL_FIELDS
    Membrane.EMIT_LABEL_FIELDS = emit
    try:
        m = Membrane(pts, labels=labels, periodic=True, bbox=bbox)
        mesh = TriMesh(pts, periodic=True)
        mesh.set_bbox(bbox[0], bbox[1])
        m.memb_smooth = mesh
        for l in order:
            Membrane.compute_densities([m], [2], SIGMAS, get_nlipdis if not isinstance(get_nlipdis, dict) else get_nlipdis[l], l)
        return {k: np.array(v) for k, v in m.attributes.items()}, {k: np.array(mesh.attributes_f64[k]) for k in m.attributes.keys()}
    finally:
        Membrane.EMIT_LABEL_FIELDS = old


def test_emitted_fields_equal_fields_on_request():
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(6400, seed=5, ntypes=3, fractions=[0.5, 0.3, 0.2])
    order = ['all', 0, 1, 2]
    a32, a64 = _densities(pts, labels, bbox, True, order)
    b32, b64 = _densities(pts, labels, bbox, False, order)
    assert sorted(a32) == sorted(b32) and len(a32) == len(order) * len(SIGMAS)
    for k in b32:
        assert np.array_equal(a32[k], b32[k]), k
        assert np.array_equal(a64[k], b64[k]), k
    # only what was asked for shows up in Membrane.attributes
    c32, _ = _densities(pts, labels, bbox, True, ['all', 1])
    assert sorted(c32) == sorted(k for k in b32 if '_all_' in k or '_1_' in k)
    for k in c32:
        assert np.array_equal(c32[k], b32[k]), k


def test_request_that_differs_from_the_emitted_field_is_recomputed():
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(6400, seed=6, ntypes=3, fractions=[0.5, 0.3, 0.2])
    # 'all' with get_nlipdis, then type 1 WITHOUT: the emitted 'density_type2_1_*' (normalised to counts) must not be used
    flags = {'all': True, 1: False}
    a32, _ = _densities(pts, labels, bbox, True, ['all', 1], get_nlipdis=flags)
    b32, _ = _densities(pts, labels, bbox, False, ['all', 1], get_nlipdis=flags)
    for k in b32:
        assert np.array_equal(a32[k], b32[k]), k
    n1 = float((labels == 1).sum())
    k1 = 'density_type2_1_k10.0'
    assert abs(float(a32[k1].astype(np.float64).sum()) - n1) > 1.0          # not normalised to the lipid count
    # a label that does not occur: the reference's empty id list = all vertices, under the name of the absent label
    c32, _ = _densities(pts, labels, bbox, True, ['all', 7])
    assert np.array_equal(c32['density_type2_7_k25.0'], c32['density_type2_all_k25.0'])
