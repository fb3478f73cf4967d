"""Periodic 2-D Delaunay on the GPU against the host triangulation
(scipy/Qhull over the tiled points, memsurfer_b200.synth.periodic_delaunay -- the stand-in for CGAL, which is
not in this image: parity with CGAL itself is unpinned).  The Delaunay triangulation of points in general
position is unique, so the face SETS must agree exactly."""
import numpy as np
import pytest


def _canonical(dfaces, doffs):
    """Every face as a row (v0,ox0,oy0, v1,..., v2,...) rotated so that its smallest (v,ox,oy) comes first
    (rotation keeps the winding); rows sorted."""
    key = dfaces.astype(np.int64) * 9 + (doffs[..., 0].astype(np.int64) + 1) * 3 + (doffs[..., 1].astype(np.int64) + 1)
    first = key.argmin(axis=1)
    idx = (first[:, None] + np.arange(3)[None, :]) % 3
    rows = np.take_along_axis(key, idx, axis=1)
    return rows[np.lexsort(rows.T[::-1])]


@pytest.mark.gpu
@pytest.mark.parametrize("n,seed", [(400, 1), (10000, 2), (250000, 3)])
def test_periodic_delaunay_gpu_vs_host(n, seed):
    from memsurfer_b200 import synth
    from memsurfer_b200.delaunay import periodic_delaunay_device
    pts, _, bbox = synth.make_leaflet(n, seed=seed)
    xy = pts[:, :2]
    gf, go = periodic_delaunay_device(xy, bbox[0], bbox[1])
    hf, ho = synth.periodic_delaunay(xy, bbox[0], bbox[1])
    assert gf.dtype == np.int64 and go.dtype == np.int8 and gf.shape[1:] == (3,) and go.shape[1:] == (3, 2)
    assert (go == 0).all(axis=2).any(axis=1).all()              # every face touches the original sheet
    a, b = _canonical(gf, go), _canonical(hf, ho)
    assert a.shape == b.shape and np.array_equal(a, b)
    # the device-side split of the trim step: same faces, same order
    ff, pf, po = periodic_delaunay_device(xy, bbox[0], bbox[1], split=True)
    full = (go == 0).all(axis=(1, 2))
    assert ff.dtype == np.uint32 and np.array_equal(ff, gf[full].astype(np.uint32))
    assert np.array_equal(pf, gf[~full]) and np.array_equal(po, go[~full])
    # counter-clockwise in the plane
    w = (bbox[1, :2] - bbox[0, :2]).astype(np.float64)
    P = xy[gf].astype(np.float64) + go.astype(np.float64) * w
    area2 = (P[:, 1, 0] - P[:, 0, 0]) * (P[:, 2, 1] - P[:, 0, 1]) - (P[:, 1, 1] - P[:, 0, 1]) * (P[:, 2, 0] - P[:, 0, 0])
    assert (area2 > 0).all()


@pytest.mark.gpu
def test_periodic_delaunay_random_points_and_mesh_flow():
    """Uniform random points (uneven cells, a few need the longer neighbour list) and the full mesh flow:
    compute_delaunay -> trim -> properties on the device triangulation equal those on the host one."""
    from memsurfer_b200 import TriMesh, synth
    from memsurfer_b200.delaunay import periodic_delaunay_device
    rng = np.random.default_rng(7)
    L = 500.0
    xy = rng.uniform(0.0, L, (20000, 2)).astype(np.float32)
    b0, b1 = np.array([0.0, 0.0], np.float32), np.array([L, L], np.float32)
    gf, go = periodic_delaunay_device(xy, b0, b1)
    hf, ho = synth.periodic_delaunay(xy, b0, b1, margin=0.3 * L)
    assert np.array_equal(_canonical(gf, go), _canonical(hf, ho))
    # a torus triangulation has exactly 2V faces: count the faces modulo translation
    key = _canonical(gf, go)
    v = key // 9; ox = (key % 9) // 3 - 1; oy = key % 3 - 1
    ox -= ox[:, :1]; oy -= oy[:, :1]                              # translate so that the first vertex is in the sheet
    mod = np.unique(np.concatenate([v, ox, oy], axis=1), axis=0)
    assert mod.shape[0] == 2 * xy.shape[0]
    import os
    pts, _, bbox = synth.make_leaflet(3600, seed=4)
    meshes = {}
    for mode in ("device", "host"):
        os.environ["MS_DELAUNAY"] = mode
        try:
            m = TriMesh(pts[:, :2], periodic=True)
            m.set_bbox(bbox[0, :2], bbox[1, :2])
            m.compute_delaunay()
            meshes[mode] = m
        finally:
            os.environ.pop("MS_DELAUNAY", None)
    d, h = meshes["device"], meshes["host"]
    assert d.faces.shape == h.faces.shape and d._trimmed_faces.shape == h._trimmed_faces.shape
    assert np.array_equal(np.sort(np.sort(d.faces, axis=1), axis=0), np.sort(np.sort(h.faces, axis=1), axis=0))
    assert np.array_equal(np.sort(d._duplicated_vidxs), np.sort(h._duplicated_vidxs))
    assert np.allclose(np.sort(d.compute_pointareas()), np.sort(h.compute_pointareas()), rtol=0, atol=0)


@pytest.mark.gpu
def test_periodic_delaunay_degenerate_inputs_are_reported():
    """An unperturbed lattice (every quadruple cocircular) and a box with far too few points: the device
    producer reports them and TriMesh.compute_delaunay falls back to the host triangulation."""
    from memsurfer_b200 import TriMesh
    from memsurfer_b200.delaunay import periodic_delaunay_device, IncompleteCells
    g = (np.arange(40) + 0.5) * 2.5
    X, Y = np.meshgrid(g, g, indexing="xy")
    xy = np.stack([X.ravel(), Y.ravel()], axis=1).astype(np.float32)
    b0, b1 = np.array([0.0, 0.0], np.float32), np.array([100.0, 100.0], np.float32)
    with pytest.raises(IncompleteCells):
        periodic_delaunay_device(xy, b0, b1)
    with pytest.raises(IncompleteCells):
        periodic_delaunay_device(xy[:7] * 0.37 + 3.0, b0, b1)
    m = TriMesh(xy, periodic=True)
    m.set_bbox(b0, b1)
    m.compute_delaunay()                                         # host fallback
    assert m.faces.shape[0] + m._periodic_faces.shape[0] >= 2 * xy.shape[0]
