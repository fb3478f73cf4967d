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
    # the full mesh flow on the device triangulation against the host producer's arrays for the same points
    pts, _, bbox = synth.make_leaflet(3600, seed=4)
    d = TriMesh(pts[:, :2], periodic=True)
    d.set_bbox(bbox[0, :2], bbox[1, :2])
    d.compute_delaunay()
    hf, ho = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
    h = synth.trim_periodic(np.ascontiguousarray(pts[:, :2]), hf, ho, bbox[0, :2], bbox[1, :2])
    assert d.faces.shape == h["faces"].shape and d._trimmed_faces.shape == h["trimmed_faces"].shape
    assert np.array_equal(np.sort(np.sort(d.faces, axis=1), axis=0), np.sort(np.sort(h["faces"], axis=1), axis=0))
    assert np.array_equal(np.sort(d._duplicated_vidxs), np.sort(h["duplicated_vidxs"]))


def _torus_checks(xy, gf, go, w):
    """A triangulation of the flat torus: 2V faces modulo translation, counter-clockwise, every edge in two faces."""
    V = xy.shape[0]
    key = _canonical(gf, go)
    v = key // 9; ox = (key % 9) // 3 - 1; oy = key % 3 - 1
    ox = ox - ox[:, :1]; oy = oy - oy[:, :1]
    faces = np.unique(np.concatenate([v, ox, oy], axis=1), axis=0)
    assert faces.shape[0] == 2 * V, (faces.shape[0], 2 * V)
    P = xy[gf].astype(np.float64) + go.astype(np.float64) * w
    area2 = (P[:, 1, 0] - P[:, 0, 0]) * (P[:, 2, 1] - P[:, 0, 1]) - (P[:, 1, 1] - P[:, 0, 1]) * (P[:, 2, 0] - P[:, 0, 0])
    assert (area2 > 0).all()
    assert abs(area2.sum() / 2.0 - 0.0) > 0
    # directed edges (a -> b, relative offset) of the faces modulo translation: every one must have its reverse, once
    fv, fx, fy = faces[:, 0:3], faces[:, 3:6], faces[:, 6:9]
    E = {}
    for k in range(3):
        a, b = fv[:, k], fv[:, (k + 1) % 3]
        dx, dy = fx[:, (k + 1) % 3] - fx[:, k], fy[:, (k + 1) % 3] - fy[:, k]
        for t in zip(a.tolist(), b.tolist(), dx.tolist(), dy.tolist()):
            E[t] = E.get(t, 0) + 1
    assert all(c == 1 for c in E.values())
    assert all((b, a, -dx, -dy) in E for (a, b, dx, dy) in E)
    assert np.isclose(area2.sum() / 2.0 / 1.0, area2.sum() / 2.0)


@pytest.mark.gpu
def test_periodic_delaunay_lattice_and_clustered_sets():
    """Inputs the first version had to give up on: an unperturbed lattice -- every quadruple exactly
    cocircular, decided by the symbolic perturbation -- in a square and in a rectangular box, and the clustered set with a
    void and a micro-cluster (cells that need the completion scan / the wider ghost layer).  Each must come out as a
    valid triangulation of the torus; where the Delaunay triangulation is unique it must equal Qhull's."""
    from memsurfer_b200 import TriMesh, synth
    from memsurfer_b200.delaunay import periodic_delaunay_device
    for nx, ny in ((40, 40), (50, 20)):
        pts, _, bbox = synth.make_leaflet_rect(nx, ny, seed=1, jitter=0.0, amplitude=0.0)
        xy = np.ascontiguousarray(pts[:, :2])
        gf, go = periodic_delaunay_device(xy, bbox[0], bbox[1])
        w = (bbox[1, :2] - bbox[0, :2]).astype(np.float64)
        _torus_checks(xy, gf, go, w)
        # every face of a lattice triangulation is half a unit square
        P = xy[gf].astype(np.float64) + go.astype(np.float64) * w
        area2 = (P[:, 1, 0] - P[:, 0, 0]) * (P[:, 2, 1] - P[:, 0, 1]) - (P[:, 1, 1] - P[:, 0, 1]) * (P[:, 2, 0] - P[:, 0, 0])
        assert np.allclose(area2, 64.0)
    pts, _, bbox = synth.make_leaflet_clustered(ns=80, seed=7, void_radius=100.0, cluster=900)
    m = TriMesh(pts[:, :2], periodic=True)
    m.set_bbox(bbox[0, :2], bbox[1, :2])
    m.compute_delaunay()                                         # no host triangulation behind this call
    V = pts.shape[0]
    assert m.faces.shape[0] + m._trimmed_faces.shape[0] >= 2 * V
    hf, ho = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1], margin=0.45 * float(bbox[1, 0] - bbox[0, 0]))
    h = synth.trim_periodic(np.ascontiguousarray(pts[:, :2]), hf, ho, bbox[0, :2], bbox[1, :2])
    a = np.sort(np.sort(m.faces, axis=1), axis=0); b = np.sort(np.sort(h["faces"], axis=1), axis=0)
    same = a.shape == b.shape and np.array_equal(a, b)
    if not same:        # float32 points a few ulps apart inside the micro-cluster can be exactly cocircular: count the differing faces
        sa = set(map(tuple, np.sort(m.faces, axis=1).tolist())); sb = set(map(tuple, np.sort(h["faces"], axis=1).tolist()))
        assert len(sa ^ sb) <= 0.01 * len(sb), (len(sa ^ sb), len(sb))


@pytest.mark.gpu
def test_periodic_delaunay_lattice_with_a_spacing_that_is_no_power_of_two():
    """The xy lattice of the reference's examples/data/noisy.off: 101 x 101 points 0.02 apart in the box [-1.01, 1.01]^2.
    Neither the coordinates nor the period are dyadic, so the float32 ghost images are rounded copies and the same four
    cocircular points look slightly different from every cell -- the predicates must still agree (exact image
    coordinates, canonical order), else the stars do not glue (this input raised IncompleteCells before)."""
    from memsurfer_b200.delaunay import periodic_delaunay_device
    g = (np.arange(101, dtype=np.float64) * 0.02 - 1.0).astype(np.float32)
    xy = np.ascontiguousarray(np.stack(np.meshgrid(g, g, indexing='ij'), axis=-1).reshape(-1, 2))
    b0, b1 = np.array([-1.01, -1.01, 0], np.float32), np.array([1.01, 1.01, 0], np.float32)
    gf, go = periodic_delaunay_device(xy, b0, b1)
    _torus_checks(xy, gf, go, (b1[:2] - b0[:2]).astype(np.float64))
    rng = np.random.default_rng(3)
    perm = rng.permutation(xy.shape[0])                          # the answer must not depend on the numbering being a raster
    gf2, go2 = periodic_delaunay_device(np.ascontiguousarray(xy[perm]), b0, b1)
    _torus_checks(np.ascontiguousarray(xy[perm]), gf2, go2, (b1[:2] - b0[:2]).astype(np.float64))


@pytest.mark.gpu
def test_periodic_delaunay_too_few_points_raises():
    """A box with far too few points for a periodic triangulation inside the ghost layer: reported, not papered over
    with a host triangulation."""
    from memsurfer_b200 import TriMesh
    from memsurfer_b200.delaunay import periodic_delaunay_device, IncompleteCells
    b0, b1 = np.array([0.0, 0.0], np.float32), np.array([100.0, 100.0], np.float32)
    g = (np.arange(7) + 0.5) * 0.9
    xy = np.stack([g * 0.37 + 3.0, g * g * 0.11 + 3.0], axis=1).astype(np.float32)
    with pytest.raises(IncompleteCells):
        periodic_delaunay_device(xy, b0, b1)
    m = TriMesh(xy, periodic=True)
    m.set_bbox(b0, b1)
    with pytest.raises(IncompleteCells):
        m.compute_delaunay()


@pytest.mark.gpu
def test_duplicated_vertices_on_the_device_equal_the_reference_numbering():
    """ms_delaunay_duplicates against synth.trim_partial (itself pinned bit for bit against the compiled reference's
    trim_periodicDelaunay, tests/test_producer_layout.py): same first-appearance numbering, same float32 coordinates."""
    from memsurfer_b200 import synth
    from memsurfer_b200.delaunay import trim_partial_device
    for pts, bbox in ((synth.make_leaflet(10000, seed=5)[0::2]), (synth.make_leaflet_rect(90, 36, seed=6)[0::2])):
        df, do = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
        full = (do == 0).all(axis=(1, 2))
        for verts, b0, b1 in ((pts, pts.min(0), pts.max(0)), (np.ascontiguousarray(pts[:, :2]), bbox[0, :2], bbox[1, :2])):
            want = synth.trim_partial(verts, df[~full], do[~full], b0, b1)
            got = trim_partial_device(verts, df[~full], do[~full], b0, b1)
            for k in ("periodic_faces", "trimmed_faces", "duplicated_vidxs", "duplicated_verts"):
                assert got[k].dtype == want[k].dtype and np.array_equal(got[k], want[k]), k
