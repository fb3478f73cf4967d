"""CPU: the producer layout every periodic test and the bench consume -- `synth.trim_periodic` / `trim_partial` (and the
wrap of `delaunay.wrap_into_box`) -- against the COMPILED reference's own TriMesh::trim_periodicDelaunay
(memsurfer/src/TriMesh.cpp:493-575) and TriMesh::wrap_vertices (:462-491), bit for bit (oracle/_ref, built in place from
/root/reference with the class opened by oracle/ref_shim.cpp)."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def ref(oracle):
    if oracle.ref() is None or not hasattr(oracle.ref(), "msref_trim_periodic"):
        pytest.skip("oracle/_ref was not built (reference tree absent)")
    return oracle


def _cases():
    from memsurfer_b200 import synth
    out = []
    pts, _, bbox = synth.make_leaflet(2500, seed=3)
    out.append(("square, own bbox = vertex extent", pts, bbox, pts.min(0), pts.max(0)))
    out.append(("square, own bbox = periodic box", pts, bbox, bbox[0], bbox[1]))
    pts, _, bbox = synth.make_leaflet_rect(60, 24, seed=4)
    out.append(("rectangular box 2.5:1", pts, bbox, pts.min(0), pts.max(0)))
    pts, _, bbox = synth.make_leaflet_clustered(ns=40, seed=5, void_radius=50.0, cluster=60)
    out.append(("clustered + void", pts, bbox, pts.min(0), pts.max(0)))
    rng = np.random.default_rng(6)
    pts = np.concatenate([rng.uniform(0, 300, (1500, 2)), rng.uniform(-5, 5, (1500, 1))], axis=1).astype(np.float32)
    bbox = np.array([[0, 0, -10], [300, 300, 10]], np.float32)
    out.append(("uniform random", pts, bbox, pts.min(0), pts.max(0)))
    return out


@pytest.mark.parametrize("case", range(5))
def test_trim_periodic_equals_compiled_reference(ref, case):
    from memsurfer_b200 import synth
    name, pts, bbox, lo, hi = _cases()[case]
    df, do = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
    assert (do != 0).any(), name
    for verts, b0, b1 in ((pts, lo, hi), (np.ascontiguousarray(pts[:, :2]), bbox[0, :2], bbox[1, :2])):      # 3-D mesh and the planar mesh
        mine = synth.trim_periodic(verts, df, do, b0, b1)
        want = ref.ref_trim_periodic(verts, df, do, b0, b1)
        for k in ("faces", "periodic_faces", "trimmed_faces", "duplicated_vidxs", "duplicated_verts"):
            assert mine[k].dtype == want[k].dtype and mine[k].shape == want[k].shape, (name, k, mine[k].shape, want[k].shape)
            assert np.array_equal(mine[k], want[k]), (name, k)
        # the device producer hands over the faces already split into inside / crossing (ms_delaunay_split):
        # the crossing part through trim_partial must give the same arrays
        full = (do == 0).all(axis=(1, 2))
        part = synth.trim_partial(verts, df[~full], do[~full], b0, b1)
        for k in ("periodic_faces", "trimmed_faces", "duplicated_vidxs", "duplicated_verts"):
            assert np.array_equal(part[k], want[k]), (name, k)
        assert np.array_equal(df[full].astype(np.uint32), want["faces"])


def test_wrap_into_box_equals_compiled_reference(ref):
    from memsurfer_b200.delaunay import wrap_into_box
    rng = np.random.default_rng(1)
    b0, b1 = np.array([10.0, -5.0], np.float32), np.array([110.0, 35.0], np.float32)
    xy = rng.uniform(-80, 200, (4000, 2)).astype(np.float32)          # up to one period outside
    xy[:4] = [[10.0, -5.0], [110.0, 35.0], [10.0, 35.0], [110.0, -5.0]]     # exactly on the faces
    inside = (xy[:, 0] > b0[0] - 100) & (xy[:, 0] < b1[0] + 100) & (xy[:, 1] > b0[1] - 40) & (xy[:, 1] < b1[1] + 40)
    xy = np.ascontiguousarray(xy[inside])
    got = wrap_into_box(xy, b0, b1, out=np.empty((xy.shape[0], 3), np.float32))[:, :2]
    want = ref.ref_wrap_vertices(xy, b0, b1, 2)
    assert np.array_equal(got, want)
    assert (got[:, 0] >= b0[0]).all() and (got[:, 0] < b1[0]).all()
