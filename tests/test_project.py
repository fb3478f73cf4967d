"""Closest point on a mesh / leaflet thickness:
oracle anchors on CPU, CUDA path vs the oracle on the GPU (bit-exact: same formulas, no FMA, ties by face id)."""
import numpy as np
import pytest


def _plane_mesh(n=12, z=0.0):
    g = np.linspace(0.0, 10.0, n)
    X, Y = np.meshgrid(g, g, indexing="ij")
    verts = np.stack([X.ravel(), Y.ravel(), np.full(X.size, z)], axis=1).astype(np.float32)
    idx = np.arange(n * n).reshape(n, n)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel(), idx[:-1, 1:].ravel()
    faces = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)]).astype(np.uint32)
    return verts, faces


def test_oracle_closest_point_anchors(oracle):
    verts, faces = _plane_mesh()
    rng = np.random.default_rng(3)
    q = np.concatenate([rng.uniform(0.5, 9.5, (200, 2)), rng.uniform(-3, 3, (200, 1))], axis=1).astype(np.float32)
    fid, bary, dist = oracle.closest_point(verts, faces, q)
    # above/below a plane: distance |z|, foot point = (x, y, 0), barycentric coordinates a partition of unity
    assert np.allclose(dist, np.abs(q[:, 2].astype(np.float64)), rtol=0, atol=1e-12)
    foot = (bary[:, :, None] * verts[faces[fid]].astype(np.float64)).sum(axis=1)
    assert np.allclose(foot[:, :2], q[:, :2], atol=1e-5) and np.allclose(foot[:, 2], 0.0, atol=1e-12)
    assert (bary >= -1e-9).all() and np.allclose(bary.sum(axis=1), 1.0, atol=1e-12)
    # outside the sheet: the closest point is on the border (edge or corner)
    q2 = np.array([[-3.0, 5.0, 4.0], [-3.0, -4.0, 0.0], [13.0, 12.0, 1.0]], np.float32)
    _, _, d2 = oracle.closest_point(verts, faces, q2)
    assert np.allclose(d2, [5.0, 5.0, np.sqrt(9.0 + 4.0 + 1.0)], atol=1e-12)
    # against dense sampling of a random triangle soup: never farther than any sampled surface point
    tv = rng.normal(size=(30, 3)).astype(np.float32)
    tf = rng.integers(0, 30, (40, 3)).astype(np.uint32)
    tf = tf[(tf[:, 0] != tf[:, 1]) & (tf[:, 1] != tf[:, 2]) & (tf[:, 0] != tf[:, 2])]
    qq = rng.normal(size=(50, 3)).astype(np.float32)
    _, _, dd = oracle.closest_point(tv, tf, qq)
    w = rng.dirichlet([1, 1, 1], 400)
    samples = np.einsum("sk,fkd->fsd", w, tv[tf].astype(np.float64)).reshape(-1, 3)
    dmin = np.sqrt(((qq[:, None, :].astype(np.float64) - samples[None]) ** 2).sum(-1)).min(axis=1)
    assert (dd <= dmin + 1e-12).all() and (dd >= 0.6 * dmin - 1e-9).all()


@pytest.mark.gpu
def test_closest_point_gpu_vs_oracle(oracle):
    from memsurfer_b200 import TriMesh, synth
    rng = np.random.default_rng(11)
    pts, _, bbox = synth.make_leaflet(2500, seed=21, z0=-20.0)
    mesh = TriMesh(pts, faces=synth.planar_delaunay(pts[:, :2]))
    lo, hi = pts.min(0), pts.max(0)
    q = np.concatenate([rng.uniform(lo[:2] - 30, hi[:2] + 30, (3000, 2)), rng.uniform(-60, 60, (3000, 1))], axis=1).astype(np.float32)
    q = np.concatenate([q, pts[:200]])                     # queries on the surface itself (distance 0, shared vertices)
    fid, bary = mesh.project_on_surface(q)
    ofid, obary, odist = oracle.closest_point(mesh.vertices, mesh.faces, q)
    assert np.array_equal(fid, ofid)
    assert np.array_equal(bary, obary)
    _, _, d64, d32 = mesh.closest_point_device(__import__("torch").from_numpy(q).to(mesh.device))
    assert np.array_equal(d64.cpu().numpy(), odist)
    assert np.array_equal(d32.cpu().numpy(), odist.astype(np.float32))
    sp, pp = mesh.project_on_surface_and_plane(q[:500])
    assert sp.dtype == np.float32 and sp.shape == (500, 3) and pp.shape == (500, 2)
    foot = (obary[:500, :, None] * mesh.vertices[mesh.faces[ofid[:500]]].astype(np.float64)).sum(axis=1)
    assert np.allclose(sp, foot, atol=1e-4) and np.allclose(pp, foot[:, :2], atol=1e-4)


@pytest.mark.gpu
def test_membrane_thickness(oracle):
    from memsurfer_b200 import Membrane, synth
    ms = []
    for z0, seed in ((20.0, 31), (-20.0, 32)):
        pts, labels, bbox = synth.make_leaflet(3600, seed=seed, z0=z0)
        m = Membrane(pts, labels=labels, periodic=True, bbox=bbox)
        m.compute_membrane_surface()
        ms.append(m)
    Membrane.compute_thickness(ms[0], ms[1], 'smooth')
    for a, b in ((ms[0], ms[1]), (ms[1], ms[0])):
        t = a.properties['thickness']
        assert t.dtype == np.float32 and t.shape == (a.memb_smooth.vertices.shape[0],)
        _, _, od = oracle.closest_point(b.memb_smooth.vertices, b.memb_smooth.faces, a.memb_smooth.vertices)
        assert np.array_equal(a.memb_smooth.attributes_f64['distance_to_surface'], od)
        assert np.array_equal(t, od.astype(np.float32))
        assert (t > 20.0).all() and (t < 80.0).all()


@pytest.mark.gpu
def test_thickness_full_size_properties():
    """BASELINE configs[2] size: closest-point distance between the two 499 849-vertex leaflets.  Size-independent
    checks: never above the vertical gap to the other surface's closest vertex, never below the z gap of the slabs."""
    import torch
    from memsurfer_b200 import TriMesh, synth
    meshes = []
    for z0, seed in ((20.0, 1234), (-20.0, 1235)):
        pts, _, bbox = synth.make_leaflet(499849, seed=seed, z0=z0)
        meshes.append(TriMesh(pts, faces=synth.planar_delaunay(pts[:, :2])))
    a, b = meshes
    d = a.compute_distance_to_surface(b)
    assert d.shape == (a.vertices.shape[0],) and np.isfinite(d).all()
    gap = a.vertices[:, 2].min() - b.vertices[:, 2].max()
    assert (d >= gap - 1e-3).all()
    # upper bound: distance to the other mesh's nearest vertex in xy on a subsample (a vertex is a surface point)
    from scipy.spatial import cKDTree
    sub = np.random.default_rng(0).choice(a.vertices.shape[0], 20000, replace=False)
    _, j = cKDTree(b.vertices[:, :2]).query(a.vertices[sub, :2])
    ub = np.linalg.norm(a.vertices[sub].astype(np.float64) - b.vertices[j].astype(np.float64), axis=1)
    assert (d[sub] <= ub + 1e-4).all()
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_overlapping_points_vs_pdist():
    """the grid-based guard finds exactly the pairs scipy's pdist finds (memsurfer/utils.py:112-113)."""
    import scipy.spatial.distance
    import torch
    from memsurfer_b200 import synth, Membrane
    from memsurfer_b200.utils import test_overlapping_points, overlapping_pairs_device
    pts, labels, bbox = synth.make_leaflet(2500, seed=8)
    assert test_overlapping_points(pts) is True
    bad = pts.copy()
    bad[17] = bad[900]; bad[1200] = bad[900]                      # a triple
    bad[33] = bad[34] + np.array([3e-6, 0, 0], np.float32)       # closer than 1e-5
    bad[50] = bad[51] + np.array([0, 5e-3, 0], np.float32)       # close but not overlapping
    pd = scipy.spatial.distance.pdist(bad)
    n = bad.shape[0]
    iu = np.triu_indices(n, 1)
    want = np.stack([iu[0][pd < 1e-5], iu[1][pd < 1e-5]], axis=1)
    got, dmin = overlapping_pairs_device(torch.from_numpy(bad).cuda(), 1e-5)
    assert np.array_equal(got, want.astype(np.int32))
    assert np.isclose(dmin, pd[pd > 1e-5].min(), rtol=0, atol=0)
    assert test_overlapping_points(bad) is False
    fixed = bad.copy()
    np.random.seed(0)          # the reference's fix draws its moves from np.random (utils.py:147-148): ~1 % of draws move < 1e-5
    test_overlapping_points(fixed, fix=True)
    assert test_overlapping_points(fixed) is True                 # every coincident pair was separated
    assert (fixed != bad).any(axis=1).sum() <= 3
    # 2-D input (ppoints) and the Membrane constructor hook
    assert test_overlapping_points(pts[:, :2].copy()) is True
    mb = Membrane(bad, labels=labels, periodic=True, bbox=bbox)   # warns, does not raise (membrane.py:84)
    assert mb.check_overlaps() is False
    assert Membrane(pts, labels=labels, periodic=True, bbox=bbox).check_overlaps() is True
