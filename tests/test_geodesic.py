"""Geodesic (type 1) KDE: the fp64 restatement against the COMPILED reference
(TriMesh::kde type 1 = compute_geodesics_fw + kde_2m, float32) on CPU; the CUDA path against the restatement
on the GPU (1e-9 relative: same relaxations, fp64, different association of the path sums)."""
import numpy as np
import pytest


def _mesh(n, seed):
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(n, seed=seed)
    return pts, synth.planar_delaunay(pts[:, :2]).astype(np.uint32), labels, bbox


def test_oracle_geodesic_kde_vs_reference(oracle):
    if oracle.ref() is None or not hasattr(oracle.ref(), "msref_kde_geodesic"):
        pytest.skip("compiled reference not built")
    pts, faces, labels, bbox = _mesh(324, 7)
    ids = np.flatnonzero(labels == 1)
    bb = np.array([bbox[0, 0], bbox[0, 1], 0, bbox[1, 0], bbox[1, 1], 0], np.float32)
    for sel, counts, box in [([], True, None), (ids, True, None), ([], False, bb), (ids, False, bb)]:
        d = oracle.kde_geodesic(pts, faces, sel, 12.0, counts, box)
        r = oracle.ref_kde_geodesic(pts, faces, sel, 12.0, counts, box)
        assert np.abs(d - r).max() <= 5e-6 * np.abs(r).max()
    # anchors: graph distances are symmetric, zero on the diagonal, never shorter than the straight line, and
    # equal to the edge length for mesh edges that are shortest paths themselves
    D = oracle.geodesics(pts, faces)
    assert np.array_equal(D, D.T) and (np.diag(D) == 0).all()
    E = np.sqrt(((pts[:, None, :].astype(np.float64) - pts[None, :, :]) ** 2).sum(-1))
    assert (D >= E - 1e-9).all()
    a, b = faces[:, 0], faces[:, 1]
    assert (D[a, b] <= E[a, b] + 1e-12).all()


@pytest.mark.gpu
def test_geodesic_kde_gpu_vs_oracle(oracle):
    import torch
    from memsurfer_b200 import TriMesh
    pts, faces, labels, bbox = _mesh(1225, 11)                # 1225 vertices: pitch 1248, 39 pivots
    mesh = TriMesh(pts, faces=faces)
    V = pts.shape[0]
    D = mesh.geodesics_device()[:V, :V].cpu().numpy()
    Do = oracle.geodesics(pts, faces)
    assert np.allclose(D, Do, rtol=1e-12, atol=0)
    ids = np.flatnonzero(labels == 2)
    for name, sel, counts in [("g_all", [], True), ("g_sel", ids, True), ("g_raw", ids, False)]:
        dens = mesh.compute_density(1, 12.0, name, counts, np.asarray(sel))
        ref = oracle.kde_geodesic(pts, faces, sel, 12.0, counts)
        assert dens.dtype == np.float32 and dens.shape == (V,)
        d64 = mesh.attributes_f64[name]
        assert np.allclose(d64, ref, rtol=1e-9, atol=1e-12)
        assert np.allclose(dens, ref.astype(np.float32), rtol=1e-6, atol=0)
    assert mesh.compute_density(1, 12.0, "g_all", True, np.empty([0])) is mesh._fields["g_all"]     # name-keyed cache


@pytest.mark.gpu
def test_geodesic_kde_periodic_through_membrane(oracle):
    """Periodic mesh (faces ++ periodic faces, wrapped edge lengths) through Membrane.compute_density(type 1)."""
    from memsurfer_b200 import Membrane, synth
    pts, labels, bbox = synth.make_leaflet(900, seed=5)
    m = Membrane(pts, labels=labels, periodic=True, bbox=bbox)
    m.compute_membrane_surface()
    mesh = m.memb_smooth
    m.compute_density(1, 15.0, "geo_1", True, [1])
    faces = np.concatenate([mesh.faces, mesh._periodic_faces]).astype(np.uint32)
    ref = oracle.kde_geodesic(mesh.vertices, faces, np.flatnonzero(labels == 1), 15.0, True, mesh.bboxl)
    assert np.allclose(mesh.attributes_f64["geo_1"], ref, rtol=1e-9, atol=1e-12)
    assert np.allclose(m.attributes["geo_1"], ref.astype(np.float32), rtol=1e-6, atol=0)
