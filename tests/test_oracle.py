"""CPU: the fp64 oracle against (a) golden vectors made by the reference's own compiled C++,
(b) the compiled reference itself when oracle/_ref is present, (c) analytic anchors."""
import numpy as np
import pytest

from memsurfer_b200 import synth


def icosphere(R=3.0, sub=3):
    t = (1.0 + 5 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], float)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                  [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]])
    v /= np.linalg.norm(v, axis=1)[:, None]
    for _ in range(sub):
        cache, nf, vl = {}, [], list(v)
        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                m = (vl[a] + vl[b]) / 2; vl.append(m / np.linalg.norm(m)); cache[k] = len(vl) - 1
            return cache[k]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        v, f = np.array(vl), np.array(nf)
    return (v * R).astype(np.float32), f.astype(np.uint32)


def test_golden_mesh_normals_and_areas(oracle, golden):
    n64 = oracle.mesh_normals(golden["verts"], golden["faces"])
    a64 = oracle.pointareas(golden["verts"], golden["faces"])
    assert np.abs(n64 - golden["ref_normals"]).max() < 5e-6          # float32 truth of the reference
    assert np.abs(a64 - golden["ref_pareas"]).max() < 1e-6 * a64.max() + 1e-9


def test_golden_kde(oracle, golden):
    v, bboxl = golden["verts"], golden["bboxl"]
    ids1 = np.where(golden["labels"] == 1)[0].astype(np.int32)
    d_all = oracle.kde(v, np.zeros(0, np.int32), 2, 0.2, True, bboxl)
    d_l1 = oracle.kde(v, ids1, 2, 0.2, True, bboxl)
    d3 = oracle.kde(v, np.zeros(0, np.int32), 3, 0.2, False, None)
    # the reference accumulates ~1e4 terms in float32: agreement to a few 1e-5 relative
    assert np.abs(d_all / golden["ref_kde_all"] - 1).max() < 2e-4
    assert np.abs(d_l1 / golden["ref_kde_l1"] - 1).max() < 2e-4
    assert np.abs(d3 / golden["ref_kde3_all_nonper"] - 1).max() < 2e-4
    assert abs(d_all.sum() - v.shape[0]) < 1e-6 and abs(d_l1.sum() - ids1.shape[0]) < 1e-6


def test_oracle_vs_compiled_reference_periodic(oracle):
    if oracle.ref() is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    pts, labels, bbox = synth.make_leaflet(2500, seed=7)
    df, do = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
    t = synth.trim_periodic(pts, df, do, pts.min(0), pts.max(0))
    verts = np.concatenate([pts, t["duplicated_verts"]]); faces = np.concatenate([t["faces"], t["trimmed_faces"]])
    V = pts.shape[0]
    assert np.abs(oracle.mesh_normals(verts, faces)[:V] - oracle.ref_need_normals(verts, faces)[:V]).max() < 5e-6
    a, ar = oracle.pointareas(verts, faces)[:V], oracle.ref_need_pointareas(verts, faces)[:V]
    assert np.abs(a / ar - 1).max() < 1e-5
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    for typ, ids in ((2, np.zeros(0, np.int32)), (3, np.where(labels == 1)[0].astype(np.int32))):
        for gc in (True, False):
            d, dr = oracle.kde(pts, ids, typ, 25.0, gc, bb), oracle.ref_kde(pts, ids, typ, 25.0, gc, bb)
            assert np.abs(d / dr - 1).max() < 1e-4


def test_anchor_plane_and_sphere(oracle):
    g = np.linspace(0, 9, 10)
    x, y = np.meshgrid(g, g)
    v = np.stack([x.ravel(), y.ravel(), np.zeros(100)], 1).astype(np.float32)
    f = synth.planar_delaunay(v[:, :2])
    n = oracle.mesh_normals(v, f)
    assert np.allclose(n, [0, 0, 1])
    assert abs(oracle.pointareas(v, f).sum() - 81.0) < 1e-9
    H, K = oracle.curvatures(v, f)
    interior = (v[:, 0] > 0) & (v[:, 0] < 9) & (v[:, 1] > 0) & (v[:, 1] < 9)
    assert np.abs(H[interior]).max() < 1e-12 and np.abs(K[interior]).max() < 1e-9
    sv, sf = icosphere(3.0, 3)
    H, K = oracle.curvatures(sv, sf)
    A = oracle.pointareas(sv, sf)
    assert abs(np.abs(H).mean() - 1 / 3.0) < 5e-3 and abs(K.mean() - 1 / 9.0) < 2e-3
    assert abs(A.sum() - 4 * np.pi * 9) < 1.0          # inscribed polyhedron: slightly less area
    # Gauss-Bonnet: sum K*dA/3 = 4 pi exactly for the angle-deficit formula on a closed mesh
    tri = sv[sf.astype(np.int64)].astype(np.float64)
    fa = 0.5 * np.linalg.norm(np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]), axis=1)
    dA = np.zeros(sv.shape[0]); np.add.at(dA, sf.ravel(), np.repeat(fa, 3))
    assert abs((K * dA / 3).sum() - 4 * np.pi) < 1e-9


def test_anchor_periodic_torus_deficit_and_kde_counts(oracle):
    pts, labels, bbox = synth.make_leaflet(1600, seed=3)
    df, do = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
    t = synth.trim_periodic(pts, df, do, pts.min(0), pts.max(0))
    verts = np.concatenate([pts, t["duplicated_verts"]]); faces = np.concatenate([t["faces"], t["trimmed_faces"]])
    V = pts.shape[0]
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    H, K = oracle.curvatures(verts, faces, (0.5 * (bb[3] - bb[0]), 0.5 * (bb[4] - bb[1])))
    tri = verts[faces.astype(np.int64)].astype(np.float64)
    fa = 0.5 * np.linalg.norm(np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]), axis=1)
    dA = np.zeros(verts.shape[0]); np.add.at(dA, faces.ravel(), np.repeat(fa, 3))
    assert abs((K[:V] * dA[:V] / 3).sum()) < 1e-5   # torus: deficits sum to 0 (up to the float32 rounding of the duplicated verts)
    raw_all = oracle.kde(pts, np.zeros(0, np.int32), 2, 15.0, False, bb, raw=True)
    raw_sum = sum(oracle.kde(pts, np.where(labels == l)[0].astype(np.int32), 2, 15.0, False, bb, raw=True)
                  for l in np.unique(labels))
    assert np.abs(raw_all / raw_sum - 1).max() < 1e-12
    shifted = pts.copy(); shifted[:, 0] += (bb[3] - bb[0])
    d0 = oracle.kde(pts, np.zeros(0, np.int32), 2, 15.0, True, bb)
    assert abs(d0.sum() - V) < 1e-8


def test_ghosts_knn_pca_orient(oracle, golden):
    v, box = golden["verts"], golden["box"]
    g, src = oracle.ghosts(v, box, 0.2)
    assert v.shape[0] + g.shape[0] == int(golden["orc_m"]) == 19881    # C1 has 9 680 ghosts
    allp = np.concatenate([v, g])
    sub = np.ascontiguousarray(allp[:3000])
    assert (oracle.knn(sub, 19, brute=True) == oracle.knn(sub, 19)).all()
    nbr = oracle.knn(allp, 19)
    assert (nbr == golden["orc_nbr"]).all() and (nbr[:, 0] == np.arange(allp.shape[0])).all()
    nrm = oracle.pca_normals(allp, nbr)
    # Jacobi eigenvector == LAPACK eigenvector up to sign
    i = 1234
    P = allp[nbr[i]].astype(np.float64); C = (P - P.mean(0)).T @ (P - P.mean(0))
    w, U = np.linalg.eigh(C)
    assert min(np.abs(U[:, 0] - nrm[i]).max(), np.abs(U[:, 0] + nrm[i]).max()) < 1e-10
    on, flags, bad = oracle.mst_orient(allp, nbr, nrm)
    assert bad == 0 and flags.all()
    pn = oracle.point_normals(v, box, 0.2, 18, +1)
    assert np.abs(pn - golden["orc_pnormals"]).max() == 0.0
    assert (pn[:, 2] > 0).mean() > 0.99
