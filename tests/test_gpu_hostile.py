"""GPU (B200): parity on inputs that are NOT a jittered square grid.

 (a) rectangular box, Lx = 2.5 Ly, through the whole drop-in flow;
 (b) an exact lattice: squared distances tie exactly (the k-NN redo path must run) and a Riemannian graph whose
     weights are massively tied and whose sign pattern is frustrated (the MST tie order decides every sign);
 (c) a clustered point set: two densities 4:1, a circular void of 10 sigma and a micro-cluster of 900 points
     inside half an Angstrom (crowded cells, ring expansion across the void, uneven Voronoi cells);
 (d) points exactly on the faces of the KDE box and outside it (the fold of the cell list);
 (e) BASELINE configs[2] at full size through the frame driver bench.py times, against the oracle.
Bars as everywhere: indices bit-exact, fp64 values within 1e-9 relative / 1e-12 absolute.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-9, 1e-12
SIGMAS = [10.0, 20.0, 30.0, 40.0]


def close(a, b, rtol=RTOL, atol=ATOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return bool((np.abs(a - b) <= atol + rtol * np.abs(b)).all()), float(np.abs(a - b).max())


def _check_membrane(oracle, m, pts, labels, bbox, sigmas, nlab):
    """Everything Membrane computed for one leaflet against the oracle; returns the measured max errors."""
    errs = {}
    want = oracle.point_normals(pts, bbox, 0.2, 18, +1)
    assert np.array_equal(m.pnormals_f64, want), "point normals differ from the oracle"
    assert np.array_equal(m.pnormals, want.astype(np.float32))
    mesh = m.memb_smooth
    verts, faces = mesh._concat_host()
    V = pts.shape[0]
    on, oa = oracle.mesh_normals(verts, faces)[:V], oracle.pointareas(verts, faces)[:V]
    assert np.array_equal(mesh.attributes_f64['pnormal'], on) and np.array_equal(mesh.attributes_f64['parea'], oa)
    H, K = oracle.curvatures(verts, faces, mesh._half_box())
    for nm, got, w in (('mean_curv', mesh.attributes['mean_curv'], H[:V]), ('gaus_curv', mesh.attributes['gaus_curv'], K[:V])):
        ok, err = close(got, w); assert ok, (nm, err)
        errs[nm] = err
    bb = mesh.bboxl
    worst = 0.0
    for s in sigmas:
        for l in ['all'] + list(range(nlab)):
            ids = np.zeros(0, np.int32) if l == 'all' else np.where(labels == l)[0].astype(np.int32)
            got = mesh.attributes_f64['density_type2_{}_k{:.1f}'.format(l, s)]
            w = oracle.kde(pts, ids, 2, float(s), True, bb)
            ok, err = close(got, w); assert ok, (s, l, err)
            worst = max(worst, float(np.abs(got / w - 1).max()))
            assert abs(got.sum() - (ids.shape[0] or V)) < 1e-6
    errs['density_rel'] = worst
    return errs


def test_a_rectangular_box_full_flow(oracle):
    """Lx = 2.5 Ly: every grid / cell / strip computation sees wx != wy (DistanceKernels.hpp:120-135)."""
    from memsurfer_b200 import Membrane, synth
    pts, labels, bbox = synth.make_leaflet_rect(125, 50, seed=11)
    assert abs((bbox[1, 0] - bbox[0, 0]) / (bbox[1, 1] - bbox[0, 1]) - 2.5) < 1e-6
    m = Membrane.compute(pts, labels, bbox, True)
    m.compute_properties('smooth')
    for l in ['all', 0, 1, 2]:
        Membrane.compute_densities([m], [2], SIGMAS, True, l)
    errs = _check_membrane(oracle, m, pts, labels, bbox, SIGMAS, 3)
    print('rectangular box max errors', errs)
    # the tall variant (Ly = 2.5 Lx): strips and fine rows swap roles
    pts2, labels2, bbox2 = synth.make_leaflet_rect(40, 100, seed=12)
    m2 = Membrane.compute(pts2, labels2, bbox2, True)
    m2.compute_properties('smooth')
    for l in ['all', 0, 1, 2]:
        Membrane.compute_densities([m2], [2], [10.0, 40.0], True, l)
    _check_membrane(oracle, m2, pts2, labels2, bbox2, [10.0, 40.0], 3)


def test_b_lattice_exact_ties_knn_and_mst(oracle):
    """Unjittered lattice with integer coordinates: exact distance ties in the k-NN (the redo path must run and
    the lists must still equal the oracle's (distance, index) order) and an MST whose weights are all tied."""
    import torch
    from memsurfer_b200 import synth
    from memsurfer_b200.membrane import knn_device, orient_normals_device, point_normals_device
    pts, labels, bbox = synth.make_leaflet_rect(48, 40, seed=3, jitter=0.0, amplitude=0.0)
    assert np.array_equal(pts[:, :2], np.round(pts[:, :2]))         # integers: squared distances are exact
    g, _ = oracle.ghosts(pts, bbox, 0.2)
    allp = np.ascontiguousarray(np.concatenate([pts, g]))
    pd = torch.from_numpy(allp).cuda()
    nbr, ties = knn_device(pd, 19)
    want_nbr = oracle.knn(allp, 19, brute=True)
    assert ties > allp.shape[0] // 2, f"the exact-tie redo path did not run ({ties} queries)"
    assert np.array_equal(nbr.cpu().numpy(), want_nbr)
    # frustrated, massively tied Riemannian graph: three unit normals 120 degrees apart in the yz plane
    # (pairwise dot = -1/2 -> weight float(0.5); equal normals -> weight 0).  Every sign is decided by the
    # tie order (weight bits, a, slot).
    rng = np.random.default_rng(5)
    c, s = -0.5, np.sqrt(3.0) / 2.0
    dirs = np.array([[0.0, 0.0, 1.0], [0.0, s, c], [0.0, -s, c]])
    nrm = np.ascontiguousarray(dirs[rng.integers(0, 3, allp.shape[0])])
    on, flags, bad = oracle.mst_orient(allp, want_nbr, nrm)
    nd = torch.from_numpy(nrm.copy()).cuda()
    bad_gpu = orient_normals_device(pd, nbr, nd)
    assert bad_gpu == bad == 0
    assert np.array_equal(nd.cpu().numpy(), on), "MST orientation differs under tied weights"
    assert (np.sign(on[:, 2]) != np.sign(nrm[:, 2])).sum() > 100           # the orientation did flip things
    # the whole chain on the lattice with a little z structure (ties stay exact in xy only)
    pts2 = pts.copy(); pts2[:, 2] += (0.5 * ((np.arange(pts.shape[0]) * 7) % 5)).astype(np.float32)
    o32, o64, info = point_normals_device(torch.from_numpy(pts2).cuda(), bbox, True, 0.2, 18, +1)
    assert np.array_equal(o64.cpu().numpy(), oracle.point_normals(pts2, bbox, 0.2, 18, +1))


def test_c_clustered_void_and_microcluster(oracle):
    """Two densities 4:1, a void of radius 10 sigma (sigma = 10), 900 points inside half an Angstrom."""
    import torch
    from memsurfer_b200 import Membrane, TriMesh, synth
    from memsurfer_b200.membrane import point_normals_device
    pts, labels, bbox = synth.make_leaflet_clustered(ns=80, seed=7, void_radius=100.0, cluster=900)
    V = pts.shape[0]
    # --- k-NN + PCA + MST: rings must expand across the void and through the crowded cell
    keep = {}
    o32, o64, info = point_normals_device(torch.from_numpy(pts).cuda(), bbox, True, 0.2, 18, +1, keep=keep)
    want, aux = oracle.point_normals(pts, bbox, 0.2, 18, +1, return_all=True)
    assert np.array_equal(keep['nbr'].cpu().numpy(), aux['nbr'])
    assert np.array_equal(o64.cpu().numpy(), want) and info['n_unoriented'] == aux['n_unoriented']
    # --- KDE: label-fused shared binning (crowded fine cells: > 256 sources in one cell) vs the oracle
    mesh = TriMesh(pts, periodic=True)
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    mesh.set_bbox(bb[:3], bb[3:])
    mesh.set_vertex_labels(labels, 3)
    mesh.density_labels_prefetch(2, [10.0, 20.0], 7.5)
    for sigma in (10.0, 20.0):
        for sel in ([0, 1, 2], [1]):
            name = f"d{sigma}_{len(sel)}"
            mesh.compute_density_labels(2, sigma, name, True, sel, cutoff_sigmas=7.5)
            ids = np.where(np.isin(labels, sel))[0].astype(np.int32)
            wantd = oracle.kde(pts, ids if len(sel) < 3 else np.zeros(0, np.int32), 2, sigma, True, bb)
            ok, err = close(mesh.attributes_f64[name], wantd); assert ok, (sigma, sel, err)
    # un-fused cut-off path with an explicit id list, and 3-D
    ids = np.where(labels == 2)[0]
    mesh.compute_density(3, 15.0, "t3", False, ids, cutoff_sigmas=7.5)
    ok, err = close(mesh.attributes_f64["t3"], oracle.kde(pts, ids.astype(np.int32), 3, 15.0, False, bb)); assert ok, err
    # --- mesh properties on the (very uneven) periodic triangulation of these points
    m = Membrane(pts, labels=labels, periodic=True, bbox=bbox, check_overlaps=False)
    m.compute_membrane_surface()
    m.compute_properties('smooth')
    msh = m.memb_smooth
    verts, faces = msh._concat_host()
    assert faces.shape[0] >= 2 * V
    assert np.array_equal(msh.attributes_f64['pnormal'], oracle.mesh_normals(verts, faces)[:V])
    oa = oracle.pointareas(verts, faces)[:V]
    fin = np.isfinite(oa)
    assert np.array_equal(msh.attributes_f64['parea'][fin], oa[fin])
    H, K = oracle.curvatures(verts, faces, msh._half_box())
    fin = np.isfinite(H[:V]) & np.isfinite(K[:V])
    assert close(msh.attributes['mean_curv'][fin], H[:V][fin])[0]
    # Gaussian curvature = 3 * (2 pi - sum of angles) / dA.  Inside the micro-cluster dA is ~1e-6 A^2 (float32
    # coordinates a few ulps apart), so one ulp of an atan2 is amplified by 1e6: the bar applies to the
    # well-conditioned quantity, the angle deficit K * dA / 3 (1e-12 absolute), and to K itself elsewhere.
    from test_gpu_configs import oracle_face_area_per_vertex
    dA = oracle_face_area_per_vertex(verts, faces, msh._half_box(), V)
    ok, err = close((msh.attributes['gaus_curv'] * dA / 3.0)[fin], (K[:V] * dA / 3.0)[fin]); assert ok, err
    big = fin & (dA > 1.0)
    assert big.sum() > V // 2 and close(msh.attributes['gaus_curv'][big], K[:V][big])[0]


def test_d_points_on_and_outside_the_box_faces(oracle):
    """Regression for the period-fold bug: vertices exactly ON both faces of the KDE box and up to
    0.3 periods OUTSIDE it; cut-off (shared binning and single sigma) and all-pairs paths against the oracle,
    whose distance is the reference's single wrap (DistanceKernels.hpp:120-135)."""
    from memsurfer_b200 import TriMesh, synth
    pts, labels, bbox = synth.make_leaflet(6400, seed=9)
    L = float(bbox[1, 0] - bbox[0, 0])
    pts = pts.copy()
    pts[0, 0], pts[1, 0], pts[2, 1], pts[3, 1] = 0.0, L, 0.0, L            # exactly on the four faces
    pts[4, :2] = (0.0, 0.0); pts[5, :2] = (L, L)                              # and on two corners
    out = np.arange(10, 400, 13)
    pts[out[0::4], 0] -= np.float32(0.3 * L)                                  # outside, each side
    pts[out[1::4], 0] += np.float32(0.3 * L)
    pts[out[2::4], 1] -= np.float32(0.3 * L)
    pts[out[3::4], 1] += np.float32(0.3 * L)
    box = np.array([0.0, 0.0, -1.0, L, L, 1.0], np.float32)                  # the nominal box, NOT the vertex extent
    mesh = TriMesh(pts, periodic=True)
    mesh.set_bbox(box[:3], box[3:])
    mesh.set_vertex_labels(labels, 3)
    mesh.density_labels_prefetch(2, [10.0, 25.0, 40.0], 7.5)
    for sigma in (10.0, 25.0, 40.0):
        for sel in ([0, 1, 2], [2]):
            ids = np.where(np.isin(labels, sel))[0].astype(np.int32)
            want = oracle.kde(pts, ids if len(sel) < 3 else np.zeros(0, np.int32), 2, sigma, True, box)
            name = f"f{sigma}_{len(sel)}"
            mesh.compute_density_labels(2, sigma, name, True, sel, cutoff_sigmas=7.5)
            ok, err = close(mesh.attributes_f64[name], want); assert ok, ("fused", sigma, sel, err)
            name = f"u{sigma}_{len(sel)}"
            mesh.compute_density(2, sigma, name, True, ids if len(sel) < 3 else np.empty([0]), cutoff_sigmas=7.5)
            ok, err = close(mesh.attributes_f64[name], want); assert ok, ("cut-off", sigma, sel, err)
            name = f"a{sigma}_{len(sel)}"
            mesh.compute_density(2, sigma, name, True, ids if len(sel) < 3 else np.empty([0]), cutoff_sigmas=0.0)
            ok, err = close(mesh.attributes_f64[name], want); assert ok, ("all pairs", sigma, sel, err)
    # ghosts + k-NN for points outside the box (no fit_points_to_box_xy): same cuts as PointSet.cpp:124-128
    import torch
    from memsurfer_b200.membrane import point_normals_device
    keep = {}
    o32, o64, info = point_normals_device(torch.from_numpy(pts).cuda(), bbox, True, 0.2, 18, +1, keep=keep)
    want, aux = oracle.point_normals(pts, bbox, 0.2, 18, +1, return_all=True)
    assert np.array_equal(keep['points_all'].cpu().numpy(), aux['points'])
    assert np.array_equal(keep['nbr'].cpu().numpy(), aux['nbr']) and np.array_equal(o64.cpu().numpy(), want)


def test_e_bench_frame_full_size_vs_oracle(oracle):
    """BASELINE configs[2]: one 499 849-point leaflet through LeafletFrame.run(SIGMAS, 18, 7.5) -- the exact call
    bench.py times -- against the oracle: point normals, vertex normals, areas and curvatures in full, all 16
    density fields on 1 024 targets each."""
    import torch
    from memsurfer_b200 import TriMesh, synth
    from memsurfer_b200.frame import LeafletFrame
    pts, labels, bbox = synth.make_leaflet(500000, seed=1234, z0=20.0, ntypes=3, fractions=[0.4, 0.3, 0.3])
    V = pts.shape[0]
    mesh = TriMesh(pts, periodic=True, label='memb_smooth')
    mesh.set_bbox(pts.min(axis=0), pts.max(axis=0))
    mesh.compute_delaunay()
    fr = LeafletFrame(pts, labels, bbox, mesh)
    fr.run(SIGMAS, 18, 7.5)
    rec = fr.summary().cpu().numpy()
    names = fr.summary_names()
    torch.cuda.synchronize()
    want = oracle.point_normals(pts, bbox, 0.2, 18, +1)
    assert np.array_equal(fr.pnormals64.cpu().numpy(), want) and fr.info['n_unoriented'] == 0
    verts, faces = mesh._concat_host()
    assert np.array_equal(fr.vn64.cpu().numpy(), oracle.mesh_normals(verts, faces)[:V])
    assert np.array_equal(fr.pa64.cpu().numpy(), oracle.pointareas(verts, faces)[:V])
    H, K = oracle.curvatures(verts, faces, mesh._half_box())
    okH, eH = close(fr.H.cpu().numpy(), H[:V]); okK, eK = close(fr.K.cpu().numpy(), K[:V])
    assert okH and okK, (eH, eK)
    bb = mesh.bboxl
    tg = np.arange(0, V, V // 1024)[:1024]
    worst = 0.0
    for s in SIGMAS:
        for l in ['all', 0, 1, 2]:
            ids = np.zeros(0, np.int32) if l == 'all' else np.where(labels == l)[0].astype(np.int32)
            name = 'density_type2_{}_k{:.1f}'.format(l, s)
            got = fr.densities[name][0].cpu().numpy()
            raw = oracle.kde_raw_targets(pts, ids, tg, 2, s, bb) / V
            i = names.index(name)
            nsel = ids.shape[0] or V
            assert abs(rec[i, 0] - nsel) < 1e-5 and abs(got.sum() - rec[i, 0]) < 1e-5
            # get_counts: density * n_sel / sum(density).  The sum runs over all V vertices, the oracle covers a
            # target subset: the field must be one constant times the oracle's raw sums, and integrate to n_sel
            ratio = got[tg] / raw
            assert np.abs(ratio / ratio[0] - 1).max() < 1e-9, (s, l)
            worst = max(worst, float(np.abs(ratio / ratio[0] - 1).max()))
    print(f"configs[2] full size: max rel dev of densities {worst:.3e}, H err {eH:.3e}, K err {eK:.3e}")
