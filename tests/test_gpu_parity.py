"""GPU (B200): the CUDA path, called through the C ABI / the drop-in classes, against the CPU
oracle on the same inputs.  Bars (BASELINE.json north_star): neighbour indices and connectivity
bit-exact; normals, areas, curvatures, densities within 1e-9 relative or 1e-12 absolute (fp64)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-9, 1e-12


def close(a, b, rtol=RTOL, atol=ATOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    err = np.abs(a - b) - (atol + rtol * np.abs(b))
    return bool((err <= 0).all()), float(np.abs(a - b).max())


@pytest.fixture(scope="module")
def leaflet():
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(6400, seed=5)
    return pts, labels, bbox


@pytest.fixture(scope="module")
def membrane(leaflet):
    from memsurfer_b200 import Membrane
    pts, labels, bbox = leaflet
    m = Membrane(pts, labels=labels, periodic=True, bbox=bbox)
    m.compute_membrane_surface()
    return m


def concat(mesh):
    return mesh._concat_host()


def test_mesh_properties_periodic(oracle, membrane):
    mesh = membrane.memb_smooth
    verts, faces = concat(mesh)
    V = mesh.vertices.shape[0]
    n32, a32 = mesh.compute_normals(), mesh.compute_pointareas()
    H, K = mesh.compute_curvatures()
    assert n32.dtype == np.float32 and n32.shape == (V, 3) and a32.dtype == np.float32 and H.dtype == np.float64
    on = oracle.mesh_normals(verts, faces)[:V]; oa = oracle.pointareas(verts, faces)[:V]
    oH, oK = oracle.curvatures(verts, faces, mesh._half_box())
    for name, got, want in (("normals", mesh.attributes_f64["pnormal"], on), ("areas", mesh.attributes_f64["parea"], oa),
                            ("mean", H, oH[:V]), ("gauss", K, oK[:V])):
        ok, err = close(got, want)
        assert ok, f"{name}: max abs err {err}"
    # normals and areas are formed in the same order as the sequential oracle: bit-exact
    assert np.array_equal(mesh.attributes_f64["pnormal"], on) and np.array_equal(mesh.attributes_f64["parea"], oa)
    assert np.array_equal(n32, on.astype(np.float32))
    # connectivity handed back untouched
    assert np.array_equal(np.concatenate([mesh.faces, mesh._trimmed_faces]), faces)


def test_mesh_properties_nonperiodic_and_degenerate(oracle):
    from memsurfer_b200 import TriMesh, synth
    rng = np.random.default_rng(0)
    v = rng.uniform(0, 10, (500, 3)).astype(np.float32); v[:, 2] *= 0.1
    f = synth.planar_delaunay(v[:, :2])
    f = np.concatenate([f, [[3, 3, 7], [5, 9, 5]]]).astype(np.uint32)      # degenerate faces
    v = np.concatenate([v, [[50, 50, 0]]]).astype(np.float32)               # isolated vertex -> (1,0,0), area 0
    mesh = TriMesh(v, faces=f)
    mesh.compute_normals(); mesh.compute_pointareas(); H, K = mesh.compute_curvatures()
    assert np.array_equal(mesh.attributes_f64["pnormal"], oracle.mesh_normals(v, f))
    assert np.array_equal(mesh.attributes_f64["pnormal"][-1], [1.0, 0.0, 0.0])
    oa = oracle.pointareas(v, f)
    assert np.array_equal(np.isnan(mesh.attributes_f64["parea"]), np.isnan(oa))
    fin = np.isfinite(oa)
    assert close(mesh.attributes_f64["parea"][fin], oa[fin])[0]
    oH, oK = oracle.curvatures(v, f)
    assert close(H, oH)[0] and close(K, oK)[0] and H[-1] == 0.0 and K[-1] == 0.0


def test_golden_noisy_off_mesh_and_kde(oracle, golden):
    from memsurfer_b200 import TriMesh
    mesh = TriMesh(golden["verts"], faces=golden["faces"], periodic=True)
    mesh.set_bbox(golden["bboxl"][:3], golden["bboxl"][3:])
    assert np.abs(mesh.compute_normals() - golden["ref_normals"]).max() < 5e-6
    assert np.abs(mesh.compute_pointareas() - golden["ref_pareas"]).max() < 1e-9 + 1e-6 * golden["ref_pareas"].max()
    ids1 = np.where(golden["labels"] == 1)[0]
    d_all = mesh.compute_density(2, 0.2, "all", True, np.empty([0]))
    d_l1 = mesh.compute_density(2, 0.2, "l1", True, ids1)
    assert d_all.dtype == np.float32
    assert np.abs(d_all / golden["ref_kde_all"] - 1).max() < 2e-4 and np.abs(d_l1 / golden["ref_kde_l1"] - 1).max() < 2e-4
    assert close(mesh.attributes_f64["all"], oracle.kde(golden["verts"], np.zeros(0, np.int32), 2, 0.2, True, golden["bboxl"]))[0]
    assert mesh.compute_density(2, 0.9, "all", True, np.empty([0])) is d_all        # name-keyed cache


@pytest.mark.parametrize("typ", [2, 3])
@pytest.mark.parametrize("periodic", [True, False])
@pytest.mark.parametrize("get_counts", [True, False])
def test_kde_vs_oracle(oracle, leaflet, typ, periodic, get_counts):
    from memsurfer_b200 import TriMesh
    pts, labels, bbox = leaflet
    mesh = TriMesh(pts, periodic=periodic)
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    if periodic:
        mesh.set_bbox(bb[:3], bb[3:])
    for sigma in (10.0, 40.0):
        for ids in (np.empty([0]), np.where(labels == 1)[0], np.array([17])):
            name = f"d{sigma}_{len(ids)}"
            mesh.compute_density(typ, sigma, name, get_counts, ids)
            want = oracle.kde(pts, ids.astype(np.int32), typ, sigma, get_counts, bb if periodic else None)
            ok, err = close(mesh.attributes_f64[name], want)
            assert ok, f"type {typ} periodic {periodic} sigma {sigma} nids {len(ids)}: max abs err {err}"
            if get_counts:
                assert abs(mesh.attributes_f64[name].sum() - (len(ids) or pts.shape[0])) < 1e-6


@pytest.mark.parametrize("typ", [2, 3])
@pytest.mark.parametrize("periodic", [True, False])
def test_kde_cutoff_cell_list_vs_oracle(oracle, leaflet, typ, periodic):
    """Cut-off (cell list) variant against the oracle's ALL-PAIRS sum: the dropped tail is below
    exp(-7^2/2) = 2.3e-11 relative, far inside the 1e-9 bar."""
    from memsurfer_b200 import TriMesh
    pts, labels, bbox = leaflet
    mesh = TriMesh(pts, periodic=periodic)
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    if periodic:
        mesh.set_bbox(bb[:3], bb[3:])
    for sigma in (10.0, 25.0, 40.0):
        for ids in (np.empty([0]), np.where(labels == 1)[0], np.array([17, 4000])):
            name = f"c{sigma}_{len(ids)}"
            mesh.compute_density(typ, sigma, name, False, ids, cutoff_sigmas=7.0)
            want = oracle.kde(pts, ids.astype(np.int32), typ, sigma, False, bb if periodic else None)
            got = mesh.attributes_f64[name]
            scale = want.max()
            err = np.abs(got - want)
            assert (err <= 1e-12 * scale + 1e-9 * want + 3e-11 * scale).all(), \
                f"type {typ} periodic {periodic} sigma {sigma} nids {len(ids)}: max err {err.max()} (scale {scale})"
            if len(ids) != 2:
                assert close(got, want)[0]


def test_kde_cutoff_matches_all_pairs_at_full_size():
    """BASELINE configs[2] leaflet size (499 849 vertices): cell list vs the all-pairs kernel."""
    from memsurfer_b200 import TriMesh, synth
    pts, labels, bbox = synth.make_leaflet(500000, seed=1234, ntypes=3, fractions=[0.4, 0.3, 0.3])
    mesh = TriMesh(pts, periodic=True)
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    mesh.set_bbox(bb[:3], bb[3:])
    ids = np.where(labels == 2)[0]
    for sigma, sel, nm in ((10.0, np.empty([0]), "a"), (40.0, ids, "b")):
        full = mesh.compute_density(2, sigma, nm + "_full", True, sel) * 1.0
        f64 = mesh.attributes_f64[nm + "_full"]
        mesh.compute_density(2, sigma, nm + "_cut", True, sel, cutoff_sigmas=7.0)
        c64 = mesh.attributes_f64[nm + "_cut"]
        ok, err = close(c64, f64)
        assert ok, f"sigma {sigma}: max abs err {err}"
        assert abs(c64.sum() - (len(sel) or pts.shape[0])) < 1e-5


@pytest.mark.parametrize("cutoff", [0.0, 8.0])
@pytest.mark.parametrize("periodic", [True, False])
def test_kde_label_fused_vs_oracle(oracle, leaflet, cutoff, periodic):
    """ms_kde_labels + ms_kde_select (one pass for every label) against the oracle's per-field sums."""
    from memsurfer_b200 import TriMesh
    pts, labels, bbox = leaflet
    labels = labels.copy(); labels[::97] = -1                      # some vertices are never a source
    nl = int(labels.max()) + 1
    mesh = TriMesh(pts, periodic=periodic)
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    if periodic:
        mesh.set_bbox(bb[:3], bb[3:])
    mesh.set_vertex_labels(labels, nl)
    for typ, sigma in ((2, 12.0), (3, 30.0)):
        for sel in ([0], [1, 2], list(range(nl))):
            for gc in (True, False):
                name = f"f{typ}_{sigma}_{sel}_{gc}"
                mesh.compute_density_labels(typ, sigma, name, gc, sel, cutoff_sigmas=cutoff)
                ids = np.where(np.isin(labels, sel))[0].astype(np.int32)
                want = oracle.kde(pts, ids, typ, sigma, gc, bb if periodic else None)
                ok, err = close(mesh.attributes_f64[name], want)
                assert ok, f"cutoff {cutoff} periodic {periodic} type {typ} sel {sel} counts {gc}: max abs err {err}"


def test_membrane_fused_densities_match_unfused(leaflet):
    from memsurfer_b200 import Membrane
    pts, labels, bbox = leaflet
    res = []
    for lab in (labels, None):
        m = Membrane(pts, labels=lab, periodic=True, bbox=bbox)
        m.compute_membrane_surface()
        Membrane.compute_densities([m], [2], [15.0], True, 'all')
        if lab is not None:
            Membrane.compute_densities([m], [2], [15.0], True, 1)
            res.append(m.memb_smooth.attributes_f64['density_type2_1_k15.0'])
            ids = np.where(labels == 1)[0]
            m.memb_smooth.compute_density(2, 15.0, 'unfused_1', True, ids, cutoff_sigmas=8.0)
            res.append(m.memb_smooth.attributes_f64['unfused_1'])
        res.append(m.memb_smooth.attributes_f64['density_type2_all_k15.0'])
    f1, u1, fall, uall = res
    assert close(f1, u1)[0] and close(fall, uall)[0]


def test_kde_large_register_blocked_path(oracle):
    """V >= 148*1024 targets takes the R=4 register-blocked kernel; oracle on a target subset."""
    from memsurfer_b200 import TriMesh, synth
    pts, labels, bbox = synth.make_leaflet(160000, seed=2)
    V = pts.shape[0]
    mesh = TriMesh(pts, periodic=True)
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    mesh.set_bbox(bb[:3], bb[3:])
    ids = np.where(labels == 2)[0]
    mesh.compute_density(2, 20.0, "all", False, np.empty([0]))
    mesh.compute_density(3, 30.0, "sub", False, ids)
    tg = np.arange(0, V, 397)
    want = oracle.kde_raw_targets(pts, np.zeros(0, np.int32), tg, 2, 20.0, bb) / V
    ok, err = close(mesh.attributes_f64["all"][tg], want); assert ok, err
    want = oracle.kde_raw_targets(pts, ids.astype(np.int32), tg, 3, 30.0, bb) / V
    ok, err = close(mesh.attributes_f64["sub"][tg], want); assert ok, err


def test_pnormals_stages(oracle, golden):
    import torch
    from memsurfer_b200 import point_normals_device
    v, box = golden["verts"], golden["box"]
    keep = {}
    o32, o64, info = point_normals_device(torch.from_numpy(v).cuda(), box, True, 0.2, 18, +1, keep=keep)
    g, src = oracle.ghosts(v, box, 0.2)
    allp = np.concatenate([v, g])
    assert info["m"] == allp.shape[0] == 19881
    assert np.array_equal(keep["points_all"].cpu().numpy(), allp)                      # ghosts: bit-exact
    assert np.array_equal(keep["ghost_src"].cpu().numpy()[v.shape[0]:], src)
    nbr = keep["nbr"].cpu().numpy()
    assert np.array_equal(nbr, golden["orc_nbr"])                                       # neighbour indices: bit-exact
    pca = keep["normals_pca"].cpu().numpy()
    opca = oracle.pca_normals(allp, golden["orc_nbr"])
    assert np.array_equal(pca, opca)                                                    # same op order, no FMA
    on, flags, bad = oracle.mst_orient(allp, golden["orc_nbr"], opca)
    assert info["n_unoriented"] == bad == 0
    assert np.array_equal(keep["normals_all"].cpu().numpy(), on)                        # MST orientation: same signs
    assert np.array_equal(o64.cpu().numpy(), golden["orc_pnormals"])
    assert np.array_equal(o32.cpu().numpy(), golden["orc_pnormals"].astype(np.float32))


def test_membrane_dropin_flow(oracle, leaflet):
    from memsurfer_b200 import Membrane
    pts, labels, bbox = leaflet
    m = Membrane.compute(pts, labels, bbox, True)
    assert m.pnormals.dtype == np.float32 and m.pnormals.shape == pts.shape
    want = oracle.point_normals(pts, bbox, 0.2, 18, +1)
    assert close(m.pnormals_f64, want)[0]
    assert (m.pnormals[:, 2] > 0).all()
    m.compute_properties('smooth'); m.compute_properties('exact')
    Membrane.compute_densities([m], [2], [10.0, 20.0], True, 'all')
    Membrane.compute_densities([m], [2], [10.0], True, 1)
    assert set(m.attributes) == {'density_type2_all_k10.0', 'density_type2_all_k20.0', 'density_type2_1_k10.0'}
    assert abs(m.attributes['density_type2_1_k10.0'].sum() - (labels == 1).sum()) < 0.5
    assert set(m.memb_smooth.attributes) >= {'pnormal', 'parea', 'mean_curv', 'gaus_curv'}


def test_nonperiodic_pnormals_small(oracle):
    import torch
    from memsurfer_b200 import point_normals_device
    rng = np.random.default_rng(3)
    p = rng.uniform(-1, 1, (700, 3)).astype(np.float32); p[:, 2] = 0.1 * np.sin(3 * p[:, 0]) + 0.01 * rng.normal(size=700)
    keep = {}
    o32, o64, info = point_normals_device(torch.from_numpy(p).cuda(), None, False, 0.2, 18, -1, keep=keep)
    assert np.array_equal(keep["nbr"].cpu().numpy(), oracle.knn(p, 19, brute=True))
    want = oracle.point_normals(p, None, 0.2, 18, -1)
    assert np.array_equal(o64.cpu().numpy(), want)
