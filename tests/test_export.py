"""Export formats of TriMesh (reference: memsurfer/trimesh.py:441-635) -- host side, no GPU needed.

The polydata content is checked against a literal restatement of the reference's per-face loop; the VTP file is
checked by reading it back (VTK itself is not in the image)."""
import logging

import numpy as np
import pytest

from memsurfer_b200 import TriMesh, synth
from memsurfer_b200 import export


def _periodic_mesh(n=900, seed=3, dim=3):
    pts, _, bbox = synth.make_leaflet(n, seed=seed)
    m = TriMesh(pts[:, :2], periodic=True)
    m.set_bbox(bbox[0, :2], bbox[1, :2])
    m.set_periodic_delaunay(*synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1]))      # host producer: no GPU in these tests
    if dim == 2:
        return m, bbox
    m3 = TriMesh(pts, periodic=True)
    m3.set_bbox(bbox[0], bbox[1])
    m3.copy_triangulation(m)
    return m3, bbox


def _reference_cells(mesh):
    """trimesh.py:455-499 as written: concatenate, then drop a face if any edge spans > half the box."""
    verts = np.concatenate((mesh.vertices, mesh._duplicated_verts), axis=0)
    faces = np.concatenate((mesh.faces, mesh._trimmed_faces), axis=0)
    pts = [(float(v[0]), float(v[1]), float(v[2]) if len(v) > 2 else 0.) for v in verts]
    xw = 0.5 * (mesh.bbox[1, 0] - mesh.bbox[0, 0])
    yw = 0.5 * (mesh.bbox[1, 1] - mesh.bbox[0, 1])
    kept = []
    for f in faces:
        is_periodic = False
        for i in range(3):
            p, q = pts[f[i]], pts[f[(i + 1) % 3]]
            is_periodic |= abs(p[0] - q[0]) > xw or abs(p[1] - q[1]) > yw
        if not is_periodic:
            kept.append([int(f[0]), int(f[1]), int(f[2])])
    return np.array(pts, np.float32), np.array(kept, np.int64).reshape(-1, 3)


@pytest.mark.parametrize("dim", [2, 3])
def test_polydata_matches_reference_loop(dim):
    m, _ = _periodic_mesh(dim=dim)
    V = m.vertices.shape[0]
    rng = np.random.default_rng(0)
    m.attributes['parea'] = rng.random(V).astype(np.float32)
    m.attributes['pnormal'] = rng.random((V, 3)).astype(np.float32)
    poly = m.as_vtkpolydata(additional_attributes={'frame': 7, 'name': 'upper',
                                                   'labels': np.array(['A', 'BB', 'C'])[rng.integers(0, 3, V)]})
    pts, cells = _reference_cells(m)
    assert np.array_equal(poly.points, pts) and poly.points.dtype == np.float32
    assert np.array_equal(poly.polys, cells)
    Vd = m._duplicated_vidxs.shape[0]
    assert poly.number_of_points() == V + Vd
    # point data is extended by the values at the duplicated vertices' originals (trimesh.py:524-525)
    assert np.array_equal(poly.point_data['parea'][V:], m.attributes['parea'][m._duplicated_vidxs])
    assert poly.point_data['pnormal'].shape == (V + Vd, 3)
    assert poly.point_data['labels'].shape == (V + Vd,) and poly.point_data['labels'].dtype.kind == 'U'
    assert poly.field_data['frame'][0] == 7 and poly.field_data['name'][0] == 'upper'
    # the trimmed faces that reach the duplicated vertices are all kept, the wrapped originals never appear
    assert cells.shape[0] == m.faces.shape[0] + m._trimmed_faces.shape[0]


def test_polydata_sizes_and_warnings(caplog):
    m, _ = _periodic_mesh()
    poly0 = m.as_vtkpolydata([])
    ncell = poly0.number_of_cells()
    with caplog.at_level(logging.WARNING):
        poly = m.as_vtkpolydata(additional_attributes={'percell': np.arange(ncell), 'bad': np.arange(5)})
    assert 'percell' in poly.cell_data and 'bad' not in poly.cell_data and 'bad' not in poly.point_data
    text = caplog.text
    assert 'Could not add "bad"' in text
    assert 'Requested attribute "pnormal" is not computed' in text   # KNOWN_ATTRIBUTES asked for, none computed


def test_polydata_without_faces_has_one_vertex_cell_per_point():
    pts, _, _ = synth.make_leaflet(100, seed=1)
    m = TriMesh(pts)
    poly = m.as_vtkpolydata([])
    assert poly.polys is None and np.array_equal(poly.verts, np.arange(pts.shape[0]))
    nonper = TriMesh(pts, faces=np.array([[0, 1, 2], [2, 1, 3]]))
    p2 = nonper.as_vtkpolydata([], {'c': np.array([1.5, 2.5])})
    assert np.array_equal(p2.polys, [[0, 1, 2], [2, 1, 3]]) and np.array_equal(p2.cell_data['c'], [1.5, 2.5])


@pytest.mark.parametrize("ascii", [False, True])
def test_write_vtp_round_trip(tmp_path, ascii):
    m, _ = _periodic_mesh(n=400)
    V = m.vertices.shape[0]
    rng = np.random.default_rng(1)
    m.attributes['mean_curv'] = rng.standard_normal(V)                 # float64, like vtkCurvatures
    m.attributes['parea'] = rng.random(V).astype(np.float32)
    m.attributes['pnormal'] = rng.random((V, 3)).astype(np.float32)
    m.attributes['shell'] = rng.integers(-1, 9, V).astype(np.int32)
    extra = {'labels': np.array(['DPPC', 'CHOL'])[rng.integers(0, 2, V)], 'frame': 12}
    fn = str(tmp_path / 'mesh.vtp')
    m.write_vtp(fn, extra, ascii=ascii)
    poly = m.as_vtkpolydata(additional_attributes=extra)
    back = export.read_vtp(fn)
    assert np.array_equal(back['Points'], poly.points)
    assert np.array_equal(back['Polys'], poly.polys)
    for k in ('mean_curv', 'parea', 'pnormal', 'shell', 'labels'):
        assert np.array_equal(back['PointData'][k], poly.point_data[k]), k
        assert back['PointData'][k].dtype == poly.point_data[k].dtype or k == 'labels'
    assert back['FieldData']['frame'][0] == 12
    head = open(fn, 'rb').read(200).decode()
    assert '<VTKFile type="PolyData" version="1.0" byte_order="LittleEndian" header_type="UInt64">' in head


def test_dataframe_and_csv(tmp_path):
    import pandas as pd
    m, _ = _periodic_mesh(n=400)
    V = m.vertices.shape[0]
    rng = np.random.default_rng(2)
    m.attributes['parea'] = rng.random(V).astype(np.float32)
    m.attributes['pnormal'] = rng.random((V, 3)).astype(np.float32)
    df = m.as_pddataframe(additional_attributes={'density': rng.random(V), 'odd': rng.random((V, 5))})
    assert list(df.columns) == ['pos_x', 'pos_y', 'pos_z', 'pnormal_x', 'pnormal_y', 'pnormal_z', 'parea', 'density']
    assert df.shape[0] == V                                            # duplicated vertices are not listed
    assert np.array_equal(df['pos_y'].to_numpy(), m.vertices[:, 1])
    fn = str(tmp_path / 'mesh.csv')
    m.write_pd(fn, {'density': df['density'].to_numpy()})
    back = pd.read_csv(fn, index_col=0)
    assert list(back.columns) == list(df.columns)
    assert np.allclose(back['parea'].to_numpy(), df['parea'].to_numpy(), rtol=1e-7)


def test_order_param_and_timer():
    from memsurfer_b200.utils import get_order_param, Timer
    rng = np.random.default_rng(5)
    a2, a3, b2, b3 = (rng.standard_normal((50, 3)) for _ in range(4))
    z = np.array([0., 0., 2.])
    p = get_order_param(a2, a3, b2, b3, z)
    assert p.shape == (50, 2)
    for col, d in ((0, a2 - a3), (1, b2 - b3)):                       # formula of utils.py:195-199, per row
        for i in range(50):
            c = np.dot(d[i], z) / (np.linalg.norm(d[i]) * 2.0)
            assert abs(p[i, col] - 0.5 * (3 * c * c - 1)) < 1e-14
    aligned = get_order_param(a3 + z, a3, b3 - z, b3)
    assert np.allclose(aligned, 1.0)                                  # parallel and anti-parallel bonds: P2 = 1
    per_lipid = get_order_param(a2, a3, b2, b3, np.tile(z, (50, 1)))
    assert np.allclose(per_lipid, p)
    with pytest.raises(AssertionError):
        get_order_param(a2[:10], a3, b2, b3)
    t = Timer(); t.end()
    assert str(t).endswith('sec.')


@pytest.mark.gpu
def test_membrane_write_all(tmp_path):
    """membrane.py:319-345: the five VTP files of a processed leaflet (no Poisson surface in this image)."""
    import os
    from memsurfer_b200 import Membrane
    pts, labels, bbox = synth.make_leaflet(2500, seed=2, ntypes=3)
    m = Membrane.compute(pts, labels, bbox, True)
    m.compute_properties('smooth')
    Membrane.compute_densities([m], [2], [10.0], True, 'all')
    prefix = str(tmp_path / 'leaflet')
    m.write_all(prefix)
    assert not os.path.exists(prefix + '_surface_poisson.vtp')
    V = pts.shape[0]
    p = export.read_vtp(prefix + '_points.vtp')
    assert np.array_equal(p['Points'], m.points) and np.array_equal(p['PointData']['normals'], m.pnormals)
    assert np.array_equal(p['Verts'], np.arange(V))
    s = export.read_vtp(prefix + '_membrane_smooth.vtp')
    Vd = m.memb_smooth._duplicated_vidxs.shape[0]
    assert s['Points'].shape[0] == V + Vd
    for k in ('pnormal', 'parea', 'mean_curv', 'gaus_curv', 'density_type2_all_k10.0', 'labels'):
        assert s['PointData'][k].shape[0] == V + Vd, k
    assert np.array_equal(s['PointData']['mean_curv'][:V], m.memb_smooth.attributes['mean_curv'])
    assert s['PointData']['mean_curv'].dtype == np.float64 and s['PointData']['parea'].dtype == np.float32
    for suffix in ('_planar', '_membrane_exact'):
        q = export.read_vtp(prefix + suffix + '.vtp')
        assert np.array_equal(q['Polys'], s['Polys'])


@pytest.mark.gpu
def test_example_ex_simple_runs_and_writes(tmp_path):
    """examples/ex_simple.py (the reference's ex_simple flow, configs[0]) end to end with string labels."""
    import importlib.util, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location('ex_simple', os.path.join(root, 'examples', 'ex_simple.py'))
    ex = importlib.util.module_from_spec(spec); spec.loader.exec_module(ex)
    m, prefix = ex.main(['--out', str(tmp_path / 'noisy')])
    s = export.read_vtp(prefix + '_membrane_smooth.vtp')
    V = m.npoints
    assert s['FieldData']['frame'][0] == 0 and s['FieldData']['time'][0] == 0.0
    assert np.array_equal(s['PointData']['labels'][:V], m.labels)
    d = s['PointData']['density_type2_all_k0.2'][:V]
    assert abs(float(d.sum()) - V) < 1e-2 * V                          # get_nlipdis: the density sums to the lipid count
    assert 'density_type2_1_k0.2' in s['PointData']
    assert np.array_equal(s['PointData']['gaus_curv'][:V], m.memb_smooth.attributes['gaus_curv'])


def test_create_logger(tmp_path):
    from memsurfer_b200.utils import create_logger
    root = logging.getLogger()
    before, level = list(root.handlers), root.level
    try:
        lg = create_logger(2, 0, 1, str(tmp_path / 'logs'), 'run')
        lg.info('hello from the test')
        lg.debug('not shown at level 2')
        for h in root.handlers:
            h.flush()
        text = open(tmp_path / 'logs' / 'run.log').read()
        assert 'hello from the test' in text and '[INFO]' in text and 'not shown' not in text
        with pytest.raises(AssertionError):
            create_logger(0, 1, 0, '', '')
    finally:
        for h in list(root.handlers):
            if h not in before:
                root.removeHandler(h); h.close()
        root.setLevel(level)


@pytest.mark.gpu
def test_example_ex_3lipid_runs_and_writes(tmp_path):
    """examples/ex_3lipid.py: two frames of a synthetic 3-type bilayer, all fields present in the files."""
    import importlib.util, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location('ex_3lipid', os.path.join(root, 'examples', 'ex_3lipid.py'))
    ex = importlib.util.module_from_spec(spec); spec.loader.exec_module(ex)
    prefix = str(tmp_path / '3lipid')
    res = ex.main(['--lipids', '5000', '--frames', '2', '--out', prefix])
    assert len(res) == 2
    for frame, (mt, mb) in enumerate(res):
        for m, side in ((mt, 'top'), (mb, 'bot')):
            s = export.read_vtp(f'{prefix}_f{frame}-{side}_membrane_smooth.vtp')
            V = m.npoints
            assert s['FieldData']['frame'][0] == frame
            names = [f'density_type2_{l}_k{k:.1f}' for l in ('all', 'DPPC', 'DIPC', 'CHOL') for k in (10, 20, 30, 40)]
            for n in names:
                assert np.array_equal(s['PointData'][n][:V], m.attributes[n]), n
            assert set(np.unique(s['PointData']['labels'])) == {'DPPC', 'DIPC', 'CHOL'}


def test_write_vtp_edge_cases(tmp_path):
    """A mesh with an empty face list, a 2-D non-periodic mesh, and a point cloud (Verts only)."""
    pts, _, _ = synth.make_leaflet(64, seed=9)
    empty = TriMesh(pts, faces=np.zeros((0, 3), np.uint32))
    fn = str(tmp_path / 'empty.vtp')
    empty.write_vtp(fn)
    back = export.read_vtp(fn)
    assert back['Points'].shape == (pts.shape[0], 3) and back['Polys'].shape == (0, 3)
    flat = TriMesh(pts[:, :2])
    flat.compute_delaunay()
    fn = str(tmp_path / 'flat.vtp')
    flat.write_vtp(fn, {'id': np.arange(pts.shape[0], dtype=np.int64)}, ascii=True)
    back = export.read_vtp(fn)
    assert np.array_equal(back['Points'][:, :2], pts[:, :2]) and not back['Points'][:, 2].any()
    assert np.array_equal(back['Polys'], flat.faces) and np.array_equal(back['PointData']['id'], np.arange(pts.shape[0]))
    cloud = TriMesh(pts)
    fn = str(tmp_path / 'cloud.vtp')
    cloud.write_vtp(fn, {'normals': np.tile([0., 0., 1.], (pts.shape[0], 1))})
    back = export.read_vtp(fn)
    assert back['Polys'] is None or back['Polys'].shape[0] == 0
    assert np.array_equal(back['Verts'], np.arange(pts.shape[0])) and back['PointData']['normals'].shape == (pts.shape[0], 3)
