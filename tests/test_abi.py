"""CPU: the C-ABI library loads and exports every symbol include/memsurfer_b200.h declares;
host-side argument validation mirrors the reference's errors.  No compute calls here."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "memsurfer_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ms_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from memsurfer_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "build with make -C memsurfer_b200/csrc"
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(L, s), f"missing export {s}"
    assert set(syms) == set(_lib.EXPORTS)
    assert _lib.lib().ms_version() >= 100
    assert _lib.lib().ms_pointset_capacity(10, 1) == 40
    assert _lib.lib().ms_kde_workspace_bytes(1000, 0) > 0


def test_membrane_and_trimesh_argument_errors():
    from memsurfer_b200 import Membrane, TriMesh
    with pytest.raises(ValueError):
        Membrane(np.zeros((5, 2)))
    with pytest.raises(ValueError):
        Membrane(np.zeros((5, 3)), periodic=True)
    with pytest.raises(ValueError):
        Membrane(np.zeros((5, 3)), labels=np.zeros(4), bbox=np.zeros((2, 3)))
    with pytest.raises(ValueError):
        TriMesh(np.zeros((5, 4)))
    with pytest.raises(ValueError):
        TriMesh(np.zeros((5, 3)), faces=np.zeros((2, 4)))
    t = TriMesh(np.zeros((5, 3)), faces=np.zeros((2, 3)))
    assert t.vertices.dtype == np.float32 and t.faces.dtype == np.uint32
    with pytest.raises(ValueError):
        t.compute_density(4, 1.0, "x", True, np.empty([0]))
    with pytest.raises(ValueError):
        t.set_bbox(np.zeros(3), np.zeros(2))
    m = Membrane(np.random.rand(10, 3), bbox=np.array([[0, 0, 0], [1, 1, 1.]]), periodic=True)
    assert m.points.dtype == np.float32 and m.bbox.dtype == np.float32 and m.blayer == 0.2
    m.points[0, 0] = -0.25
    assert m.fit_points_to_box_xy() == 1 and abs(m.points[0, 0] - 0.75) < 1e-6


def test_no_cpu_fallback_without_cuda():
    import torch
    from memsurfer_b200 import Membrane, _lib
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    m = Membrane(np.random.rand(50, 3), bbox=np.array([[0, 0, 0], [1, 1, 1.]]), periodic=True)
    with pytest.raises(_lib.MemSurferCudaError):
        m.compute_pnormals()


def test_periodic_trim_layout():
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(900, seed=11)
    df, do = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
    assert ((do == 0).all(axis=2).sum(axis=1) >= 1).all()
    t = synth.trim_periodic(pts, df, do, pts.min(0), pts.max(0))
    V = pts.shape[0]
    assert t["faces"].dtype == np.uint32 and t["faces"].max() < V
    assert t["trimmed_faces"].max() == V + t["duplicated_verts"].shape[0] - 1
    assert ((t["trimmed_faces"] >= V).sum(axis=1) >= 1).all() and ((t["trimmed_faces"] < V).sum(axis=1) >= 1).all()
    assert (t["periodic_faces"] < V).all() and t["periodic_faces"].shape == t["trimmed_faces"].shape
    # every original vertex of a periodic triangulation has a closed fan: 2 * #edges == 3 * #faces locally
    deg = np.bincount(np.concatenate([t["faces"], t["trimmed_faces"]]).ravel(), minlength=V)[:V]
    assert deg.min() >= 3 and abs(deg.mean() - 6.0) < 0.05
    w = pts.max(0) - pts.min(0)
    d = t["duplicated_verts"] - pts[t["duplicated_vidxs"]]
    assert np.all((np.abs(np.abs(d[:, 0]) - w[0]) < 1e-2) | (np.abs(d[:, 0]) < 1e-6))
