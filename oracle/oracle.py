"""TEST INFRASTRUCTURE ONLY -- ctypes front-end for the CPU oracle.

`libmsoracle.so`  : fp64 C restatement of the hot path (oracle/ms_oracle.c).
`_ref/libmsref.so`: the reference's own float32 C++ compiled in place from
                    /root/reference (oracle/ref_shim.cpp), when it was built.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline leg and
--impl reference) may import this module.  The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build():
    """Compile the C restatement (and oracle/_ref when /root/reference exists)."""
    subprocess.run(["make", "-s", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libmsoracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_ghosts.restype = C.c_int64
        L.orc_ghosts.argtypes = [_f32p, C.c_int64, _f32p, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_mesh_normals.argtypes = [_f32p, C.c_int64, _u32p, C.c_int64, _f64p]
        L.orc_pointareas.argtypes = [_f32p, C.c_int64, _u32p, C.c_int64, _f64p]
        L.orc_curvatures.argtypes = [_f32p, C.c_int64, _u32p, C.c_int64, C.c_double, C.c_double, _f64p, _f64p]
        L.orc_face_keep.argtypes = [_f32p, _u32p, C.c_int64, C.c_double, C.c_double, _u8p]
        L.orc_kde.argtypes = [_f32p, C.c_int64, _i32p, C.c_int64, C.c_int, C.c_int, _f32p, C.c_float, C.c_int, _f64p]
        L.orc_kde_raw.argtypes = [_f32p, C.c_int64, _i32p, C.c_int64, C.c_int, C.c_int, _f32p, C.c_float, _f64p]
        L.orc_kde_raw_targets.argtypes = [_f32p, C.c_int64, _i32p, C.c_int64, _i32p, C.c_int64, C.c_int, C.c_int, _f32p, C.c_float, _f64p]
        L.orc_knn_brute.argtypes = [_f32p, C.c_int64, C.c_int, _i32p, C.c_void_p]
        L.orc_knn_grid.argtypes = [_f32p, C.c_int64, C.c_int, _i32p, C.c_void_p]
        L.orc_pca_normals.argtypes = [_f32p, C.c_int64, _i32p, C.c_int, _f64p]
        L.orc_mst_orient.restype = C.c_int64
        L.orc_mst_orient.argtypes = [_f32p, C.c_int64, _i32p, C.c_int, _f64p, _u8p]
        L.orc_closest_point.argtypes = [_f32p, C.c_int64, _u32p, C.c_int64, _f32p, C.c_int64, _i32p, _f64p, _f64p]
        L.orc_geodesics.argtypes = [_f32p, C.c_int64, _u32p, C.c_int64, C.c_int, _f32p, _f64p]
        L.orc_kde_geodesic.argtypes = [_f64p, C.c_int64, _i32p, C.c_int64, C.c_float, C.c_int, _f64p]
        L.orc_need_neighbors.restype = C.c_int64
        L.orc_need_neighbors.argtypes = [_u32p, C.c_int64, C.c_int64, _i32p, _i32p, C.c_int64]
        L.orc_shells.argtypes = [_i32p, _i32p, C.c_int64, _i32p, C.c_int64, _i32p]
        _LIB = L
    return _LIB


def ref():
    """The compiled reference (float32), or None when oracle/_ref was not built."""
    global _REF
    if _REF is None:
        path = os.path.join(_HERE, "_ref", "libmsref.so")
        if not os.path.exists(path):
            return None
        R = C.CDLL(path)
        R.msref_need_normals.argtypes = [_f32p, C.c_int, _u32p, C.c_int, _f32p]
        R.msref_need_pointareas.argtypes = [_f32p, C.c_int, _u32p, C.c_int, _f32p]
        R.msref_kde.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, _f32p, C.c_int,
                                _i32p, C.c_int, _f32p]
        if hasattr(R, "msref_kde_geodesic"):
            R.msref_kde_geodesic.argtypes = [_f32p, C.c_int, _u32p, C.c_int, C.c_int, C.c_float, C.c_int, _f32p, C.c_int,
                                             _i32p, C.c_int, _f32p]
        if hasattr(R, "msref_need_neighbors"):
            R.msref_need_neighbors.restype = C.c_int64
            R.msref_need_neighbors.argtypes = [_f32p, C.c_int, _u32p, C.c_int, _i32p, C.c_int64]
        if hasattr(R, "msref_trim_periodic"):
            _i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
            _i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
            R.msref_trim_periodic.argtypes = [_f32p, C.c_int, C.c_int, _f32p, C.c_int, _i32p, _i8p, C.c_int,
                                              _i32p, _i32p, _i32p, _i32p, _f32p, _i64p]
            R.msref_wrap_vertices.argtypes = [_f32p, C.c_int, C.c_int, _f32p, C.c_int, C.c_int, _f32p]
        _REF = R
    return _REF


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _faces(f):
    return np.ascontiguousarray(f, dtype=np.uint32).reshape(-1, 3)


# ---------------------------------------------------------------------------- fp64 restatement
def ghosts(points, bbox, blayer=0.2):
    """PointSet.cpp:94-147 -> (ghost points (G,3) f32, source index (G,) i32)."""
    p = _f32(points)
    box = _f32(np.asarray(bbox).reshape(-1))
    n = lib().orc_ghosts(p, p.shape[0], box, blayer, None, None)
    out = np.empty((n, 3), np.float32)
    src = np.empty((n,), np.int32)
    lib().orc_ghosts(p, p.shape[0], box, blayer, out.ctypes.data, src.ctypes.data)
    return out, src


def mesh_normals(verts, faces):
    v, f = _f32(verts), _faces(faces)
    out = np.empty((v.shape[0], 3), np.float64)
    lib().orc_mesh_normals(v, v.shape[0], f, f.shape[0], out)
    return out


def pointareas(verts, faces):
    v, f = _f32(verts), _faces(faces)
    out = np.empty((v.shape[0],), np.float64)
    lib().orc_pointareas(v, v.shape[0], f, f.shape[0], out)
    return out


def curvatures(verts, faces, half_box=None):
    """half_box = (xw, yw) enables the trimesh.py:484-499 face filter."""
    v, f = _f32(verts), _faces(faces)
    xw, yw = (-1.0, -1.0) if half_box is None else (float(half_box[0]), float(half_box[1]))
    H = np.empty((v.shape[0],), np.float64)
    K = np.empty((v.shape[0],), np.float64)
    rc = lib().orc_curvatures(v, v.shape[0], f, f.shape[0], xw, yw, H, K)
    assert rc == 0
    return H, K


def kde(verts, ids, type, sigma, get_counts, bbox=None, raw=False):
    v = _f32(verts)
    ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(-1)
    periodic = bbox is not None
    box = _f32(np.asarray(bbox).reshape(-1)) if periodic else np.zeros(6, np.float32)
    out = np.empty((v.shape[0],), np.float64)
    if raw:
        lib().orc_kde_raw(v, v.shape[0], ids, ids.shape[0], type, int(periodic), box, sigma, out)
    else:
        lib().orc_kde(v, v.shape[0], ids, ids.shape[0], type, int(periodic), box, sigma, int(get_counts), out)
    return out


def kde_raw_targets(verts, ids, targets, type, sigma, bbox=None):
    """Un-normalised pair sums sfactor*sum_i exp(...) for the listed target vertices only."""
    v = _f32(verts)
    ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(-1)
    tg = np.ascontiguousarray(targets, dtype=np.int32).reshape(-1)
    periodic = bbox is not None
    box = _f32(np.asarray(bbox).reshape(-1)) if periodic else np.zeros(6, np.float32)
    out = np.empty((tg.shape[0],), np.float64)
    lib().orc_kde_raw_targets(v, v.shape[0], ids, ids.shape[0], tg, tg.shape[0], type, int(periodic), box, sigma, out)
    return out


def knn(points, k1, brute=False, with_d2=False):
    p = _f32(points)
    nbr = np.empty((p.shape[0], k1), np.int32)
    d2 = np.empty((p.shape[0], k1), np.float64) if with_d2 else None
    fn = lib().orc_knn_brute if brute else lib().orc_knn_grid
    fn(p, p.shape[0], k1, nbr, d2.ctypes.data if with_d2 else None)
    return (nbr, d2) if with_d2 else nbr


def pca_normals(points, nbr):
    p = _f32(points)
    nbr = np.ascontiguousarray(nbr, dtype=np.int32)
    out = np.empty((p.shape[0], 3), np.float64)
    lib().orc_pca_normals(p, p.shape[0], nbr, nbr.shape[1], out)
    return out


def mst_orient(points, nbr, normals):
    """Returns (oriented normals (M,3) f64, oriented flags (M,) u8, n_unoriented)."""
    p = _f32(points)
    nbr = np.ascontiguousarray(nbr, dtype=np.int32)
    n = np.array(normals, dtype=np.float64, order="C", copy=True)
    flags = np.zeros((p.shape[0],), np.uint8)
    bad = lib().orc_mst_orient(p, p.shape[0], nbr, nbr.shape[1], n, flags)
    return n, flags, int(bad)


def point_normals(points, bbox=None, blayer=0.2, knbrs=18, ndir_hint=+1, return_all=False):
    """Membrane.compute_pnormals (membrane.py:124-157) end to end, fp64."""
    p = _f32(points)
    n = p.shape[0]
    allp = p
    if bbox is not None:
        g, _ = ghosts(p, bbox, blayer)
        allp = np.ascontiguousarray(np.concatenate([p, g], axis=0))
    nbr = knn(allp, knbrs + 1)
    nrm = pca_normals(allp, nbr)
    nrm, flags, bad = mst_orient(allp, nbr, nrm)
    out = nrm[:n].copy()
    imax = int(np.argmax(np.abs(out[:, 2].astype(np.float32))))
    if out[imax, 2] * ndir_hint < 0:
        out *= -1.0
    if return_all:
        return out, dict(points=allp, nbr=nbr, normals_all=nrm, oriented=flags, n_unoriented=bad)
    return out


# ---------------------------------------------------------------------------- compiled reference (float32)
def ref_need_normals(verts, faces):
    v, f = _f32(verts), _faces(faces)
    out = np.empty((v.shape[0], 3), np.float32)
    assert ref().msref_need_normals(v, v.shape[0], f, f.shape[0], out) == 0
    return out


def ref_need_pointareas(verts, faces):
    v, f = _f32(verts), _faces(faces)
    out = np.empty((v.shape[0],), np.float32)
    assert ref().msref_need_pointareas(v, v.shape[0], f, f.shape[0], out) == 0
    return out


def ref_kde(verts, ids, type, sigma, get_counts, bbox=None):
    v = _f32(verts)
    ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(-1)
    periodic = bbox is not None
    box = _f32(np.asarray(bbox).reshape(-1)) if periodic else np.zeros(6, np.float32)
    out = np.empty((v.shape[0],), np.float32)
    assert ref().msref_kde(v, v.shape[0], type, int(get_counts), sigma, int(periodic), box, 6, ids, ids.shape[0], out) == 0
    return out


def closest_point(verts, faces, queries):
    """(face id i32, barycentric (nq,3) f64, distance f64) of the closest mesh point to every query."""
    verts = np.ascontiguousarray(verts, np.float32); faces = np.ascontiguousarray(faces, np.uint32)
    queries = np.ascontiguousarray(queries, np.float32)
    nq = queries.shape[0]
    fid = np.zeros(nq, np.int32); bary = np.zeros((nq, 3), np.float64); dist = np.zeros(nq, np.float64)
    lib().orc_closest_point(verts, verts.shape[0], faces, faces.shape[0], queries, nq, fid, bary, dist)
    return fid, bary, dist


def need_neighbors(faces, nv):
    """(off (nv+1,), lists) int32: TriMesh::need_neighbors restated (TriMesh.cpp:197-255)."""
    faces = _faces(faces)
    off = np.zeros(nv + 1, np.int32)
    lists = np.zeros(max(1, 6 * faces.shape[0]), np.int32)
    tot = lib().orc_need_neighbors(faces, faces.shape[0], nv, off, lists, lists.shape[0])
    assert tot >= 0
    return off, lists[:tot].copy()


def shells(off, lists, ref_pts):
    """(nv,) int32 shell index by the FIFO breadth-first loop of trimesh.py:392-404 (-1: unreached)."""
    off = np.ascontiguousarray(off, np.int32); lists = np.ascontiguousarray(lists, np.int32)
    ref = np.ascontiguousarray(ref_pts, np.int32).reshape(-1)
    nv = off.shape[0] - 1
    out = np.zeros(nv, np.int32)
    lib().orc_shells(off, lists, nv, ref, ref.shape[0], out)
    return out


def ref_need_neighbors(verts, faces):
    """The compiled reference's TriMesh::need_neighbors(): linearised [counts..., lists...] (int32)."""
    R = ref()
    verts = _f32(verts); faces = _faces(faces)
    out = np.zeros(verts.shape[0] + 6 * faces.shape[0] + 1, np.int32)
    n = R.msref_need_neighbors(verts, verts.shape[0], faces, faces.shape[0], out, out.shape[0])
    assert n >= 0
    return out[:n].copy()


def geodesics(verts, faces, bbox=None):
    """(V,V) f64 graph distances along the mesh edges (TriMesh::compute_geodesics_fw restated)."""
    verts = _f32(verts); faces = _faces(faces)
    nv = verts.shape[0]
    D = np.zeros((nv, nv), np.float64)
    bb = np.zeros(6, np.float32) if bbox is None else np.ascontiguousarray(bbox, np.float32).reshape(-1)
    lib().orc_geodesics(verts, nv, faces, faces.shape[0], int(bbox is not None), bb, D)
    return D


def kde_geodesic(verts, faces, ids, sigma, get_counts, bbox=None):
    """(V,) f64 type-1 density (kde_2m + normalisation) on the geodesic distances of `faces`."""
    D = geodesics(verts, faces, bbox)
    ids = np.ascontiguousarray(ids, np.int32).reshape(-1)
    out = np.zeros(D.shape[0], np.float64)
    lib().orc_kde_geodesic(D, D.shape[0], ids, ids.shape[0], float(sigma), int(bool(get_counts)), out)
    return out


def ref_kde_geodesic(verts, faces, ids, sigma, get_counts, bbox=None):
    """The compiled reference's TriMesh::kde(type 1) (float32)."""
    R = ref()
    verts = _f32(verts); faces = _faces(faces)
    ids = np.ascontiguousarray(ids, np.int32).reshape(-1)
    bb = np.zeros(6, np.float32) if bbox is None else np.ascontiguousarray(bbox, np.float32).reshape(-1)
    out = np.zeros(verts.shape[0], np.float32)
    rc = R.msref_kde_geodesic(verts, verts.shape[0], faces, faces.shape[0], int(bool(get_counts)), float(sigma),
                              int(bbox is not None), bb, 6, ids, ids.shape[0], out)
    assert rc == 0
    return out


def ref_trim_periodic(verts, dfaces, doffs, bbox0, bbox1):
    """The COMPILED reference's TriMesh::trim_periodicDelaunay (TriMesh.cpp:493-575) on a given periodic Delaunay
    (mDelaunayFaces set through the opened class, oracle/ref_shim.cpp): dict with the five arrays of
    trimesh.py:421-436 in the reference's dtypes."""
    R = ref()
    v = _f32(verts)
    nv, dim = v.shape
    bb = np.ascontiguousarray(np.concatenate([np.asarray(bbox0, np.float32)[:dim], np.asarray(bbox1, np.float32)[:dim]]), np.float32)
    df = np.ascontiguousarray(dfaces, np.int32).reshape(-1, 3)
    do = np.ascontiguousarray(doffs, np.int8).reshape(-1, 3, 2)
    nf = df.shape[0]
    faces, pf, tf = (np.zeros((nf, 3), np.int32) for _ in range(3))
    dids = np.zeros(3 * nf, np.int32)
    dverts = np.zeros((3 * nf, dim), np.float32)
    counts = np.zeros(3, np.int64)
    rc = R.msref_trim_periodic(v, nv, dim, bb, bb.shape[0], df, do, nf, faces, pf, tf, dids, dverts, counts)
    assert rc == 0, rc
    n0, n1, n2 = (int(x) for x in counts)
    return dict(faces=faces[:n0].astype(np.uint32), periodic_faces=pf[:n1].astype(np.uint32), trimmed_faces=tf[:n1].astype(np.uint32),
                duplicated_vidxs=dids[:n2].astype(np.uint32), duplicated_verts=dverts[:n2].copy())


def ref_wrap_vertices(verts, bbox0, bbox1, wrap_dim=2):
    """The compiled reference's TriMesh::wrap_vertices (TriMesh.cpp:462-491)."""
    R = ref()
    v = _f32(verts)
    nv, dim = v.shape
    bb = np.ascontiguousarray(np.concatenate([np.asarray(bbox0, np.float32)[:dim], np.asarray(bbox1, np.float32)[:dim]]), np.float32)
    out = np.zeros_like(v)
    assert R.msref_wrap_vertices(v, nv, dim, bb, bb.shape[0], int(wrap_dim), out) == 0
    return out
