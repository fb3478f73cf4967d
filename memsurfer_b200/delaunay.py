"""Periodic 2-D Delaunay on the GPU (first slice).

Replaces the CGAL call of TriMesh::periodicDelaunay (memsurfer/src/TriMesh_cgal.cpp:160-281) for the mesh
producer: ghost images (ms_ghosts) -> k nearest neighbours (ms_knn) -> one Voronoi cell per point
(ms_delaunay_periodic).  Output in the layout of `synth.periodic_delaunay` (the host stand-in): original
vertex ids + periodic offsets of every face with at least one vertex in the original sheet.
"""
import ctypes as C
import logging

import numpy as np
import torch

from . import _lib

LOGGER = logging.getLogger(__name__)


class IncompleteCells(RuntimeError):
    """The cell-by-cell construction could not finish: Voronoi cells that do not close inside the ghost layer (too few
    points for the box; `retry` = a wider ghost layer may help) or stars that do not glue to a triangulation."""

    def __init__(self, msg, retry=False):
        super().__init__(msg)
        self.retry = retry


def wrap_into_box(xy, b0, b1, out=None):
    """TriMesh::wrap_vertices (memsurfer/src/TriMesh.cpp:462-491, called at TriMesh_cgal.cpp:191): x, y folded into
    [box0, box1) by at most one period, float32 arithmetic.  Returns an (n, 3) float32 array with z = 0."""
    xy = np.asarray(xy)[:, :2]
    b0, b1 = np.asarray(b0, np.float32)[:2], np.asarray(b1, np.float32)[:2]
    w = (b1 - b0).astype(np.float32)
    pts = _lib.pinned_empty((xy.shape[0], 3), np.float32) if out is None else out
    pts[:, :2] = xy
    for d in range(2):
        lo = pts[:, d] < b0[d]
        pts[lo, d] += w[d]
        hi = pts[:, d] >= b1[d]
        pts[hi, d] -= w[d]
    pts[:, 2] = 0.0
    return pts


def periodic_delaunay_device(xy, box0, box1, device='cuda:0', knbrs=23, blayer=0.2, split=False):
    """xy (V,2) planar points, box0/box1 the periodic box.  Returns (dfaces (F,3) int64, doffs (F,3,2) int8);
    faces are counter-clockwise and ordered by their smallest vertex id.  split=True: returns instead
    (faces (F0,3) uint32 inside the sheet, pfaces (Fp,3) int64, poffs (Fp,3,2) int8 crossing the box) --
    the first step of the trim, done on the device."""
    _lib.require_cuda()
    L = _lib.lib()
    dev = torch.device(device)
    xy = np.asarray(xy)[:, :2]
    V = xy.shape[0]
    b0 = np.asarray(box0, np.float32)[:2]
    b1 = np.asarray(box1, np.float32)[:2]
    w = (b1 - b0).astype(np.float32)
    pts = wrap_into_box(xy, b0, b1)
    st = _lib.stream_ptr(dev)
    pd = _lib.h2d(pts, dev)
    cap = int(L.ms_pointset_capacity(V, 1))
    pts_all = torch.empty((cap, 3), dtype=torch.float32, device=dev)
    gsrc = torch.empty(cap, dtype=torch.int32, device=dev)
    info = torch.empty(_lib.INFO_WORDS, dtype=torch.int64, device=dev)
    ws = _lib.workspace(L.ms_ghosts_workspace_bytes(V), dev)
    box = _lib.box6(np.array([b0[0], b0[1], -1.0, b1[0], b1[1], 1.0], np.float32))
    _lib.check(L.ms_ghosts(pd.data_ptr(), V, 1, box, float(blayer), pts_all.data_ptr(), gsrc.data_ptr(),
                           info.data_ptr(), ws.data_ptr(), ws.numel(), st), 'ms_ghosts')
    m = int(info[_lib.INFO_M].item())          # the producer is not the per-frame path: it may wait for the count
    k1 = min(int(knbrs) + 1, 32, m)
    nbr = torch.empty((m, k1), dtype=torch.int32, device=dev)
    ws = _lib.workspace(L.ms_knn_workspace_bytes(m, k1), dev)
    _lib.check(L.ms_knn(pts_all.data_ptr(), m, info.data_ptr(), k1, nbr.data_ptr(), ws.data_ptr(), ws.numel(), st), 'ms_knn')
    fcap = 2 * m + 64
    dfaces = torch.empty((fcap, 3), dtype=torch.int32, device=dev)
    doffs = torch.empty((fcap, 3, 2), dtype=torch.int8, device=dev)
    nf, bad = C.c_int64(0), C.c_int64(0)
    ws = _lib.workspace(L.ms_delaunay_periodic_workspace_bytes(V), dev)
    _lib.check(L.ms_delaunay_periodic(pts_all.data_ptr(), m, V, gsrc.data_ptr(), nbr.data_ptr(), k1,
                                      float(w[0]), float(w[1]), float(blayer) * float(min(w)),
                                      dfaces.data_ptr(), doffs.data_ptr(), fcap,
                                      C.byref(nf), C.byref(bad), ws.data_ptr(), ws.numel(), st), 'ms_delaunay_periodic')
    if bad.value < 0:
        raise IncompleteCells(f'the Voronoi cells do not add up to a triangulation of the torus ({-bad.value} star corners off 6 V)')
    if bad.value > 0:
        raise IncompleteCells(f'{bad.value} Voronoi cells cannot be closed inside the ghost layer '
                              f'({blayer} box widths): too few points for this box', retry=blayer < 0.4)
    F = int(nf.value)
    if not split:
        return _lib.d2h(dfaces[:F]).astype(np.int64), _lib.d2h(doffs[:F])
    full = torch.empty((F, 3), dtype=torch.int32, device=dev)
    pfaces = torch.empty((F, 3), dtype=torch.int32, device=dev)
    poffs = torch.empty((F, 3, 2), dtype=torch.int8, device=dev)
    nfull = C.c_int64(0)
    ws = _lib.workspace(L.ms_delaunay_split_workspace_bytes(F), dev)
    _lib.check(L.ms_delaunay_split(dfaces.data_ptr(), doffs.data_ptr(), F, full.data_ptr(), pfaces.data_ptr(), poffs.data_ptr(),
                                   C.byref(nfull), ws.data_ptr(), ws.numel(), st), 'ms_delaunay_split')
    F0 = int(nfull.value)
    return (_lib.d2h(full[:F0]).view(np.uint32), _lib.d2h(pfaces[:F - F0]).astype(np.int64), _lib.d2h(poffs[:F - F0]))


def trim_partial_device(verts, pf, po, bbox0, bbox1, device='cuda:0'):
    """Device twin of `synth.trim_partial` (TriMesh.cpp:513-560 for the faces that cross the box): duplicated vertices
    numbered in first-appearance order by ms_delaunay_duplicates.  pf (Fp,3) vertex ids, po (Fp,3,2) offsets (host);
    returns the same dict (periodic_faces, trimmed_faces, duplicated_vidxs, duplicated_verts)."""
    _lib.require_cuda()
    L = _lib.lib()
    dev = torch.device(device)
    verts = np.asarray(verts, np.float32)
    V, dim = verts.shape
    pf = np.ascontiguousarray(pf, np.int32).reshape(-1, 3)
    po = np.ascontiguousarray(po, np.int8).reshape(-1, 3, 2)
    nf = pf.shape[0]
    out = dict(periodic_faces=pf.astype(np.uint32))
    if nf == 0:
        out.update(trimmed_faces=np.zeros((0, 3), np.uint32), duplicated_vidxs=np.zeros(0, np.uint32), duplicated_verts=np.zeros((0, dim), np.float32))
        return out
    boxw = (np.asarray(bbox1, np.float32)[:dim] - np.asarray(bbox0, np.float32)[:dim]).astype(np.float32)
    cv = _lib.h2d(np.ascontiguousarray(verts[pf.reshape(-1)]), dev)
    pfd, pod = _lib.h2d(pf, dev), _lib.h2d(po, dev)
    tf = torch.empty((nf, 3), dtype=torch.int32, device=dev)
    dids = torch.empty(3 * nf, dtype=torch.int32, device=dev)
    dverts = torch.empty((3 * nf, dim), dtype=torch.float32, device=dev)
    ws = _lib.workspace(L.ms_delaunay_duplicates_workspace_bytes(nf), dev)
    nd = C.c_int64(0)
    with torch.cuda.device(dev):
        _lib.check(L.ms_delaunay_duplicates(cv.data_ptr(), V, dim, pfd.data_ptr(), pod.data_ptr(), nf, float(boxw[0]), float(boxw[1]),
                                            tf.data_ptr(), dids.data_ptr(), dverts.data_ptr(), C.byref(nd), ws.data_ptr(), ws.numel(),
                                            _lib.stream_ptr(dev)), 'ms_delaunay_duplicates')
    n = int(nd.value)
    out.update(trimmed_faces=_lib.d2h(tf).view(np.uint32), duplicated_vidxs=_lib.d2h(dids[:n]).view(np.uint32),
               duplicated_verts=_lib.d2h(dverts[:n]))
    return out
