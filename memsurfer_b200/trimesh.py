"""memsurfer_b200.TriMesh -- drop-in for memsurfer.trimesh.TriMesh on the hot path.

Same constructor, attribute names, method names, return dtypes and memoisation as the reference
class (memsurfer/trimesh.py:27-361); the arithmetic runs in hand-written sm_100a kernels through
the C ABI (include/memsurfer_b200.h).  What stays on the host, as in the reference design, is the
mesh PRODUCER: `compute_delaunay` here is a scipy stand-in for the CGAL call
(memsurfer/src/TriMesh_cgal.cpp:160-281) and `copy_triangulation` re-trims with this mesh's own
bbox exactly like TriMesh.hpp:301-304 + TriMesh.cpp:493-575.

Extras that the reference does not have: `attributes_f64` (fp64 twins of the float32 results),
`to_device()` (pre-stage inputs in HBM) and `*_device` accessors returning torch tensors.
"""
import logging

import numpy as np
import torch

from . import _lib, synth

LOGGER = logging.getLogger(__name__)


class LazyF64(dict):
    """name -> fp64 twin of a float32 result.  The values stay in HBM until somebody asks for them:
    `park(name, tensor)` registers the device tensor, the first lookup copies it to the host."""

    def __init__(self):
        super().__init__()
        self._parked = {}

    def park(self, name, tensor):
        self._parked[name] = (tensor, torch.cuda.current_stream(tensor.device))
        super().pop(name, None)

    def __missing__(self, name):
        t, st = self._parked.pop(name)        # KeyError if unknown, like a dict
        with torch.cuda.stream(st):           # the stream that produced it
            v = _lib.d2h(t)
        self[name] = v
        return v

    def __contains__(self, name):
        return super().__contains__(name) or name in self._parked

    def keys(self):
        return list(super().keys()) + list(self._parked.keys())


class TriMesh(object):
    """Triangular mesh (2D or 3D vertices) with GPU-computed per-vertex properties."""

    KNOWN_ATTRIBUTES = ['pnormal', 'parea', 'mean_curv', 'gaus_curv', 'shell']

    # ---------------------------------------------------------------------- constructor
    def __init__(self, vertices, **kwargs):
        vertices = np.asarray(vertices)
        if vertices.ndim != 2 or vertices.shape[1] not in [2, 3]:
            raise ValueError(f'TriMesh needs 2D or 3D vertices: got {vertices.shape}')
        self.vertices, _ = _lib.staged(vertices, np.float32)       # page-locked float32 inputs are used where they are

        if 'faces' in kwargs and kwargs['faces'] is not None:
            faces = np.asarray(kwargs['faces'])
            if faces.ndim != 2 or faces.shape[1] != 3:
                raise ValueError(f'TriMesh needs triangles: got {faces.shape}')
            self.faces = faces.astype(np.uint32)
        else:
            self.faces = None

        self.periodic = kwargs.get('periodic', False)
        self.label = kwargs.get('label', 'TriMeshPeriodic' if self.periodic else 'TriMesh')
        self.device = torch.device(kwargs.get('device', 'cuda:0'))

        self.bbox = None
        self.bboxl = None
        self._periodic_faces = None
        self._trimmed_faces = None
        self._duplicated_vidxs = None
        self._duplicated_verts = None
        self._parameterized_verts = None
        self._delaunay = None            # (dfaces, doffs) of the periodic Delaunay (host producer)

        self.attributes = _lib.LazyResults()     # numpy on access; downloads may still be in flight
        self.attributes_f64 = LazyF64()   # fp64 twins, fetched from the device on first access
        self._fields = _lib.LazyResults() # KDE fields cached by name (TriMesh_kde.cpp:455-460)
        self._emitted = {}                # name -> (get_nlipids, label selection) of fields written ahead of their request
        self._label_index = None         # optional per-vertex integer labels (label-fused KDE)
        self._nlabels = 0
        self._dev = {}                   # device-resident buffers
        self._stream = None              # home stream (set by the Membrane that drives this mesh)

    def tag(self):
        return f'[{self.label}, periodic={self.periodic}]'

    def __str__(self):
        nf = self.faces.shape[0] if self.faces is not None else 0
        return f'{self.tag()}: {self.vertices.shape[0]} verts, {nf} faces. ' \
               f'bbox = {self.vertices.min(axis=0)}, {self.vertices.max(axis=0)}'

    __repr__ = __str__

    # ---------------------------------------------------------------------- bbox / triangulation (host producer side)
    def set_bbox(self, bb0, bb1):
        bb0, bb1 = np.asarray(bb0), np.asarray(bb1)
        if bb0.shape[0] == 2 and bb1.shape[0] == 2:
            self.bboxl = np.array([bb0[0], bb0[1], bb1[0], bb1[1]])
            self.bbox = np.array([[bb0[0], bb0[1], 0.], [bb1[0], bb1[1], 0.]])
        elif bb0.shape[0] == 3 and bb1.shape[0] == 3:
            self.bboxl = np.array([bb0[0], bb0[1], bb0[2], bb1[0], bb1[1], bb1[2]])
            self.bbox = np.array([[bb0[0], bb0[1], bb0[2]], [bb1[0], bb1[1], bb1[2]]])
        else:
            raise ValueError(f'TriMesh got incorrect bbox: {bb0}, {bb1}')
        self.bbox = self.bbox.astype(np.float32)
        self.bboxl = self.bboxl.astype(np.float32)     # memsurfer.i takes float32 (INPLACE_ARRAY1)
        self._invalidate()

    def _invalidate(self):
        self._dev.clear()

    def set_periodic_triangulation(self, faces, periodic_faces, trimmed_faces, duplicated_vidxs, duplicated_verts):
        """Inject the arrays `_fetch_faces` (trimesh.py:421-436) would fetch from the C++ object."""
        self.faces, _ = _lib.staged(np.asarray(faces).reshape(-1, 3), np.uint32)
        self._periodic_faces = np.ascontiguousarray(periodic_faces, dtype=np.uint32).reshape(-1, 3)
        self._trimmed_faces, _ = _lib.staged(np.asarray(trimmed_faces).reshape(-1, 3), np.uint32)
        self._duplicated_vidxs = np.ascontiguousarray(duplicated_vidxs, dtype=np.uint32)
        self._duplicated_verts, _ = _lib.staged(np.asarray(duplicated_verts).reshape(-1, self.vertices.shape[1]), np.float32)
        self._invalidate()

    def _retrim(self):
        if self.bbox is None:
            raise ValueError(f'{self.tag()} periodic mesh needs a bbox (set_bbox) before triangulation')
        b0, b1 = self.bbox[0, :self.vertices.shape[1]], self.bbox[1, :self.vertices.shape[1]]
        if len(self._delaunay) == 3:                  # device producer: already split into inside / crossing faces
            faces, pf, po = self._delaunay
            if torch.cuda.is_available():      # duplicated-vertex numbering of the crossing faces on the device
                from .delaunay import trim_partial_device
                with torch.cuda.stream(self._stream or torch.cuda.current_stream(self.device)):
                    t = trim_partial_device(self.vertices, pf, po, b0, b1, self.device)
            else:
                t = synth.trim_partial(self.vertices, pf, po, b0, b1)
            t['faces'] = faces
        else:
            dfaces, doffs = self._delaunay
            t = synth.trim_periodic(self.vertices, dfaces, doffs, b0, b1)
        self.set_periodic_triangulation(t['faces'], t['periodic_faces'], t['trimmed_faces'],
                                        t['duplicated_vidxs'], t['duplicated_verts'])

    def set_periodic_delaunay(self, dfaces, doffs):
        """Inject a periodic Delaunay made by another producer (CGAL in the reference, `synth.periodic_delaunay` in the
        CPU tests): dfaces (F,3) original vertex ids, doffs (F,3,2) periodic offsets -- the PeriodicVertex triples of
        TriMesh_cgal.cpp:223-233.  Trimmed with this mesh's bbox like copy_periodicDelaunay (TriMesh.hpp:301-304)."""
        if not self.periodic:
            raise ValueError(f'{self.tag()} is not periodic')
        self._delaunay = (np.asarray(dfaces, np.int64).reshape(-1, 3), np.asarray(doffs, np.int8).reshape(-1, 3, 2))
        self._retrim()

    def compute_delaunay(self):
        """TriMesh::delaunay / periodicDelaunay (memsurfer/src/TriMesh_cgal.cpp:111-281; CGAL in the reference).  Periodic
        meshes are triangulated on the GPU (memsurfer_b200.delaunay: ghost images -> k-NN -> one Voronoi cell per point,
        robust in-circle predicate).  Cells that do not close inside the 0.2-box-width ghost layer (very sparse or
        very uneven point sets) get one more try with a 0.45-box-width layer; after that IncompleteCells is raised --
        there is no host triangulation behind this call.  Non-periodic meshes (not on the per-frame path of the
        reference's drivers, which are periodic) use the planar host producer `synth.planar_delaunay`."""
        if self.faces is not None:
            return
        if not self.periodic:
            self.faces = synth.planar_delaunay(self.vertices[:, :2])
            self._invalidate()
            return
        if self.bbox is None:
            raise ValueError(f'{self.tag()} periodic Delaunay needs a bbox')
        _lib.require_cuda()
        from .delaunay import periodic_delaunay_device, IncompleteCells
        self._delaunay = None
        with torch.cuda.stream(self._stream or torch.cuda.current_stream(self.device)):
            try:
                self._delaunay = periodic_delaunay_device(self.vertices[:, :2], self.bbox[0], self.bbox[1], self.device, split=True)
            except IncompleteCells as e:
                if not e.retry:
                    raise
                LOGGER.info(f'{self.tag()} {e}: once more with a 0.45-box-width ghost layer')
                self._delaunay = periodic_delaunay_device(self.vertices[:, :2], self.bbox[0], self.bbox[1], self.device, split=True,
                                                          blayer=0.45, knbrs=31)
        self._retrim()

    def copy_triangulation(self, mesh):
        if self.periodic != mesh.periodic:
            raise ValueError(f'Cannot copy a (periodic={self.periodic}) mesh into ({mesh.periodic})')
        if not self.periodic:
            self.faces = mesh.faces.copy()
            self._invalidate()
            return
        self._delaunay = mesh._delaunay
        self._retrim()

    def copy_densities(self, mesh):
        """Adopt the cached density fields of another mesh (trimesh.py:125-126 -> TriMesh::set_fields,
        TriMesh.hpp:277-284: every field whose name STARTS with 'density')."""
        for name in list(mesh._fields.keys()):
            if name.find('density') == 0:
                self._fields[name] = mesh._fields[name]
                if name in mesh.attributes_f64:
                    self.attributes_f64[name] = mesh.attributes_f64[name]

    def parameterization_bbox(self):
        """(min, max) of the parameterised vertices (trimesh.py:131-134)."""
        bb0 = self._parameterized_verts.min(axis=0)
        bb1 = self._parameterized_verts.max(axis=0)
        return bb0, bb1

    # ---------------------------------------------------------------------- device staging
    def _concat_host(self):
        """verts ++ dup verts (z = 0 for planar meshes) and faces ++ trimmed faces, page-locked."""
        include_periodic = self.periodic and (self._duplicated_vidxs is not None)
        V, dim = self.vertices.shape
        Vd = self._duplicated_verts.shape[0] if include_periodic else 0
        F0 = self.faces.shape[0]
        Ft = self._trimmed_faces.shape[0] if include_periodic else 0
        verts = _lib.pinned_empty((V + Vd, 3), np.float32)
        faces = _lib.pinned_empty((F0 + Ft, 3), np.uint32)
        if dim == 2:
            verts[:, 2] = 0.0
        verts[:V, :dim] = self.vertices
        faces[:F0] = self.faces
        if include_periodic:
            verts[V:, :dim] = self._duplicated_verts
            faces[F0:] = self._trimmed_faces
        return verts, faces

    @_lib.home_stream
    def to_device(self, pinned=None):
        """Stage verts ++ dup verts and faces ++ trimmed faces in HBM (TriMesh.hpp:360-364 layout).  The pieces
        are uploaded straight into slices of the device arrays (no concatenated host copy)."""
        _lib.require_cuda()
        if 'verts' in self._dev:
            return self._dev
        if self.faces is None:
            raise ValueError(f'{self.tag()} has no faces (call compute_delaunay or pass faces=)')
        include_periodic = self.periodic and (self._duplicated_vidxs is not None)
        V, dim = self.vertices.shape
        d = self._dev
        if dim != 3:
            verts, faces = self._concat_host()
            d['verts'] = _lib.h2d(verts, self.device)
            d['faces'] = _lib.h2d(faces.view(np.int32), self.device)
            d['_host'] = (verts, faces)          # keep the page-locked staging alive until the copy is done
        else:
            Vd = self._duplicated_verts.shape[0] if include_periodic else 0
            F0 = self.faces.shape[0]
            Ft = self._trimmed_faces.shape[0] if include_periodic else 0
            host = [self.vertices, self.faces]
            with _lib.uploading(self.device) as keep:          # on the upload stream: under the kernels already queued
                dv = keep(torch.empty((V + Vd, 3), dtype=torch.float32, device=self.device))
                df = keep(torch.empty((F0 + Ft, 3), dtype=torch.int32, device=self.device))
                dv[:V].copy_(torch.from_numpy(self.vertices), non_blocking=True)
                df[:F0].copy_(torch.from_numpy(self.faces.view(np.int32)), non_blocking=True)
                if Vd:
                    dv[V:].copy_(torch.from_numpy(self._duplicated_verts), non_blocking=True)
                    host.append(self._duplicated_verts)
                if Ft:
                    df[F0:].copy_(torch.from_numpy(self._trimmed_faces.view(np.int32)), non_blocking=True)
                    host.append(self._trimmed_faces)
            d['verts'], d['faces'], d['_host'] = dv, df, tuple(host)
        d['nv'] = V
        return d

    def _half_box(self):
        include_periodic = self.periodic and (self._duplicated_vidxs is not None)
        if not include_periodic:
            return -1.0, -1.0
        # trimesh.py:484-485 : float32 arithmetic on the float32 bbox
        xw = np.float32(0.5) * (self.bbox[1, 0] - self.bbox[0, 0])
        yw = np.float32(0.5) * (self.bbox[1, 1] - self.bbox[0, 1])
        return float(xw), float(yw)

    @_lib.home_stream
    def _topology_device(self):
        d = self.to_device()
        if 'csr_off' in d:
            return d
        L = _lib.lib()
        nv_all, nv, nf = d['verts'].shape[0], d['nv'], d['faces'].shape[0]
        dev = self.device
        d['csr_off'] = torch.empty(nv + 1, dtype=torch.int32, device=dev)
        d['csr_adj'] = torch.empty(max(3 * nf, 1), dtype=torch.int32, device=dev)
        d['keep'] = torch.empty(max(nf, 1), dtype=torch.uint8, device=dev)
        ws = _lib.workspace(L.ms_mesh_topology_workspace_bytes(nv, nf), dev)
        xw, yw = self._half_box()
        _lib.check(L.ms_mesh_topology(d['verts'].data_ptr(), nv_all, nv, d['faces'].data_ptr(), nf, xw, yw,
                                      d['csr_off'].data_ptr(), d['csr_adj'].data_ptr(), d['keep'].data_ptr(),
                                      ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)), 'ms_mesh_topology')
        return d

    @_lib.home_stream
    def properties_device(self):
        """One fused launch: vertex normals, point areas, mean and Gaussian curvature (device tensors)."""
        d = self._topology_device()
        if 'vn64' in d:
            return d
        L = _lib.lib()
        nv_all, nv, nf = d['verts'].shape[0], d['nv'], d['faces'].shape[0]
        dev = self.device
        d['vn64'] = torch.empty((nv, 3), dtype=torch.float64, device=dev)
        d['vn32'] = torch.empty((nv, 3), dtype=torch.float32, device=dev)
        d['pa64'] = torch.empty(nv, dtype=torch.float64, device=dev)
        d['pa32'] = torch.empty(nv, dtype=torch.float32, device=dev)
        d['H'] = torch.empty(nv, dtype=torch.float64, device=dev)
        d['K'] = torch.empty(nv, dtype=torch.float64, device=dev)
        ws = _lib.workspace(L.ms_mesh_properties_workspace_bytes(nv, nf), dev)
        _lib.check(L.ms_mesh_properties(d['verts'].data_ptr(), nv_all, nv, d['faces'].data_ptr(), nf,
                                        d['csr_off'].data_ptr(), d['csr_adj'].data_ptr(), d['keep'].data_ptr(),
                                        d['vn64'].data_ptr(), d['vn32'].data_ptr(), d['pa64'].data_ptr(),
                                        d['pa32'].data_ptr(), d['H'].data_ptr(), d['K'].data_ptr(),
                                        ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)), 'ms_mesh_properties')
        return d

    # ---------------------------------------------------------------------- computation of attributes
    @_lib.home_stream
    def _properties_lazy(self):
        """Queue the fused properties launch and the four downloads; nothing is waited for here."""
        if 'pnormal' in self.attributes:
            return
        d = self.properties_device()
        for key, t32, t64 in (('pnormal', d['vn32'], d['vn64']), ('parea', d['pa32'], d['pa64']),
                              ('mean_curv', d['H'], d['H']), ('gaus_curv', d['K'], d['K'])):
            self.attributes.park(key, t32)
            self.attributes_f64.park(key, t64)

    def compute_normals(self):
        """(V,3) float32 vertex normals (trimesh.py:244-261 -> TriMesh.cpp:99-133)."""
        self._properties_lazy()
        return self.attributes['pnormal']

    def compute_pointareas(self):
        """(V,) float32 mixed Voronoi vertex areas (trimesh.py:264-281 -> TriMesh.cpp:140-190)."""
        self._properties_lazy()
        return self.attributes['parea']

    def compute_curvatures(self):
        """((V,) f64 mean, (V,) f64 Gaussian) as vtkCurvatures gives them (trimesh.py:284-316)."""
        self._properties_lazy()
        return self.attributes['mean_curv'], self.attributes['gaus_curv']

    # ---------------------------------------------------------------------- closest point / thickness
    @_lib.home_stream
    def closest_point_device(self, queries_dev):
        """Closest point of THIS mesh's own faces to every query (device tensors in, device tensors out):
        (face id i32 (n,), barycentric f64 (n,3), distance f64 (n,), distance f32 (n,))."""
        _lib.require_cuda()
        L = _lib.lib()
        d = self.to_device()
        dev = self.device
        V, F0 = self.vertices.shape[0], self.faces.shape[0]
        if self.vertices.shape[1] != 3:
            raise ValueError(f'{self.tag()} project_on_surface() requires a mesh in 3D!')
        nq = int(queries_dev.shape[0])
        fid = torch.empty(nq, dtype=torch.int32, device=dev)
        bary = torch.empty((nq, 3), dtype=torch.float64, device=dev)
        d64 = torch.empty(nq, dtype=torch.float64, device=dev)
        d32 = torch.empty(nq, dtype=torch.float32, device=dev)
        lo, hi = self.vertices.min(axis=0), self.vertices.max(axis=0)
        box4 = (_lib.C.c_float * 4)(float(lo[0]), float(lo[1]), float(hi[0]), float(hi[1]))
        ws = self._workspace('proj_ws', L.ms_closest_point_workspace_bytes(V, F0))
        _lib.check(L.ms_closest_point(d['verts'].data_ptr(), V, d['faces'].data_ptr(), F0, box4,
                                      queries_dev.data_ptr(), nq, fid.data_ptr(), bary.data_ptr(), d64.data_ptr(),
                                      d32.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)), 'ms_closest_point')
        return fid, bary, d64, d32

    @_lib.home_stream
    def project_on_surface(self, points):
        """(face ids (n,) int, barycentric coordinates (n,3) f64) of the closest surface point of every
        point (TriMesh::project_on_surface, TriMesh_cgal.cpp:295-363, without the float packing)."""
        points = np.asarray(points)
        if points.ndim != 2 or points.shape[1] != 3:
            raise ValueError(f'{self.tag()} project_on_surface() needs 3D points: got {points.shape}')
        q = _lib.h2d(_lib.pinned_copy(points, np.float32), self.device)
        fid, bary, _, _ = self.closest_point_device(q)
        return _lib.d2h(fid).astype(int), _lib.d2h(bary)

    def set_parameterization(self, pverts):
        """Inject the producer's planar parameterisation (the reference computes it with CGAL, trimesh.py:129-156)."""
        pverts = np.asarray(pverts, np.float32).reshape(-1, 2)
        if pverts.shape[0] != self.vertices.shape[0]:
            raise ValueError(f'{self.tag()} expects one parameterized vertex per vertex')
        self._parameterized_verts = pverts

    def parameterize(self, xy=False):
        """Planar parameterisation of the vertices (trimesh.py:136-156; `xy=True` is TriMesh::parameterize_xy, the
        projection on the xy plane).  The CGAL producer behind `xy=False` is a host step that is not available
        here: unless a parameterisation was injected (`set_parameterization`), the xy projection is used too."""
        if xy or self._parameterized_verts is None:
            if not xy:
                LOGGER.info(f'{self.tag()} no injected parameterization: using the xy projection')
            self._parameterized_verts = np.ascontiguousarray(self.vertices[:, :2], np.float32)
        return self._parameterized_verts

    def project_on_surface_and_plane(self, points):
        """Project points on the triangulation and on its parameterisation (trimesh.py:158-190): returns
        (spoints (n,3) f32, ppoints (n,2) f32).  The per-point Python loop of the reference is three
        vectorised accumulations in the same order and dtypes."""
        self.parameterize()
        fids, fbarys = self.project_on_surface(points)
        npoints = fids.shape[0]
        spoints = np.zeros((npoints, 3), dtype=np.float32)
        ppoints = np.zeros((npoints, 2), dtype=np.float32)
        f = self.faces[fids].astype(np.int64)
        for j in range(3):
            spoints += fbarys[:, j, None] * self.vertices[f[:, j]]
            ppoints += fbarys[:, j, None] * self._parameterized_verts[f[:, j]]
        return spoints, ppoints

    def compute_distance_to_surface(self, other):
        """(V,) float32 distance of every vertex of this mesh to the other mesh's surface (trimesh.py:319-321 ->
        TriMesh::distance_to_other_mesh, TriMesh_cgal.cpp:390-425)."""
        d = self.to_device()
        q = d['verts'][:self.vertices.shape[0]]
        if other.device != self.device:
            q = q.to(other.device)
        if self._stream is not None and other._stream is not self._stream:     # the queries were staged on this mesh's stream
            (other._stream or torch.cuda.current_stream(other.device)).wait_stream(self._stream)
        _, _, d64, d32 = other.closest_point_device(q)
        with torch.cuda.stream(other._stream or torch.cuda.current_stream(other.device)):    # where the result was produced
            self.attributes_f64.park('distance_to_surface', d64)
            return _lib.d2h(d32)

    # ---------------------------------------------------------------------- vertex neighbours / shells
    @_lib.home_stream
    def neighbors_device(self):
        """(nbr_off (V+1,) i32, nbr (total,) i32) device tensors: TriMesh::need_neighbors (TriMesh.cpp:197-255) on
        the mesh's own faces, lists in the reference's order."""
        _lib.require_cuda()
        if self.faces is None:
            raise ValueError(f'{self.tag()} has no faces (call compute_delaunay or pass faces=)')
        d = self._dev
        if 'nbr' not in d:
            L = _lib.lib()
            dev = self.device
            V, F = self.vertices.shape[0], self.faces.shape[0]
            faces = _lib.h2d(np.ascontiguousarray(self.faces, np.uint32).view(np.int32), dev)
            off = torch.empty(V + 1, dtype=torch.int32, device=dev)
            nbr = torch.empty(max(1, 6 * F), dtype=torch.int32, device=dev)
            ws = _lib.workspace(L.ms_mesh_neighbors_workspace_bytes(V, F), dev)
            total = _lib.C.c_int64(0)
            _lib.check(L.ms_mesh_neighbors(faces.data_ptr(), F, V, off.data_ptr(), nbr.data_ptr(), _lib.C.byref(total),
                                           ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)), 'ms_mesh_neighbors')
            d['nbr'] = (off, nbr[:int(total.value)].clone())
        return d['nbr']

    def need_neighbors(self):
        """Linearised neighbour lists as TriMesh::need_neighbors returns them (TriMesh.cpp:241-254): the first V
        values are the list lengths, followed by the lists."""
        off, nbr = self.neighbors_device()
        with torch.cuda.stream(self._stream or torch.cuda.current_stream(self.device)):
            off_h, nbr_h = _lib.d2h(off), _lib.d2h(nbr)
        return np.concatenate([np.diff(off_h), nbr_h]).astype(np.int32)

    @_lib.home_stream
    def compute_shells(self, ref_pts):
        """(V,) int shell index = hop distance to the nearest reference vertex (trimesh.py:364-416)."""
        key = 'shell'
        if key in self.attributes:
            return self.attributes[key]
        assert isinstance(ref_pts, (list, np.ndarray))
        assert len(ref_pts) > 0
        L = _lib.lib()
        dev = self.device
        V = self.vertices.shape[0]
        off, nbr = self.neighbors_device()
        ref = _lib.h2d(_lib.pinned_copy(np.asarray(ref_pts).reshape(-1), np.int32), dev)
        shells = torch.empty(V, dtype=torch.int32, device=dev)
        ws = _lib.workspace(L.ms_mesh_shells_workspace_bytes(V), dev)
        bad, top = _lib.C.c_int64(0), _lib.C.c_int(0)
        _lib.check(L.ms_mesh_shells(off.data_ptr(), nbr.data_ptr(), V, ref.data_ptr(), int(ref.numel()), shells.data_ptr(),
                                    _lib.C.byref(bad), _lib.C.byref(top), ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)),
                   'ms_mesh_shells')
        out = _lib.d2h(shells)
        if bad.value > 0:
            wrong = np.where(out < 0)[0]
            LOGGER.error(f'Failed to assign shell id for {wrong.shape[0]} vertices: '
                         f'ids = {wrong}, verts = {self.vertices[wrong]}')
            assert out.min() == 0, 'Failed to assign shell id for some vertices'
        self.attributes[key] = out.astype(int)
        return self.attributes[key]

    # ---------------------------------------------------------------------- KDE
    def _kde_box(self, cutoff_sigmas):
        if self.periodic:
            if self.bboxl is None:
                raise ValueError(f'{self.tag()} periodic density needs a bbox')
            b = np.asarray(self.bboxl, np.float32)
            if b.shape[0] == 4:
                b = np.array([b[0], b[1], 0, b[2], b[3], 0], np.float32)
            return _lib.box6(b)
        if cutoff_sigmas > 0:
            # non-periodic cut-off: the cell list only needs the extent of the vertices (two passes over them: kept)
            if getattr(self, '_extent6', None) is None:
                lo, hi = self.vertices.min(axis=0), self.vertices.max(axis=0)
                self._extent6 = np.array([lo[0], lo[1], 0, hi[0], hi[1], 0], np.float32)
            return _lib.box6(self._extent6)
        return None

    def _kde_verts_device(self):
        d = self._dev
        if 'kverts' not in d:
            dev = self.device
            if 'verts' in d:
                d['kverts'] = d['verts'][:self.vertices.shape[0]]      # already resident (first V rows)
            elif self.vertices.shape[1] == 3:
                d['kverts'] = _lib.h2d(self.vertices, dev)
            else:
                v = _lib.pinned_empty((self.vertices.shape[0], 3), np.float32)
                v[:, :2] = self.vertices; v[:, 2] = 0.0
                d['kverts'] = _lib.h2d(v, dev)
                d['_khost'] = v
        return d['kverts']

    def _workspace(self, key, nbytes):
        ws = self._dev.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = self._dev[key] = _lib.workspace(nbytes, self.device)
        return ws

    @_lib.home_stream
    def geodesics_device(self):
        """(pitch, pitch) f64 device tensor of graph distances along the mesh edges (TriMesh::compute_geodesics_fw,
        TriMesh_kde.cpp:196-332; periodic meshes: faces ++ periodic faces, :487-489).  Computed once per mesh."""
        _lib.require_cuda()
        d = self._dev
        if 'geodesics' not in d:
            L = _lib.lib()
            dev = self.device
            if self.faces is None:
                raise ValueError(f'{self.tag()} has no faces (call compute_delaunay or pass faces=)')
            faces = np.ascontiguousarray(self.faces, np.uint32)
            if self.periodic and self._periodic_faces is not None and self._periodic_faces.shape[0] > 0:
                faces = np.concatenate([faces, np.ascontiguousarray(self._periodic_faces, np.uint32)])
            V = self.vertices.shape[0]
            pitch = int(L.ms_geodesic_pitch(V))
            D = torch.empty((pitch, pitch), dtype=torch.float64, device=dev)
            fd = _lib.h2d(faces.view(np.int32), dev)
            _lib.check(L.ms_geodesics(self._kde_verts_device().data_ptr(), V, fd.data_ptr(), faces.shape[0],
                                      int(bool(self.periodic)), self._kde_box(0.0), D.data_ptr(), _lib.stream_ptr(dev)),
                       'ms_geodesics')
            d['geodesics'] = D
        return d['geodesics']

    @_lib.home_stream
    def density_geodesic_device(self, sigma, get_nlipids, pidxs_dev=None):
        """Type-1 density on device-resident inputs; returns (f64 tensor, f32 tensor)."""
        L = _lib.lib()
        dev = self.device
        D = self.geodesics_device()
        nv = self.vertices.shape[0]
        nids = 0 if pidxs_dev is None else int(pidxs_dev.numel())
        o64 = torch.empty(nv, dtype=torch.float64, device=dev)
        o32 = torch.empty(nv, dtype=torch.float32, device=dev)
        ws = self._workspace('kde_geo_ws', L.ms_kde_geodesic_workspace_bytes(nv))
        _lib.check(L.ms_kde_geodesic(D.data_ptr(), nv, _lib.ptr(pidxs_dev) if nids else None, nids, float(sigma),
                                     int(bool(get_nlipids)), o64.data_ptr(), o32.data_ptr(), ws.data_ptr(), ws.numel(),
                                     _lib.stream_ptr(dev)), 'ms_kde_geodesic')
        return o64, o32

    @_lib.home_stream
    def density_device(self, type, sigma, get_nlipids, pidxs_dev=None, cutoff_sigmas=0.0):
        """Run ms_kde on device-resident inputs; returns (f64 tensor, f32 tensor)."""
        _lib.require_cuda()
        L = _lib.lib()
        dev = self.device
        kverts = self._kde_verts_device()
        nv = kverts.shape[0]
        nids = 0 if pidxs_dev is None else int(pidxs_dev.numel())
        o64 = torch.empty(nv, dtype=torch.float64, device=dev)
        o32 = torch.empty(nv, dtype=torch.float32, device=dev)
        ws = self._workspace('kde_ws', L.ms_kde_workspace_bytes(nv, nids))
        _lib.check(L.ms_kde(kverts.data_ptr(), nv, _lib.ptr(pidxs_dev) if nids else None, nids,
                            int(type), int(bool(self.periodic)), self._kde_box(cutoff_sigmas), float(sigma),
                            int(bool(get_nlipids)), float(cutoff_sigmas), o64.data_ptr(), o32.data_ptr(),
                            ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)), 'ms_kde')
        return o64, o32

    KDE_RAW_CACHE_BYTES = 4 << 30      # raw per-label pair sums kept per mesh (oldest (type, sigma, cut-off) evicted first)

    def _kde_raw_store(self, keys, block):
        """Cache the raw per-label sums of one ms_kde_labels(_multi) call: block (nsigma, nlabels, nv); keys[i] = the
        (type, sigma, cut-off) of sigma i.  Every key maps to (block holder, i); the holder also keeps the per-row
        totals once they are needed.  Oldest blocks are evicted first when the cache outgrows its byte budget."""
        cache = self._dev.setdefault('kde_raw', {})
        holder = {'raw': block, 'totals': None}
        for i, k in enumerate(keys):
            cache[k] = (holder, i)
        seen, total = [], 0
        for k, (h, _) in cache.items():
            if not any(h is x for x in seen):
                seen.append(h); total += h['raw'].numel() * 8
        for h in seen:
            if total <= self.KDE_RAW_CACHE_BYTES or h is holder:
                continue
            total -= h['raw'].numel() * 8
            for k in [k for k, (hh, _) in cache.items() if hh is h]:
                del cache[k]
        return holder

    def _kde_totals(self, holder):
        """(nsigma * nlabels, 256) partial sums of every row of the block (one launch, cached with the block)."""
        if holder['totals'] is None:
            L = _lib.lib()
            raw = holder['raw']
            nrows, nv = raw.shape[0] * raw.shape[1], raw.shape[2]
            tot = torch.empty(int(L.ms_kde_label_totals_bytes(nrows)) // 8, dtype=torch.float64, device=self.device)
            _lib.check(L.ms_kde_label_totals(raw.data_ptr(), nrows, nv, tot.data_ptr(), _lib.stream_ptr(self.device)), 'ms_kde_label_totals')
            holder['totals'] = tot
        return holder['totals']

    def clear_density_cache(self):
        """Drop the raw per-label pair sums kept for later compute_density_labels calls (device memory)."""
        self._dev.pop('kde_raw', None)

    def set_vertex_labels(self, label_index, nlabels, counts=None, index32=None):
        """Integer label in [0, nlabels) per vertex (negative: never a KDE source).  Enables the label-fused
        KDE: one pass evaluates every vertex pair once and keeps the sums apart by the source's label.
        index32: the same index as a ready int32 array (page-locked if it is to be uploaded by DMA); used as it is."""
        label_index = np.asarray(label_index)
        if label_index.shape != (self.vertices.shape[0],):
            raise ValueError(f'{self.tag()} expects one label per vertex')
        if index32 is not None and index32.shape == label_index.shape and index32.dtype == np.int32:
            self._label_index = index32
        else:
            self._label_index = _lib.pinned_copy(label_index, np.int32)
        self._nlabels = int(nlabels)
        self._label_counts = np.asarray(counts) if counts is not None else \
            np.bincount(label_index[label_index >= 0], minlength=self._nlabels)
        self._dev.pop('labels', None)
        self._dev.pop('kde_raw', None)

    @_lib.home_stream
    def density_labels_device(self, type, sigma, cutoff_sigmas=0.0):
        """(nlabels, V) f64 device tensor of raw per-label pair sums (cached per (type, sigma, cut-off))."""
        _lib.require_cuda()
        L = _lib.lib()
        dev = self.device
        d = self._dev
        cache = d.setdefault('kde_raw', {})
        key = (int(type), float(sigma), float(cutoff_sigmas))
        if key in cache:
            h, i = cache[key]
            return h['raw'][i]
        kverts = self._kde_verts_device()
        nv = kverts.shape[0]
        if 'labels' not in d:
            d['labels'] = _lib.h2d(self._label_index, dev)
        raw = torch.empty((1, self._nlabels, nv), dtype=torch.float64, device=dev)
        ws = self._workspace('kde_ws', L.ms_kde_labels_workspace_bytes(nv, self._nlabels))
        _lib.check(L.ms_kde_labels(kverts.data_ptr(), nv, d['labels'].data_ptr(), self._nlabels, int(type),
                                   int(bool(self.periodic)), self._kde_box(cutoff_sigmas), float(sigma),
                                   float(cutoff_sigmas), raw.data_ptr(), ws.data_ptr(), ws.numel(),
                                   _lib.stream_ptr(dev)), 'ms_kde_labels')
        self._kde_raw_store([key], raw)
        return raw[0]

    @_lib.home_stream
    def density_labels_prefetch(self, type, sigmas, cutoff_sigmas=0.0, after=None, emit=None):
        """Raw per-label sums of SEVERAL sigmas in one ms_kde_labels_multi call (the vertices are binned once);
        fills the cache density_labels_device reads.  No-op (returns False) for sigmas that are already cached.
        `after`: a CUDA event the launches wait for.
        `emit` = (sets, get_nlipids, namer): also write, sigma by sigma, the normalised fields of the label sets
        `sets` (lists of label indices) under the names namer(sigma, k) and start their downloads inside the call
        (ms_kde_labels_multi_fields): a field that is asked for later by that name, with the same selection and the
        same get_nlipids, is already on its way to the host; any other request recomputes it."""
        _lib.require_cuda()
        L = _lib.lib()
        dev = self.device
        d = self._dev
        cache = d.setdefault('kde_raw', {})
        todo = [float(s) for s in sigmas if (int(type), float(s), float(cutoff_sigmas)) not in cache]
        if len(todo) < 2 or self._label_index is None:
            return False
        if after is not None:                 # see Membrane.compute_densities: one leaflet's pair sums after the other's
            torch.cuda.current_stream(dev).wait_event(after)
        kverts = self._kde_verts_device()
        nv = kverts.shape[0]
        if 'labels' not in d:
            d['labels'] = _lib.h2d(self._label_index, dev)
        raw = torch.empty((len(todo), self._nlabels, nv), dtype=torch.float64, device=dev)
        ws = self._workspace('kde_ws', L.ms_kde_labels_multi_workspace_bytes(nv, self._nlabels))
        sg = (_lib.C.c_float * len(todo))(*todo)
        if emit is not None and len(emit[0]) * len(todo) <= 64:
            sets, get_nlipids, namer = emit
            C = _lib.C
            ns, nsets = len(todo), len(sets)
            masks, nsels = [], []
            for sel_ in sets:
                masks.append(int(sum(1 << int(l) for l in sel_)))
                nsels.append(float(self._label_counts[np.asarray(sel_, dtype=np.int64)].sum()))
            o32 = torch.empty((ns, nsets, nv), dtype=torch.float32, device=dev)
            o64 = torch.empty((ns, nsets, nv), dtype=torch.float64, device=dev)
            _lib.release_finished()
            host = torch.empty((ns, nsets, nv), dtype=torch.float32, pin_memory=True)
            tot = torch.empty(int(L.ms_kde_label_totals_bytes(ns * self._nlabels)) // 8, dtype=torch.float64, device=dev)
            cs = _lib.copy_stream(dev)
            _lib.check(L.ms_kde_labels_multi_fields(kverts.data_ptr(), nv, d['labels'].data_ptr(), self._nlabels, int(type),
                                                    int(bool(self.periodic)), self._kde_box(cutoff_sigmas), sg, ns, float(cutoff_sigmas),
                                                    raw.data_ptr(), nsets, (C.c_uint64 * nsets)(*masks), (C.c_double * nsets)(*nsels),
                                                    int(bool(get_nlipids)), tot.data_ptr(), o64.data_ptr(), o32.data_ptr(),
                                                    host.data_ptr(), cs.cuda_stream, ws.data_ptr(), ws.numel(),
                                                    _lib.stream_ptr(dev)), 'ms_kde_labels_multi_fields')
            ev = torch.cuda.Event()
            ev.record(cs)                       # after the last block's copy
            o32.record_stream(cs)
            _lib.keep_until(ev, host)           # the copies were issued by the library: keep the block out of torch's cache till then
            for i, s_ in enumerate(todo):
                for k, sel_ in enumerate(sets):
                    n_ = namer(s_, k)
                    self._fields.adopt(n_, _lib.Pending(host[i, k], ev))
                    self.attributes_f64.park(n_, o64[i, k])
                    self._emitted[n_] = (bool(get_nlipids), tuple(int(l) for l in sel_))
            h = self._kde_raw_store([(int(type), s_, float(cutoff_sigmas)) for s_ in todo], raw)
            h['totals'] = tot                   # the per-row partial sums of this block are already there
        else:
            _lib.check(L.ms_kde_labels_multi(kverts.data_ptr(), nv, d['labels'].data_ptr(), self._nlabels, int(type),
                                             int(bool(self.periodic)), self._kde_box(cutoff_sigmas), sg, len(todo),
                                             float(cutoff_sigmas), raw.data_ptr(), ws.data_ptr(), ws.numel(),
                                             _lib.stream_ptr(dev)), 'ms_kde_labels_multi')
            self._kde_raw_store([(int(type), s_, float(cutoff_sigmas)) for s_ in todo], raw)
        return True

    @_lib.home_stream
    def density_select_device(self, raw, select, get_nlipids):
        """Normalised density of the union of the selected labels from the raw per-label sums."""
        L = _lib.lib()
        dev = self.device
        nv = raw.shape[1]
        sel = np.zeros(self._nlabels, np.uint8)
        sel[np.asarray(select, dtype=np.int64)] = 1
        nsel = float(self._label_counts[sel.astype(bool)].sum())
        o64 = torch.empty(nv, dtype=torch.float64, device=dev)
        o32 = torch.empty(nv, dtype=torch.float32, device=dev)
        ws = self._workspace('kde_sel_ws', L.ms_kde_select_workspace_bytes(nv))
        _lib.check(L.ms_kde_select(raw.data_ptr(), nv, self._nlabels, sel.tobytes(), nsel, int(bool(get_nlipids)),
                                   o64.data_ptr(), o32.data_ptr(), ws.data_ptr(), ws.numel(),
                                   _lib.stream_ptr(dev)), 'ms_kde_select')
        return o64, o32

    def _drop_stale_emitted(self, name, get_nlipids, select):
        """A field written ahead of its request (density_labels_prefetch(emit=...)) only stands for a request with the
        same normalisation and the same label selection; anything else is computed afresh, like the reference would."""
        e = self._emitted.get(name)
        if e is not None and e != (bool(get_nlipids), tuple(int(l) for l in np.asarray(select).reshape(-1))):
            self._emitted.pop(name, None)
            if name in self._fields:
                del self._fields[name]

    @_lib.home_stream
    def compute_density_labels_multi(self, type, sigmas, names, get_nlipids, select, cutoff_sigmas=0.0):
        """compute_density_labels for several sigmas of one label set at once: the fields that are not cached yet are
        normalised and written by ONE ms_kde_fields launch per block of raw sums (all sigmas of a prefetch share one).
        Returns the stored values (downloads possibly still in flight) in the order of `names`."""
        if type not in (2, 3):
            return [self.compute_density_labels(type, s_, n_, get_nlipids, select, cutoff_sigmas, lazy=True) for s_, n_ in zip(sigmas, names)]
        L = _lib.lib()
        dev = self.device
        sel = np.zeros(self._nlabels, bool)
        sel[np.asarray(select, dtype=np.int64)] = True
        mask = int(sum(1 << int(l) for l in np.flatnonzero(sel)))
        nsel = float(self._label_counts[sel].sum())
        for n_ in names:
            self._drop_stale_emitted(n_, get_nlipids, select)
        todo = [(float(s_), n_) for s_, n_ in zip(sigmas, names) if n_ not in self._fields]
        by_block = {}
        for s_, n_ in todo:
            self.density_labels_device(type, s_, cutoff_sigmas)             # makes sure the raw sums exist
            h, i = self._dev['kde_raw'][(int(type), s_, float(cutoff_sigmas))]
            by_block.setdefault(id(h), (h, []))[1].append((i, n_))
        for h, items in by_block.values():
            raw = h['raw']
            nv, nf = raw.shape[2], len(items)
            tot = self._kde_totals(h)
            row0 = (_lib.C.c_int32 * nf)(*[i * self._nlabels for i, _ in items])
            masks = (_lib.C.c_uint64 * nf)(*([mask] * nf))
            nsels = (_lib.C.c_double * nf)(*([nsel] * nf))
            o64 = torch.empty((nf, nv), dtype=torch.float64, device=dev)
            o32 = torch.empty((nf, nv), dtype=torch.float32, device=dev)
            _lib.check(L.ms_kde_fields(raw.data_ptr(), tot.data_ptr(), nv, self._nlabels, nf, row0, masks, nsels,
                                       int(bool(get_nlipids)), o64.data_ptr(), o32.data_ptr(), _lib.stream_ptr(dev)), 'ms_kde_fields')
            rows = _lib.d2h_async_rows(o32)             # one download for the block, a view per field
            for k, (_, n_) in enumerate(items):
                self._fields.adopt(n_, rows[k])
                self.attributes_f64.park(n_, o64[k])
        return [self._fields.raw(n_) for n_ in names]

    @_lib.home_stream
    def compute_density_labels(self, type, sigma, name, get_nlipids, select, cutoff_sigmas=0.0, lazy=False):
        """Label-fused twin of compute_density: `select` = label indices (set_vertex_labels) whose vertices
        are the sources.  Same result as compute_density with the ids of those vertices."""
        if type < 1 or type > 3:
            raise ValueError('Invalid density type, {}. Should be 1 (geodesic), 2 (2D) or 3 (3D)'.format(type))
        if type == 1:                 # geodesic: no label-fused kernel, the ids of the selected labels go to kde_2m
            sel = np.zeros(self._nlabels, bool); sel[np.asarray(select, dtype=np.int64)] = True
            pidxs = np.empty([0]) if sel.all() else np.where(sel[np.maximum(self._label_index, 0)] & (self._label_index >= 0))[0]
            return self.compute_density(1, sigma, name, get_nlipids, pidxs, cutoff_sigmas, lazy=lazy)
        self._drop_stale_emitted(name, get_nlipids, select)
        if name not in self._fields:
            raw = self.density_labels_device(type, sigma, cutoff_sigmas)
            o64, o32 = self.density_select_device(raw, select, get_nlipids)
            self._fields.park(name, o32)
            self.attributes_f64.park(name, o64)
            if LOGGER.isEnabledFor(logging.INFO):
                dens = self._fields[name]
                LOGGER.info(f'{self.tag()} density [{name}] range = [{dens.min()}, {dens.max()}], sum = {dens.sum()}')
        if lazy:                  # hand out the download in flight (or the finished array)
            return self._fields.raw(name)
        return self._fields[name]

    @_lib.home_stream
    def compute_density(self, type, sigma, name, get_nlipids, pidxs, cutoff_sigmas=0.0, lazy=False):
        """(V,) float32 density (trimesh.py:324-361 -> TriMesh::kde TriMesh_kde.cpp:443-532).

        type 2: 2-D Gaussian on (x,y); type 3: 3-D Gaussian; type 1: Gaussian of the graph distance along the
        mesh edges (Floyd-Warshall, TriMesh_kde.cpp:196-332 -- O(V^3) work and a V x V matrix: small meshes).
        """
        if type < 1 or type > 3:
            raise ValueError('Invalid density type, {}. Should be 1 (geodesic), 2 (2D) or 3 (3D)'.format(type))
        if name not in self._fields:                  # TriMesh_kde.cpp:455-460: name-keyed cache
            pidxs = np.asarray(pidxs)
            pd = None
            if pidxs.shape[0] > 0:
                pd = _lib.h2d(_lib.pinned_copy(pidxs, np.int32), self.device)
            if type == 1:                             # geodesic: Floyd-Warshall distances + kde_2m
                o64, o32 = self.density_geodesic_device(sigma, get_nlipids, pd)
            else:
                o64, o32 = self.density_device(type, sigma, get_nlipids, pd, cutoff_sigmas)
            self._fields.park(name, o32)
            self.attributes_f64.park(name, o64)
            if LOGGER.isEnabledFor(logging.INFO):
                dens = self._fields[name]
                LOGGER.info(f'{self.tag()} density [{name}] range = [{dens.min()}, {dens.max()}], sum = {dens.sum()}')
        if lazy:
            return self._fields.raw(name)
        return self._fields[name]

    # ---------------------------------------------------------------------- export (host)
    def as_vtkpolydata(self, include_attributes=KNOWN_ATTRIBUTES, additional_attributes={}):
        """Content of the vtkPolyData of trimesh.py:441-567 as an `export.PolyData` (VTK is not required):
        vertices + duplicated vertices, faces + trimmed faces minus the ones spanning the box, attributes
        sorted into point / cell / field data by their length."""
        from . import export
        return export.as_polydata(self, include_attributes, additional_attributes)

    def as_pddataframe(self, include_attributes=KNOWN_ATTRIBUTES, additional_attributes={}):
        """pandas.DataFrame of positions and per-vertex attributes (trimesh.py:570-617)."""
        from . import export
        return export.as_pddataframe(self, include_attributes, additional_attributes)

    def write_vtp(self, filename, additional_attrs={}, ascii=False):
        """VTK XML PolyData file (trimesh.py:620-628), written without VTK."""
        from . import export
        export.write_vtp(self.as_vtkpolydata(TriMesh.KNOWN_ATTRIBUTES, additional_attrs), filename, ascii=ascii)
        LOGGER.info(f'{self.tag()} Written to ({filename})')

    def write_pd(self, filename, additional_attrs={}):
        """CSV through pandas (trimesh.py:630-635)."""
        self.as_pddataframe(TriMesh.KNOWN_ATTRIBUTES, additional_attrs).to_csv(filename)
        LOGGER.info(f'{self.tag()} Written to ({filename})')
