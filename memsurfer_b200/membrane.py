"""memsurfer_b200.Membrane -- drop-in for memsurfer.Membrane on the per-frame hot path.

Mirrors memsurfer/membrane.py:27-345: same constructor kwargs, attributes, method names, return
dtypes and memoisation.  compute_pnormals / compute_properties / compute_density(-ies) run on the
B200 through the C ABI; the surface PRODUCER (Poisson reconstruction + CGAL parameterisation /
projection / Delaunay, membrane.py:160-234) stays a host step: here `compute_membrane_surface`
is a documented stand-in (planar projection + scipy Delaunay) because pypoisson/CGAL are not
available in this image; real producers can inject their arrays with `set_membrane_surface`.
"""
import logging

import numpy as np
import torch

from . import _lib
from .trimesh import TriMesh

LOGGER = logging.getLogger(__name__)


class PointNormalsInfo(dict):
    """info of point_normals_device: `m` (points incl. ghost images), `n_unoriented`, `knn_ties`, `mst_rounds`.  The
    numbers live in an 8 x int64 status block on the device (include/memsurfer_b200.h, MS_INFO_*) whose download
    was queued behind the kernels; the first look at any key waits for it -- the chain itself never does."""

    def __init__(self, pending):
        super().__init__()
        self._pending = pending

    def _resolve(self):
        if self._pending is not None:
            h = self._pending.result()
            self._pending = None
            dict.update(self, m=int(h[_lib.INFO_M]), n_unoriented=int(h[_lib.INFO_UNORIENTED]),
                        knn_ties=int(h[_lib.INFO_KNN_TIES]), mst_rounds=int(h[_lib.INFO_MST_ROUNDS]),
                        knn_candidates=int(h[_lib.INFO_KNN_CANDS]))
            if int(h[_lib.INFO_ERROR]) != 0:
                raise _lib.MemSurferCudaError('point normals: non-finite point coordinates')
        return self

    def __getitem__(self, k):
        return dict.__getitem__(self._resolve(), k)

    def get(self, k, default=None):
        return dict.get(self._resolve(), k, default)

    def __contains__(self, k):
        return dict.__contains__(self._resolve(), k)

    def __repr__(self):
        return dict.__repr__(self._resolve())


class _InfoOnDevice(object):
    """Stand-in for a download in flight: the status block stays on the device until somebody looks at it (used while
    a CUDA graph is being captured, where a copy on a side stream would leave the capture unjoined)."""

    def __init__(self, info):
        self.info = info

    def result(self):
        return self.info.cpu().numpy()


def point_normals_device(points_dev, bbox, periodic, blayer, knbrs, ndir_hint=+1, keep=None, download_info=True):
    """Device pipeline of Membrane.compute_pnormals (membrane.py:124-157).

    points_dev: (n,3) float32 CUDA tensor.  Returns (pnormals f32 (n,3), pnormals f64 (n,3), info)
    with info = PointNormalsInfo(m=#points incl. ghosts, n_unoriented=..., ...).  Nothing is waited for: the
    five C-ABI calls are queued back to back on the current stream (the number of ghost images stays on the
    device; buffers are sized by ms_pointset_capacity).  If `keep` is a dict the intermediate device buffers
    (points_all, ghost_src, nbr, normals_all, normals_pca) are stored in it for tests (this does synchronise).
    """
    _lib.require_cuda()
    L = _lib.lib()
    dev = points_dev.device
    st = _lib.stream_ptr(dev)
    n = int(points_dev.shape[0])
    k1 = int(knbrs) + 1
    cap = int(L.ms_pointset_capacity(n, int(bool(periodic))))
    pts_all = torch.empty((cap, 3), dtype=torch.float32, device=dev)
    gsrc = torch.empty(cap, dtype=torch.int32, device=dev)
    info = torch.empty(_lib.INFO_WORDS, dtype=torch.int64, device=dev)
    ws = _lib.workspace(max(L.ms_ghosts_workspace_bytes(n), L.ms_knn_workspace_bytes(cap, k1), L.ms_orient_workspace_bytes(cap, k1)), dev)
    box = _lib.box6(np.asarray(bbox, np.float32).reshape(-1)) if periodic else None
    _lib.check(L.ms_ghosts(points_dev.data_ptr(), n, int(bool(periodic)), box, float(blayer) if periodic else 0.0,
                           pts_all.data_ptr(), gsrc.data_ptr(), info.data_ptr(), ws.data_ptr(), ws.numel(), st), 'ms_ghosts')
    nbr = torch.empty((cap, k1), dtype=torch.int32, device=dev)
    _lib.check(L.ms_knn(pts_all.data_ptr(), cap, info.data_ptr(), k1, nbr.data_ptr(), ws.data_ptr(), ws.numel(), st), 'ms_knn')
    nrm = torch.empty((cap, 3), dtype=torch.float64, device=dev)
    _lib.check(L.ms_pca_normals(pts_all.data_ptr(), cap, info.data_ptr(), nbr.data_ptr(), k1, nrm.data_ptr(), st), 'ms_pca_normals')
    if keep is not None:
        keep['normals_pca'] = nrm.clone()
    _lib.check(L.ms_orient_normals(pts_all.data_ptr(), cap, info.data_ptr(), nbr.data_ptr(), k1, nrm.data_ptr(),
                                   ws.data_ptr(), ws.numel(), st), 'ms_orient_normals')
    o32 = torch.empty((n, 3), dtype=torch.float32, device=dev)
    o64 = torch.empty((n, 3), dtype=torch.float64, device=dev)
    _lib.check(L.ms_finish_pnormals(nrm.data_ptr(), n, int(ndir_hint), o32.data_ptr(), o64.data_ptr(),
                                    ws.data_ptr(), ws.numel(), st), 'ms_finish_pnormals')
    pinfo = PointNormalsInfo(_lib.d2h_async(info) if download_info else _InfoOnDevice(info))
    if keep is not None:
        m = pinfo['m']
        keep.update(points_all=pts_all[:m], ghost_src=gsrc[:m], nbr=nbr[:m], normals_all=nrm[:m], normals_pca=keep['normals_pca'][:m])
    return o32, o64, pinfo


def _info_block(m, dev):
    return torch.tensor([int(m)] + [0] * (_lib.INFO_WORDS - 1), dtype=torch.int64, device=dev)


def knn_device(points_dev, k1):
    """(m,k1) int32 neighbour lists of an (m,3) float32 CUDA tensor through ms_knn, plus the number of queries
    that met an exact distance tie (diagnostic).  Stage 2 of point_normals_device on its own."""
    L = _lib.lib()
    dev = points_dev.device
    st = _lib.stream_ptr(dev)
    m = int(points_dev.shape[0])
    nbr = torch.empty((m, int(k1)), dtype=torch.int32, device=dev)
    info = _info_block(m, dev)
    ws = _lib.workspace(L.ms_knn_workspace_bytes(m, int(k1)), dev)
    _lib.check(L.ms_knn(points_dev.data_ptr(), m, info.data_ptr(), int(k1), nbr.data_ptr(), ws.data_ptr(), ws.numel(), st), 'ms_knn')
    return nbr, int(info[_lib.INFO_KNN_TIES].item())


def orient_normals_device(points_dev, nbr_dev, normals_dev, phase_times=None):
    """MST orientation (stage 4) of given (m,3) float64 normals in place; returns the number of points that are
    not connected to the source.  phase_times: a list that receives the microseconds of every phase of the kernel."""
    import ctypes as C
    L = _lib.lib()
    dev = points_dev.device
    m, k1 = int(nbr_dev.shape[0]), int(nbr_dev.shape[1])
    info = _info_block(m, dev)
    ws = _lib.workspace(L.ms_orient_workspace_bytes(m, k1), dev)
    _lib.check(L.ms_orient_normals(points_dev.data_ptr(), m, info.data_ptr(), nbr_dev.data_ptr(), k1, normals_dev.data_ptr(),
                                   ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)), 'ms_orient_normals')
    if phase_times is not None:
        buf, n = (C.c_double * 64)(), C.c_int(0)
        _lib.check(L.ms_orient_phase_times(ws.data_ptr(), ws.numel(), m, k1, buf, 64, C.byref(n), _lib.stream_ptr(dev)), 'ms_orient_phase_times')
        phase_times.extend(buf[i] * 1e-3 for i in range(n.value))
    return int(info[_lib.INFO_UNORIENTED].item())


_label_pool = None


def _label_table_of(lab, pinned):
    """(distinct values, per-point index into them, counts, the index as int32 [page-locked if `pinned`]) of a label
    array.  Pure numpy that releases the GIL: Membrane.__init__ starts it on a worker thread, so that the ~2 ms it takes
    for 500 000 points are spent while the point normals are being queued, not in front of the first KDE launch."""
    if lab.dtype.kind in 'iu' and lab.size > 0 and lab.min() >= 0 and lab.max() < 65536:
        # small non-negative integers: counting is much cheaper than the sort inside np.unique
        counts = np.bincount(lab)
        present = counts > 0
        values = np.flatnonzero(present).astype(lab.dtype)
        inverse = lab if present.all() else (np.cumsum(present) - 1)[lab]      # 0..L-1 all in use: the labels are their own index
        counts = counts[present]
    else:
        values, inverse, counts = np.unique(lab, return_inverse=True, return_counts=True)
    index32 = _lib.pinned_empty(inverse.shape, np.int32) if pinned else np.empty(inverse.shape, np.int32)
    np.copyto(index32, inverse, casting='unsafe')
    return values, inverse, counts, index32


def _label_table_async(lab):
    global _label_pool
    if _label_pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _label_pool = ThreadPoolExecutor(2, thread_name_prefix="ms_labels")
    return _label_pool.submit(_label_table_of, lab, True)


class Membrane(object):
    """Create and analyse membrane surfaces (GPU hot path, host surface producer)."""

    # KDE cut-off in sigmas used by compute_density / compute_densities.  The reference sums all
    # V x n_ids pairs (TriMesh_kde.cpp:344-392); the mass of a 2-D Gaussian beyond c sigma is
    # exp(-c^2/2) of the whole: 6.1e-13 at c = 7.5 -- three orders under the 1e-9 relative parity bar
    # AND under the 1e-12 absolute bar for the O(1) values of normalised densities (SURVEY 8(d) accepts
    # 7 sigma = 2.3e-11), so the default skips those pairs through a cell list (the C ABI falls back
    # to all pairs when the window does not fit the box).  Set to 0 for the reference's full sum.
    KDE_CUTOFF_SIGMAS = 7.5

    def __init__(self, points, **kwargs):
        points = np.asarray(points)
        if points.ndim != 2 or points.shape[1] != 3:
            raise ValueError(f'Membrane needs 3D points: got {points.shape}')
        self.points = points
        self.npoints = self.points.shape[0]

        self.periodic = kwargs.get('periodic', False)
        if self.periodic:
            self.blayer = kwargs.get('boundary_layer', 0.2)
            if 'bbox' not in kwargs:
                raise ValueError('Periodic membrane needs 3D bounding box: ndarray (2 ,3)')
        if 'bbox' in kwargs:
            self.bbox = np.asarray(kwargs['bbox'])
            if self.bbox.shape[0] not in [2, 3]:
                raise ValueError('Membrane needs 3D bounding box: ndarray (2 ,3)')
        else:
            # the reference fails with AttributeError at membrane.py:80 when bbox is missing;
            # a non-periodic membrane simply does not need one here
            self.bbox = np.stack([points.min(axis=0), points.max(axis=0)])

        if 'labels' in kwargs and kwargs['labels'] is not None:
            self.labels = np.asarray(kwargs['labels'])
            if self.labels.shape[0] != self.npoints or len(self.labels.shape) > 1:
                raise ValueError('Membrane expects one label per point')
        else:
            self.labels = None

        self.device = torch.device(kwargs.get('device', 'cuda:0'))
        # float32 copy as membrane.py:79, made in page-locked memory so that the upload is a DMA.  An input that
        # already is float32 and page-locked (`memsurfer_b200.pinned`) is used where it is: it is never written
        # to (fit_points_to_box_xy copies first if it has to move a point).
        self.points, self._points_shared = _lib.staged(self.points, np.float32)
        self.bbox = self.bbox.astype(np.float32)
        # the reference copies the points into its PointSet HERE (membrane.py:81), i.e. before any
        # later fit_points_to_box_xy(): normals are computed on these coordinates.  The PointSet copy
        # lives in HBM: the upload starts now and self.points is not touched before it has landed.
        self._pset_dev, self._pset_event, self._stream, self._overlap = None, None, None, None
        self._outside, self._points_at_upload = None, True      # self.points still equals what was uploaded
        self._pset_nbytes = self.points.nbytes
        self._label_future = None
        if torch.cuda.is_available():
            self._stream = _lib.next_stream(self.device)
            self._upload_and_check(kwargs.get('check_overlaps', True))
            if self.labels is not None:
                self._label_future = _label_table_async(self.labels)

        self._pnormals = None            # float32 numpy once fetched
        self._pnormals_pending = None    # its download in flight (_lib.Pending)
        self._pnormals_f64 = None
        self._pnormals_f64_dev = None
        self.spoints = None
        self.ppoints = None
        self.surf_poisson = None
        self.memb_planar = None
        self.memb_smooth = None
        self.memb_exact = None
        self.attributes = _lib.LazyResults()      # numpy on access; downloads may still be in flight
        self.properties = {}
        self.info = {}

    @_lib.home_stream
    def _upload_and_check(self, check_overlaps):
        self._pset_dev = _lib.h2d(self.points, self.device)
        self._pset_event = torch.cuda.Event()
        self._pset_event.record(torch.cuda.current_stream(self.device))
        # fit_points_to_box_xy (membrane.py:100-121) walks the points four times on the host; the device counts the
        # coordinates outside the box right behind the upload, and the host only walks them if there are any
        self._outside = None
        if self.periodic:
            cnt = torch.empty(1, dtype=torch.int64, device=self.device)
            _lib.check(_lib.lib().ms_points_outside_box(self._pset_dev.data_ptr(), self.npoints, _lib.box6(self.bbox.reshape(-1)),
                                                        cnt.data_ptr(), _lib.stream_ptr(self.device)), 'ms_points_outside_box')
            self._outside = _lib.d2h_async(cnt)
            # the count stands for self.points only while nobody writes to the array: hand out a read-only view
            # (fit_points_to_box_xy swaps in a writable copy when it has to move a point)
            self.points = self.points.view()
            self.points.flags.writeable = False
        # membrane.py:84 -- warn about coincident input points (grid-based on the GPU instead of pdist).  The
        # search is only queued here (cell list over the given box, no read-back); the verdict is fetched at
        # the first point where this membrane synchronises anyway (compute_pnormals) or by check_overlaps().
        self._overlap = None
        if check_overlaps:
            from .utils import OverlapCheck
            b = self.bbox
            self._overlap = OverlapCheck(self._pset_dev, (b[0, 0], b[0, 1], b[1, 0], b[1, 1]))

    def check_overlaps(self):
        """True if no two input points are closer than 1e-5 (the reference's constructor check, membrane.py:84)."""
        return True if self._overlap is None else self._overlap.report()

    def _adopt(self, mesh):
        """The meshes this membrane drives run on its stream."""
        if mesh is not None and mesh._stream is None:
            mesh._stream = self._stream
        return mesh

    @property
    def pnormals(self):
        """(npoints,3) float32 point normals, None before compute_pnormals (memsurfer/membrane.py:62,141).  The
        kernels and the download are only QUEUED by Membrane.compute(); the first look at the array waits."""
        if self._pnormals is None and self._pnormals_pending is not None:
            self._pnormals = self._pnormals_pending.result()
            self._pnormals_pending = None
            self._after_pnormals()
        return self._pnormals

    @pnormals.setter
    def pnormals(self, value):
        self._pnormals, self._pnormals_pending = value, None

    def _after_pnormals(self):
        info = self.info.get('pnormals')
        if info is not None and info['n_unoriented'] > 0:
            # the reference silently returns a permuted array here (PointSet.cpp:176-193)
            LOGGER.warning(f"PointSet::need_normals failed for {info['n_unoriented']} points: "
                           "Some normals may be incorrectly oriented!")
        self.check_overlaps()
        if self._pnormals.shape != (self.npoints, 3):
            raise ValueError(f'PointSet computed incorrect normals (got {self._pnormals.size} values')

    @property
    def pnormals_f64(self):
        """fp64 twin of `pnormals` (not part of the reference API; fetched from the device on first use)."""
        if self._pnormals_f64 is None and self._pnormals_f64_dev is not None:
            with torch.cuda.stream(self._stream):
                self._pnormals_f64 = _lib.d2h(self._pnormals_f64_dev)
            self._pnormals_f64_dev = None
        return self._pnormals_f64

    # ---------------------------------------------------------------------- host helpers
    def fit_points_to_box_xy(self):
        """Wrap x,y into the box (membrane.py:100-121); returns the number of adjusted coordinates."""
        nadjusted = 0
        boxw = self.bbox[1, :] - self.bbox[0, :]
        outside = getattr(self, '_outside', None)
        if outside is not None and self._points_at_upload:
            if int(outside.result()[0]) == 0:      # counted on the device right behind the upload
                return 0
        self._points_at_upload = False
        if self._points_shared or not self.points.flags.writeable:      # the caller's own array / the read-only view
            self.points, self._points_shared = self.points.copy(), False
        for d in range(2):
            lo = self.points[:, d] < self.bbox[0, d]
            self.points[lo, d] += boxw[d]
            hi = self.points[:, d] > self.bbox[1, d]
            self.points[hi, d] -= boxw[d]
            nadjusted += int(lo.sum()) + int(hi.sum())
        return nadjusted

    # ---------------------------------------------------------------------- point normals (GPU)
    @_lib.home_stream
    def _queue_pnormals(self, knbrs=18, ndir_hint=+1):
        """Queue ghosts -> k-NN -> PCA -> MST orientation -> float32 hand-off and the download of the result on
        this membrane's stream; returns at once."""
        if self._pnormals is not None or self._pnormals_pending is not None:
            return
        _lib.require_cuda()
        o32, o64, info = point_normals_device(self._pset_dev, self.bbox, self.periodic,
                                              getattr(self, 'blayer', 0.2), knbrs, ndir_hint)
        self.info['pnormals'] = info
        self._pnormals_pending = _lib.d2h_async(o32)
        self._pnormals_f64_dev, self._pnormals_f64 = o64, None      # fp64 twin stays in HBM until asked for

    def compute_pnormals(self, knbrs=18, ndir_hint=+1, wait=True):
        """(npoints,3) float32 normals by (k+1)-NN PCA + MST orientation (membrane.py:124-157).  wait=False (not in
        the reference) only queues the kernels and the download and returns None; `self.pnormals` waits when read."""
        self._queue_pnormals(knbrs, ndir_hint)
        return self.pnormals if wait else None

    # ---------------------------------------------------------------------- surface producer (host)
    def compute_approx_surface(self, exactness_level=10):
        """Reference: screened Poisson reconstruction via pypoisson (membrane.py:160-183), a host
        producer that is unavailable in this image.  Kept as an explicit no-op stage: the point
        normals it would consume are computed, the smooth surface falls back to the points."""
        self._queue_pnormals()
        if self.surf_poisson is None:
            LOGGER.warning('compute_approx_surface: screened Poisson reconstruction (pypoisson) is a HOST producer that is '
                           'not available here; assign a TriMesh to Membrane.surf_poisson (or use set_membrane_surface) '
                           'to get a smoothed surface -- otherwise memb_smooth is built on the raw points')

    @_lib.home_stream
    def set_membrane_surface(self, spoints, ppoints, planar_bbox):
        """Inject a producer's output: points projected on the smooth surface and on the plane."""
        self.spoints = np.asarray(spoints, np.float32)
        self.ppoints = np.asarray(ppoints, np.float32)
        self._planar_bbox = (np.asarray(planar_bbox[0], np.float32), np.asarray(planar_bbox[1], np.float32))
        self._build_meshes()

    @_lib.home_stream
    def compute_membrane_surface(self):
        """memsurfer/membrane.py:186-234.  With a smoothing surface in `self.surf_poisson` (a TriMesh: the output of a
        Poisson reconstruction, assigned by the caller -- pypoisson itself is a host producer that is not in this
        image) the reference flow runs as written: parameterize it, project the points on it and on its plane
        (closest-point kernel), planar bbox = the parameterisation's.  Without one the points are their own
        smooth surface: spoints = points, ppoints = their (x, y), planar bbox = the periodic box (logged)."""
        if self.surf_poisson is not None:
            self.surf_poisson.parameterize()
            self.spoints, self.ppoints = self.surf_poisson.project_on_surface_and_plane(self.points)
            bb0, bb1 = self.surf_poisson.parameterization_bbox()
            self._planar_bbox = (np.asarray(bb0, np.float32), np.asarray(bb1, np.float32))
        else:
            LOGGER.warning('compute_membrane_surface: no smoothing surface (Membrane.surf_poisson is None): '
                           'memb_smooth is built on the raw points and equals memb_exact')
            self.spoints = self.points.copy()
            self.ppoints = self.points[:, :2].copy()
            self._planar_bbox = (self.bbox[0, :2].copy(), self.bbox[1, :2].copy())
        # membrane.py:199-200 -- separate coincident projections before triangulating
        from .utils import test_overlapping_points
        test_overlapping_points(self.spoints, fix=True, device=self.device)
        test_overlapping_points(self.ppoints, fix=True, device=self.device)
        self._build_meshes()

    def _build_meshes(self):
        dev = self.device
        self.memb_planar = TriMesh(self.ppoints, periodic=self.periodic, label='memb_planar', device=dev)
        self.memb_smooth = TriMesh(self.spoints, periodic=self.periodic, label='memb_smooth', device=dev)
        self.memb_exact = TriMesh(self.points, periodic=self.periodic, label='memb_exact', device=dev)
        for mesh in (self.memb_planar, self.memb_smooth, self.memb_exact):
            self._adopt(mesh)
        if self.periodic:
            self.memb_planar.set_bbox(*self._planar_bbox)
            self.memb_smooth.set_bbox(self.spoints.min(axis=0), self.spoints.max(axis=0))
            self.memb_exact.set_bbox(self.points.min(axis=0), self.points.max(axis=0))
        self.memb_planar.compute_delaunay()
        self.memb_smooth.copy_triangulation(self.memb_planar)
        self.memb_exact.copy_triangulation(self.memb_planar)
        if self.periodic:
            self.memb_planar.set_bbox(self.ppoints.min(axis=0), self.ppoints.max(axis=0))

    # ---------------------------------------------------------------------- properties / densities (GPU)
    @_lib.home_stream
    def compute_properties(self, mtype='smooth'):
        """Vertex normals, point areas and curvatures of the chosen mesh (membrane.py:237-247).  The results
        land in mesh.attributes; the launch and the downloads are only queued here."""
        assert mtype == 'smooth' or mtype == 'exact'
        mesh = self._adopt(self.memb_exact if mtype == 'exact' else self.memb_smooth)
        mesh._properties_lazy()

    MAX_FUSED_LABELS = 64

    def _label_table(self):
        """Distinct label values and the per-point index into them (computed once -- in the background since the
        constructor when a GPU is present, see _label_table_of)."""
        if getattr(self, '_label_values', None) is None:
            fut, self._label_future = getattr(self, '_label_future', None), None
            table = fut.result() if fut is not None else _label_table_of(self.labels, False)
            self._label_values, self._label_inverse, self._label_counts, self._label_index32 = table
        return self._label_values, self._label_inverse

    @_lib.home_stream
    def compute_density(self, type, sigma, name, get_nlipdis, labels=[], cutoff_sigmas=None):
        """Density of the points carrying one of `labels` (all points if empty) on memb_smooth
        (membrane.py:252-274).  With point labels available the label-fused kernel is used: the first
        field of a given (type, sigma) evaluates every vertex pair once and later fields of other
        labels are only a normalisation pass."""
        if cutoff_sigmas is None:
            cutoff_sigmas = self.KDE_CUTOFF_SIGMAS
        if type < 1 or type > 3:
            raise ValueError(f'Invalid density type, {type}. Should be 1 (geodesic), 2 (2D) or 3 (3D)')
        mesh = self._adopt(self.memb_smooth)
        if len(labels) > 0 and self.labels is None:
            raise ValueError('Cannot compute density of selected labels, because point labels are not available')
        fused = self.labels is not None and type != 1
        if fused:
            values, inverse = self._label_table()
            fused = values.shape[0] <= self.MAX_FUSED_LABELS
        if fused:
            if mesh._label_index is None:
                mesh.set_vertex_labels(inverse, values.shape[0], self._label_counts, index32=self._label_index32)
            select = np.arange(values.shape[0]) if len(labels) == 0 else np.where(np.isin(values, labels))[0]
            if select.shape[0] == 0:
                # none of the requested labels occurs in this leaflet: the reference's np.where(np.in1d(...)) is
                # empty, and an empty id list means ALL vertices (membrane.py:273-274, TriMesh_kde.cpp: ids.empty())
                select = np.arange(values.shape[0])
            self.attributes.adopt(name, mesh.compute_density_labels(type, sigma, name, get_nlipdis, select, cutoff_sigmas, lazy=True))
            return
        if len(labels) == 0:
            self.attributes.adopt(name, mesh.compute_density(type, sigma, name, get_nlipdis, np.empty([0]), cutoff_sigmas, lazy=True))
            return
        lidxs = np.where(np.isin(self.labels, labels))[0]
        self.attributes.adopt(name, mesh.compute_density(type, sigma, name, get_nlipdis, lidxs, cutoff_sigmas, lazy=True))

    # compute_densities is called once per label ('all', then every lipid type: examples/ex_3lipid.py:165-175).  The first
    # call computes the pair sums of every label anyway; with this switch it also writes -- sigma by sigma, while the
    # next sigma is being summed -- the fields of 'all' and of every single label and starts their downloads, so that
    # the later calls find their results on the way to the host instead of queueing them behind the last kernel.
    # Costs (1 + #labels) * #sigmas * 4 V bytes of device->host traffic whether or not the fields are asked for.
    EMIT_LABEL_FIELDS = True

    @_lib.home_stream
    def _prefetch_densities(self, types, sigmas, after=None, get_nlipdis=None):
        """Queue the pair sums of all sigmas; returns an event that marks their end (None if nothing was queued)."""
        if self.labels is None or self.memb_smooth is None or len(sigmas) < 2:
            return None
        values, inverse = self._label_table()
        if values.shape[0] > self.MAX_FUSED_LABELS:
            return None
        mesh = self._adopt(self.memb_smooth)
        if mesh._label_index is None:
            mesh.set_vertex_labels(inverse, values.shape[0], self._label_counts, index32=self._label_index32)
        queued = False
        for t in types:
            if t in (2, 3):
                emit = None
                if self.EMIT_LABEL_FIELDS and get_nlipdis is not None:
                    nl = values.shape[0]
                    sets = [list(range(nl))] + ([[k] for k in range(nl)] if nl > 1 else [])
                    tags = ['all'] + ([values[k] for k in range(nl)] if nl > 1 else [])
                    emit = (sets, bool(get_nlipdis),
                            lambda s_, k, t=t, tags=tags: 'density_type{0}_{1}_k{2:.1f}'.format(t, tags[k], s_))
                queued |= bool(mesh.density_labels_prefetch(t, sigmas, self.KDE_CUTOFF_SIGMAS, after=after, emit=emit))
        if not queued:
            return None
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(mesh.device))
        return ev

    @staticmethod
    def compute_thickness(a, b, mtype='smooth'):
        """Leaflet thickness: distance of every vertex of one membrane's surface to the other's
        (membrane.py:277-286 -> TriMesh::distance_to_other_mesh); stored in .properties['thickness']."""
        assert mtype in ['smooth', 'exact']
        for m in (a, b):
            m._adopt(m.memb_exact); m._adopt(m.memb_smooth)
        if mtype == 'exact':
            a.properties['thickness'] = a.memb_exact.compute_distance_to_surface(b.memb_exact)
            b.properties['thickness'] = b.memb_exact.compute_distance_to_surface(a.memb_exact)
        else:
            a.properties['thickness'] = a.memb_smooth.compute_distance_to_surface(b.memb_smooth)
            b.properties['thickness'] = b.memb_smooth.compute_distance_to_surface(a.memb_smooth)

    def _fused_select(self, labels):
        """Label indices (into the distinct label values) selected by `labels`, or None when the label-fused KDE
        does not apply (no point labels, or too many distinct ones)."""
        if self.labels is None:
            return None
        values, inverse = self._label_table()
        if values.shape[0] > self.MAX_FUSED_LABELS:
            return None
        select = np.arange(values.shape[0]) if len(labels) == 0 else np.where(np.isin(values, labels))[0]
        if select.shape[0] == 0:           # absent label: empty id list = all vertices (membrane.py:273-274)
            select = np.arange(values.shape[0])
        return select

    @_lib.home_stream
    def _compute_densities_fused(self, t, sigmas, names, get_nlipdis, labels):
        """All sigmas of one (type, label set) through one multi-field launch; False if the fused path does not apply."""
        if t not in (2, 3) or self.memb_smooth is None:
            return False
        select = self._fused_select(labels)
        if select is None:
            return False
        mesh = self._adopt(self.memb_smooth)
        if mesh._label_index is None:
            mesh.set_vertex_labels(self._label_table()[1], self._label_values.shape[0], self._label_counts, index32=self._label_index32)
        vals = mesh.compute_density_labels_multi(t, sigmas, names, get_nlipdis, select, self.KDE_CUTOFF_SIGMAS)
        for n_, v in zip(names, vals):
            self.attributes.adopt(n_, v)
        return True

    @staticmethod
    def compute_densities(membranes, types, sigmas, get_nlipdis, l):
        labels = [] if l == 'all' else [l]
        # label-fused path: all sigmas of a type on one binning of the vertices.  The membranes of a call run on their
        # own streams; their pair sums are chained (each waits for the previous one's) so that the leaflets FINISH one
        # after the other: the result downloads of the first then run under the kernels of the second instead of all
        # downloads piling up behind the last kernel of the frame (the persistent pair kernels fill the GPU either way).
        done = {}                      # per device
        for m in membranes:
            dkey = str(getattr(m, 'device', None))
            ev = m._prefetch_densities(types, sigmas, after=done.get(dkey), get_nlipdis=get_nlipdis)
            if ev is not None:
                done[dkey] = ev
        for t in types:
            names = ['density_type{0}_{1}_k{2:.1f}'.format(t, l, s) for s in sigmas]
            for m in membranes:
                if len(labels) > 0 and m.labels is None:
                    raise ValueError('Cannot compute density of selected labels, because point labels are not available')
                if m._compute_densities_fused(t, sigmas, names, get_nlipdis, labels):
                    continue
                for s, name in zip(sigmas, names):       # the reference's loop (membrane.py:289-297)
                    m.compute_density(t, s, name, get_nlipdis, labels)

    @staticmethod
    def compute(positions, labels, bbox, periodic):
        knbrs = 18
        m = Membrane(positions, labels=labels, periodic=periodic, bbox=bbox)
        m.fit_points_to_box_xy()
        m._queue_pnormals(knbrs)            # compute_pnormals without waiting for the result (pnormals is lazy)
        m.compute_approx_surface()
        m.compute_membrane_surface()
        return m

    # ---------------------------------------------------------------------- export (host)
    def write_all(self, outprefix, params={}):
        """VTP files of the points and of every surface, as membrane.py:319-345 writes them
        (`_points`, `_surface_poisson` when a Poisson surface exists, `_planar`, `_membrane_exact`,
        `_membrane_smooth`); written by memsurfer_b200.export, VTK is not required."""
        params = dict(params)
        pparams = {'normals': self.pnormals}
        if self.labels is not None and self.labels.shape != (0, 0):
            pparams['labels'] = self.labels
        if 'shells' in self.memb_smooth.attributes:
            pparams['shells'] = self.memb_smooth.attributes['shells']
        TriMesh(self.points, device=self.device).write_vtp(outprefix + '_points.vtp', pparams)
        if self.surf_poisson is not None:
            self.surf_poisson.write_vtp(outprefix + '_surface_poisson.vtp')
        for key in list(self.attributes.keys()):
            params[key] = self.attributes[key]
        if 'labels' in pparams.keys():
            params['labels'] = pparams['labels']
        self.memb_planar.write_vtp(outprefix + '_planar.vtp', params)
        self.memb_exact.write_vtp(outprefix + '_membrane_exact.vtp', params)
        self.memb_smooth.write_vtp(outprefix + '_membrane_smooth.vtp', params)
