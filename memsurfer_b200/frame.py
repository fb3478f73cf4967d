"""Per-frame driver of the hot path on device-resident inputs, and frame sharding across GPUs.

One *frame*  is, for each leaflet: `compute_pnormals(k=18)`
(memsurfer/membrane.py:124-157), `compute_properties` (membrane.py:237-247) and
`compute_densities(types=[2], sigmas, get_nlipdis=True)` for label 'all' and every lipid type
(membrane.py:289-297, examples/ex_3lipid.py:169-175).  `LeafletFrame` keeps the inputs of one
leaflet in HBM and issues exactly those C-ABI calls (include/memsurfer_b200.h) on the current
stream; nothing is copied back except the small per-field summary records.

Independent frames share no state (examples/ex_3lipid.py:142-182), so a trajectory is sharded as
frame f -> rank f mod world_size with no data-path collective; ranks exchange only the summary
records (one all_gather of a few hundred bytes per frame; NCCL on GPUs, gloo in the CPU tests).
"""
import numpy as np
import torch

from . import _lib
from .membrane import point_normals_device

SUMMARY_FIELDS = ("parea", "mean_curv", "gaus_curv")


def shard_frames(nframes, rank, world_size):
    """Frame ids owned by `rank`: round-robin f = rank, rank + W, ... (contract of section 8(e))."""
    return list(range(int(rank), int(nframes), int(world_size)))


class LeafletFrame(object):
    """Device-resident working set of one leaflet of one frame."""

    def __init__(self, points, labels, bbox, mesh, device="cuda:0", blayer=0.2):
        """points (N,3) f32, labels (N,), bbox (2,3): the Membrane inputs; mesh: a TriMesh whose
        (periodic) triangulation is already set by the host producer."""
        _lib.require_cuda()
        self.device = torch.device(device)
        self.bbox = np.asarray(bbox, np.float32)
        self.blayer = float(blayer)
        self.periodic = bool(mesh.periodic)
        self.mesh = mesh
        self.n = int(points.shape[0])
        verts, faces = mesh._concat_host()
        self.nv, self.nv_all, self.nf = int(mesh.vertices.shape[0]), int(verts.shape[0]), int(faces.shape[0])
        self.h2d_bytes = 0
        self.points = self._up(np.ascontiguousarray(points, np.float32))
        self.verts = self._up(verts)
        self.faces = self._up(faces.view(np.int32))
        self.kverts = self.verts[:self.nv]
        self.half_box = mesh._half_box()
        self.bboxl = None if mesh.bboxl is None else np.asarray(mesh.bboxl, np.float32)
        self.label_ids = {}
        self.label_values, self.label_index, self.label_counts = None, None, None
        if labels is not None:
            labels = np.asarray(labels)
            self.label_values, inv = np.unique(labels, return_inverse=True)
            self.label_counts = np.bincount(inv, minlength=self.label_values.shape[0])
            self.label_index = self._up(np.ascontiguousarray(inv, np.int32))
            for l in self.label_values:
                ids = np.where(labels == l)[0].astype(np.int32)
                self.label_ids[l.item() if hasattr(l, "item") else l] = self._up(ids)
        dev = self.device
        L = _lib.lib()
        nv, nf = self.nv, self.nf
        self.csr_off = torch.empty(nv + 1, dtype=torch.int32, device=dev)
        self.csr_adj = torch.empty(max(3 * nf, 1), dtype=torch.int32, device=dev)
        self.keep = torch.empty(max(nf, 1), dtype=torch.uint8, device=dev)
        self.topo_ws = _lib.workspace(L.ms_mesh_topology_workspace_bytes(nv, nf), dev)
        self.prop_ws = _lib.workspace(L.ms_mesh_properties_workspace_bytes(nv, nf), dev)
        self.vn64 = torch.empty((nv, 3), dtype=torch.float64, device=dev)
        self.vn32 = torch.empty((nv, 3), dtype=torch.float32, device=dev)
        self.pa64 = torch.empty(nv, dtype=torch.float64, device=dev)
        self.pa32 = torch.empty(nv, dtype=torch.float32, device=dev)
        self.H = torch.empty(nv, dtype=torch.float64, device=dev)
        self.K = torch.empty(nv, dtype=torch.float64, device=dev)
        nl = 1 if self.label_values is None else int(self.label_values.shape[0])
        self.kde_ws = _lib.workspace(max(L.ms_kde_workspace_bytes(nv, 0), L.ms_kde_labels_workspace_bytes(nv, nl)), dev)
        self.kde_sel_ws = _lib.workspace(L.ms_kde_select_workspace_bytes(nv), dev)
        self.kde_raw = torch.empty((nl, nv), dtype=torch.float64, device=dev)
        self.sum_ws = _lib.workspace(L.ms_field_summary_workspace_bytes(24), dev)
        self.densities = {}
        self._field_blocks = {}
        self.pnormals32 = self.pnormals64 = None
        self.info = {}

    def _up(self, arr):
        self.h2d_bytes += arr.nbytes
        return torch.from_numpy(arr).to(self.device, non_blocking=True)

    # ------------------------------------------------------------------ the three stages
    def run_pnormals(self, knbrs=18, ndir_hint=+1):
        self.pnormals32, self.pnormals64, self.info = point_normals_device(
            self.points, self.bbox, self.periodic, self.blayer, knbrs, ndir_hint,
            download_info=not torch.cuda.is_current_stream_capturing())

    def run_properties(self):
        L = _lib.lib()
        st = _lib.stream_ptr(self.device)
        xw, yw = self.half_box
        _lib.check(L.ms_mesh_topology(self.verts.data_ptr(), self.nv_all, self.nv, self.faces.data_ptr(), self.nf,
                                      xw, yw, self.csr_off.data_ptr(), self.csr_adj.data_ptr(), self.keep.data_ptr(),
                                      self.topo_ws.data_ptr(), self.topo_ws.numel(), st), 'ms_mesh_topology')
        _lib.check(L.ms_mesh_properties(self.verts.data_ptr(), self.nv_all, self.nv, self.faces.data_ptr(), self.nf,
                                        self.csr_off.data_ptr(), self.csr_adj.data_ptr(), self.keep.data_ptr(),
                                        self.vn64.data_ptr(), self.vn32.data_ptr(), self.pa64.data_ptr(),
                                        self.pa32.data_ptr(), self.H.data_ptr(), self.K.data_ptr(),
                                        self.prop_ws.data_ptr(), self.prop_ws.numel(), st),
                   'ms_mesh_properties')

    def run_density(self, name, type, sigma, get_nlipids, ids_dev=None, cutoff_sigmas=0.0):
        L = _lib.lib()
        nids = 0 if ids_dev is None else int(ids_dev.numel())
        out = self.densities.get(name)
        if out is None:
            out = self.densities[name] = (torch.empty(self.nv, dtype=torch.float64, device=self.device),
                                          torch.empty(self.nv, dtype=torch.float32, device=self.device))
        box = self._box(cutoff_sigmas)
        _lib.check(L.ms_kde(self.kverts.data_ptr(), self.nv, _lib.ptr(ids_dev) if nids else None, nids, int(type),
                            int(self.periodic), box, float(sigma), int(bool(get_nlipids)), float(cutoff_sigmas),
                            out[0].data_ptr(), out[1].data_ptr(), self.kde_ws.data_ptr(), self.kde_ws.numel(),
                            _lib.stream_ptr(self.device)), 'ms_kde')
        return out

    def run_densities(self, sigmas, types=(2,), get_nlipids=True, cutoff_sigmas=0.0, fused=True):
        """'all' + one field per label, named as Membrane.compute_densities names them.  fused: one
        ms_kde_labels pass per (type, sigma) + one ms_kde_select per field; else one ms_kde per field."""
        L = _lib.lib()
        st = _lib.stream_ptr(self.device)
        fused = fused and self.label_index is not None and self.label_values.shape[0] <= 64
        raws = {}
        if fused and len(sigmas) > 1 and len(types) > 1:        # checked before anything is launched
            raise NotImplementedError('run_densities: one density type per call when sigmas share a binning')
        if fused and len(sigmas) > 1:      # all sigmas of a type on one binning of the vertices
            nl = int(self.label_values.shape[0])
            need = len(sigmas) * nl * self.nv
            if getattr(self, 'kde_raw_multi', None) is None or self.kde_raw_multi.numel() < need:
                self.kde_raw_multi = torch.empty(need, dtype=torch.float64, device=self.device)
            import ctypes as C
            sg = (C.c_float * len(sigmas))(*[float(x) for x in sigmas])
            for t in types:
                _lib.check(L.ms_kde_labels_multi(self.kverts.data_ptr(), self.nv, self.label_index.data_ptr(), nl, int(t),
                                                 int(self.periodic), self._box(cutoff_sigmas), sg, len(sigmas), float(cutoff_sigmas),
                                                 self.kde_raw_multi.data_ptr(), self.kde_ws.data_ptr(), self.kde_ws.numel(), st),
                           'ms_kde_labels_multi')
                for i, s in enumerate(sigmas):
                    raws[(t, float(s))] = self.kde_raw_multi[i * nl * self.nv:(i + 1) * nl * self.nv]
        import ctypes as C
        for t in types:
            if not fused:
                for s in sigmas:
                    self.run_density('density_type{0}_{1}_k{2:.1f}'.format(t, 'all', s), t, s, get_nlipids, None, cutoff_sigmas)
                    for l, ids in self.label_ids.items():
                        self.run_density('density_type{0}_{1}_k{2:.1f}'.format(t, l, s), t, s, get_nlipids, ids, cutoff_sigmas)
                continue
            nl = int(self.label_values.shape[0])
            if (t, float(sigmas[0])) in raws:
                raw_all = self.kde_raw_multi                     # (nsigma, nl, nv) from the shared binning
            else:                                                # one ms_kde_labels pass per sigma into the same layout
                need = len(sigmas) * nl * self.nv
                if getattr(self, 'kde_raw_multi', None) is None or self.kde_raw_multi.numel() < need:
                    self.kde_raw_multi = torch.empty(need, dtype=torch.float64, device=self.device)
                raw_all = self.kde_raw_multi
                for i, s in enumerate(sigmas):
                    _lib.check(L.ms_kde_labels(self.kverts.data_ptr(), self.nv, self.label_index.data_ptr(), nl, int(t),
                                               int(self.periodic), self._box(cutoff_sigmas), float(s), float(cutoff_sigmas),
                                               raw_all.data_ptr() + 8 * i * nl * self.nv, self.kde_ws.data_ptr(), self.kde_ws.numel(), st),
                               'ms_kde_labels')
            # every field of this type in two launches: per-row totals, then all normalised fields
            nrows = len(sigmas) * nl
            tot_words = int(L.ms_kde_label_totals_bytes(nrows)) // 8
            if getattr(self, 'kde_totals', None) is None or self.kde_totals.numel() < tot_words:
                self.kde_totals = torch.empty(tot_words, dtype=torch.float64, device=self.device)
            _lib.check(L.ms_kde_label_totals(raw_all.data_ptr(), nrows, self.nv, self.kde_totals.data_ptr(), st), 'ms_kde_label_totals')
            names, row0, masks, nsels = [], [], [], []
            for i, s in enumerate(sigmas):
                for li, l in enumerate(['all'] + [v.item() for v in self.label_values]):
                    names.append('density_type{0}_{1}_k{2:.1f}'.format(t, l, s))
                    row0.append(i * nl)
                    masks.append((1 << nl) - 1 if li == 0 else 1 << (li - 1))
                    nsels.append(float(self.label_counts.sum() if li == 0 else self.label_counts[li - 1]))
            for b in range(0, len(names), 64):
                nb_ = min(64, len(names) - b)
                blk = self._field_blocks.get((t, b))
                if blk is None or blk[0].shape[0] != nb_:
                    blk = self._field_blocks[(t, b)] = (torch.empty((nb_, self.nv), dtype=torch.float64, device=self.device),
                                                        torch.empty((nb_, self.nv), dtype=torch.float32, device=self.device))
                _lib.check(L.ms_kde_fields(raw_all.data_ptr(), self.kde_totals.data_ptr(), self.nv, nl, nb_,
                                           (C.c_int32 * nb_)(*row0[b:b + nb_]), (C.c_uint64 * nb_)(*masks[b:b + nb_]),
                                           (C.c_double * nb_)(*nsels[b:b + nb_]), int(bool(get_nlipids)),
                                           blk[0].data_ptr(), blk[1].data_ptr(), st), 'ms_kde_fields')
                for k in range(nb_):
                    self.densities[names[b + k]] = (blk[0][k], blk[1][k])

    def run(self, sigmas, knbrs=18, cutoff_sigmas=0.0, fused=True):
        self.run_pnormals(knbrs)
        self.run_properties()
        self.run_densities(sigmas, cutoff_sigmas=cutoff_sigmas, fused=fused)

    # ------------------------------------------------------------------ the frame as one CUDA graph
    def capture(self, sigmas, knbrs=18, cutoff_sigmas=0.0):
        """Record run() + summary() of this leaflet as ONE CUDA graph (the chain has no host round trip and every
        buffer keeps its address) and return it; replay() then costs one launch instead of ~46.  Valid while the
        shapes, the box and the sigmas stay what they were at capture time -- new coordinates of the same system
        (an NVT trajectory) are copied into self.points / self.verts before a replay; a changed box means a new
        capture.  The first call runs the frame once eagerly (workspaces, function attributes)."""
        L = _lib.lib()
        self.run(sigmas, knbrs, cutoff_sigmas)
        self.summary()
        torch.cuda.current_stream(self.device).synchronize()
        g = torch.cuda.CUDAGraph()
        L.ms_launch_count(1)
        with torch.cuda.graph(g):
            self.run(sigmas, knbrs, cutoff_sigmas)
            rec = self.summary()
        self.graph, self.graph_summary, self.graph_launches = g, rec, int(L.ms_launch_count(1))
        return g

    def replay(self):
        """Launch the captured frame on the current stream; returns the (F,4) summary tensor (overwritten by the next replay)."""
        self.graph.replay()
        return self.graph_summary

    def _box(self, cutoff_sigmas):
        if self.periodic:
            b = self.bboxl
            if b.shape[0] == 4:
                b = np.array([b[0], b[1], 0, b[2], b[3], 0], np.float32)
            return _lib.box6(b)
        if cutoff_sigmas > 0:
            if getattr(self, '_extent', None) is None:           # once per frame object: two passes over the vertices
                v = self.mesh.vertices
                lo, hi = v.min(axis=0), v.max(axis=0)
                self._extent = np.array([lo[0], lo[1], 0, hi[0], hi[1], 0], np.float32)
            return _lib.box6(self._extent)
        return None

    def evaluated_pairs(self, sigmas, types=(2,), cutoff_sigmas=0.0, fused=True, split=False):
        """Pairs the KDE kernels actually evaluate in run_densities (diagnostic; synchronises).  split=True: (all pairs,
        pairs summed in fp64) -- the cut-off kernel sums the pairs beyond 5 sigma in fp32."""
        import ctypes as C
        L = _lib.lib()
        total = near = 0.0
        out = (C.c_double * 2)()
        fused = fused and self.label_index is not None and self.label_values.shape[0] <= 64
        if fused and len(sigmas) > 1:          # the sigmas share one binning (ms_kde_labels_multi)
            sg = (C.c_float * len(sigmas))(*[float(x) for x in sigmas])
            per = (C.c_double * (2 * len(sigmas)))()
            for t in types:
                _lib.check(L.ms_kde_labels_multi_pairs(self.kverts.data_ptr(), self.nv, self.label_index.data_ptr(),
                                                       int(self.label_values.shape[0]), int(t), int(self.periodic), self._box(cutoff_sigmas),
                                                       sg, len(sigmas), float(cutoff_sigmas), per, self.kde_ws.data_ptr(),
                                                       self.kde_ws.numel(), _lib.stream_ptr(self.device)), 'ms_kde_labels_multi_pairs')
                total += sum(per[:len(sigmas)]); near += sum(per[len(sigmas):])
            return (total, near) if split else total
        for t in types:
            for s in sigmas:
                if fused:
                    sets = [(None, 0, self.label_index, int(self.label_values.shape[0]))]
                else:
                    sets = [(None, 0, None, 1)] + [(ids, int(ids.numel()), None, 1) for ids in self.label_ids.values()]
                for ids, nids, lab, nl in sets:
                    _lib.check(L.ms_kde_evaluated_pairs(self.kverts.data_ptr(), self.nv, _lib.ptr(ids) if nids else None, nids,
                                                        _lib.ptr(lab), nl, int(t), int(self.periodic), self._box(cutoff_sigmas),
                                                        float(s), float(cutoff_sigmas), out, self.kde_ws.data_ptr(),
                                                        self.kde_ws.numel(), _lib.stream_ptr(self.device)),
                               'ms_kde_evaluated_pairs')
                    total += out[0]; near += out[1]
        return (total, near) if split else total

    def pairs_per_run(self, sigmas, types=(2,)):
        """All-pairs-equivalent KDE pair count of run_densities."""
        per_sigma = self.nv * self.nv + sum(self.nv * int(i.numel()) for i in self.label_ids.values())
        return per_sigma * len(sigmas) * len(types)

    # ------------------------------------------------------------------ summary record
    def summary(self):
        """(F,4) f64 device tensor: {sum, sum|x|, min, max} of parea, mean_curv, gaus_curv and every
        density field (sorted by name)."""
        import ctypes as C
        L = _lib.lib()
        st = _lib.stream_ptr(self.device)
        fields = [self.pa64, self.H, self.K] + [self.densities[k][0] for k in sorted(self.densities)]
        nf = len(fields)
        out = torch.empty((nf, 4), dtype=torch.float64, device=self.device)
        need = L.ms_field_summary_workspace_bytes(nf)
        if self.sum_ws.numel() < need:
            self.sum_ws = _lib.workspace(need, self.device)
        ptrs = (C.c_void_p * nf)(*[f.data_ptr() for f in fields])
        _lib.check(L.ms_field_summary(ptrs, nf, self.nv, out.data_ptr(), self.sum_ws.data_ptr(), self.sum_ws.numel(), st),
                   'ms_field_summary')
        return out

    def summary_names(self):
        return list(SUMMARY_FIELDS) + sorted(self.densities)


def gather_summaries(local, group=None):
    """all_gather of one rank's summary tensor (any shape, same on every rank) -> (W, ...) tensor.
    With torch.distributed not initialised this is the identity (W = 1)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return local.unsqueeze(0)
    W = dist.get_world_size(group)
    out = [torch.empty_like(local) for _ in range(W)]
    dist.all_gather(out, local.contiguous(), group=group)
    return torch.stack(out)
