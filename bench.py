#!/usr/bin/env python
"""bench.py -- membrane frames/sec of the per-frame hot path (normals + curvature + KDE).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One step = one frame: for each of the 2 leaflets of a synthetic periodic
undulating bilayer (BASELINE.json configs[2]: 1M lipids, ~500k points per leaflet)
compute_pnormals(k=18), compute_properties (vertex normals, point areas, mean + Gaussian
curvature) and compute_densities(types=[2], sigmas=[10,20,30,40], get_nlipdis=True) for 'all' and
each of the 3 lipid types.  The mesh producer (periodic Delaunay + trim) runs once in set-up and is
not part of the timed region; it is reported beside it (`mesh_producer`).

 * `value`: frames/s with the inputs resident in HBM (C-ABI calls on device pointers).
 * `e2e`:   the same frame through the drop-in Membrane/TriMesh API from page-locked HOST arrays:
            host->device copies of points/mesh/labels and device->host copies of every result inside
            the timing.
 * `roofline`: the dominant kernel (KDE pair-sum) against the fp64 issue peak measured in the same
            process by a DFMA probe; `roofline_geometry` (mesh properties), `roofline_knn` and
            `roofline_orient` against the measured HBM copy bandwidth (MEASURED_PEAKS.json).
 * `parity`: max errors of this run's own outputs against the CPU oracle (outside the timed region).
 * `cpu_baseline`: the reference's own C++ (oracle/_ref, single-threaded as shipped) for vertex
            normals, point areas and KDE, the fp64 restatement for the CGAL/VTK parts, timed on
            this box's host cores on a bounded sample (see `sample`).
 * `allpairs`: the same frame with the reference's literal all-pairs KDE sum (like-for-like with the CPU arm).
 * `trajectory`: BASELINE configs[3] (100k-lipid, 8-type frames sharded over the ranks; strong scaling).
N > 1: every rank runs the same pipeline on its own frame (weak scaling, frames sharded across
GPUs); the only collective is the all_gather of the per-frame summary records.  Ranks pin
themselves to their own host cores before any thread pool starts.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SIGMAS = [10.0, 20.0, 30.0, 40.0]
FRACTIONS = [0.4, 0.3, 0.3]
METRIC = "membrane_frames_per_sec"
# fp64 FLOP per evaluated 2-D periodic pair / algorithmic bytes per unit
FLOP_PER_PAIR_2D = 32.0
FP64_INSTR_PER_PAIR = {"kde_pairs": 14, "kde_groups": 10}
GEOM_BYTES_PER_VERTEX = 84.0
KNN_BYTES_PER_POINT = (36.0, 112.0)             # without / with the neighbour indices written out
ORIENT_BYTES_PER_POINT = lambda k1: 4.0 * k1 + 24.0 + 24.0      # neighbour row + normal in + normal out
MESH_KERNELS = ("face_records", "edge_mean", "vertex_gather", "pad_verts", "corner_terms", "vertex_gather8")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--lipids", type=int, default=1000000, help="lipids per frame (both leaflets)")
    ap.add_argument("--kde-cutoff", type=float, default=7.5,
                    help="KDE cut-off in sigmas (the drop-in's default 7.5: dropped tail <= exp(-28.1) = 6.1e-13 relative; "
                         "0 = all pairs, the reference's literal sum)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--split-streams", action="store_true",
                    help="experiment: the point-normal chain of a leaflet on a stream of its own, beside its mesh + KDE work")
    ap.add_argument("--no-graphs", action="store_true", help="trajectory arm: queue every frame launch by launch instead of one CUDA graph per leaflet")
    ap.add_argument("--no-extras", action="store_true", help="skip the all-pairs, trajectory, type-3 and producer side measurements")
    ap.add_argument("--trajectory-frames", type=int, default=512, help="total frames of the configs[3] measurement (all ranks)")
    ap.add_argument("--no-pin", action="store_true", help="do not bind the rank to its own host cores")
    ap.add_argument("--single-stream", action="store_true", help="device arm: the two leaflets of a frame one after the other "
                                                                 "(default: each on its own stream)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------- host placement
def pin_rank(local_rank, local_world):
    """Bind this process to its own slice of the host cores (before numpy/torch start thread pools).  Returns
    (cores, numa node of the first core)."""
    try:
        avail = sorted(os.sched_getaffinity(0))
        per = max(1, len(avail) // max(1, local_world))
        mine = avail[local_rank * per:(local_rank + 1) * per] or avail
        os.sched_setaffinity(0, mine)
    except Exception:
        return None, None
    node = None
    try:
        base = "/sys/devices/system/node"
        for d in sorted(os.listdir(base)):
            if not d.startswith("node"):
                continue
            cpus = set()
            for part in open(os.path.join(base, d, "cpulist")).read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            if mine[0] in cpus:
                node = int(d[4:])
                break
    except Exception:
        pass
    return mine, node


# ---------------------------------------------------------------------------------------------- inputs
def make_leaflet_inputs(npoints, seed, z0, ntypes, fractions, phase_shift, device):
    """One leaflet + its periodic mesh in the reference layout.  device=None: host producer (scipy stand-in for
    CGAL); else the GPU periodic Delaunay (TriMesh.compute_delaunay) -- either way outside the timed frame."""
    import numpy as np
    from memsurfer_b200 import synth
    pts, labels, bbox = synth.make_leaflet(npoints, seed=seed, z0=z0, ntypes=ntypes, fractions=fractions, phase_shift=phase_shift)
    lo, hi = pts.min(axis=0), pts.max(axis=0)
    if device is None:
        dfaces, doffs = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
        tri = synth.trim_periodic(pts, dfaces, doffs, lo, hi)
    else:
        from memsurfer_b200 import TriMesh
        planar = TriMesh(pts[:, :2], periodic=True, label="memb_planar", device=device)
        planar.set_bbox(bbox[0, :2], bbox[1, :2])
        planar.compute_delaunay()
        m3 = TriMesh(pts, periodic=True, label="memb_smooth", device=device)
        m3.set_bbox(lo, hi)                                   # the producer's bbox (spoints min/max, membrane.py:214-217)
        m3.copy_triangulation(planar)
        tri = dict(faces=m3.faces, periodic_faces=m3._periodic_faces, trimmed_faces=m3._trimmed_faces,
                   duplicated_vidxs=m3._duplicated_vidxs, duplicated_verts=m3._duplicated_verts)
    return dict(points=pts, labels=labels, bbox=bbox, tri=tri, lo=lo, hi=hi)


def make_frame_inputs(lipids, frame, device=None, ntypes=3, fractions=FRACTIONS):
    """Two leaflets of frame `frame` + their periodic meshes."""
    return [make_leaflet_inputs(lipids // 2, 1234 + li, z0, ntypes, fractions, 0.05 * frame, device)
            for li, z0 in enumerate((20.0, -20.0))]


def pin_inputs(lf):
    """The frame's input arrays in page-locked host memory (the e2e arm uploads from there by DMA)."""
    import numpy as np
    from memsurfer_b200 import pinned
    out = dict(lf)
    out["points"] = pinned(lf["points"], np.float32)
    t = lf["tri"]
    out["tri"] = dict(faces=pinned(t["faces"], np.uint32), periodic_faces=t["periodic_faces"],
                      trimmed_faces=pinned(t["trimmed_faces"], np.uint32), duplicated_vidxs=t["duplicated_vidxs"],
                      duplicated_verts=pinned(t["duplicated_verts"], np.float32))
    return out


def new_mesh(lf, device):
    from memsurfer_b200 import TriMesh
    t = lf["tri"]
    mesh = TriMesh(lf["points"], periodic=True, label="memb_smooth", device=device)
    mesh.set_bbox(lf["lo"], lf["hi"])
    mesh.set_periodic_triangulation(t["faces"], t["periodic_faces"], t["trimmed_faces"],
                                    t["duplicated_vidxs"], t["duplicated_verts"])
    return mesh


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler(object):
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        import numpy as np
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------- CPU baseline
def cpu_reference_baseline(lf, kde_sample_verts=16384, all_threads=False, keep=None):
    """Reference CPU path on a bounded sample of the frame workload -> (frames/s, details).  keep: dict that
    receives the oracle's full-size outputs (for the parity key)."""
    import numpy as np
    from oracle import oracle as O
    import ctypes
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        ncores = len(os.sched_getaffinity(0)) or 1
        gomp.omp_set_num_threads(ncores if all_threads else 1)
    except OSError:
        ncores = 1
    pts, bbox, tri, labels = lf["points"], lf["bbox"], lf["tri"], lf["labels"]
    V = pts.shape[0]
    verts = np.ascontiguousarray(np.concatenate([pts, tri["duplicated_verts"]]), np.float32)
    faces = np.ascontiguousarray(np.concatenate([tri["faces"], tri["trimmed_faces"]]), np.uint32)
    R = O.ref()
    t = {}
    t0 = time.perf_counter(); pn = O.point_normals(pts, bbox, 0.2, 18, +1); t["pnormals_port"] = time.perf_counter() - t0
    if R is not None:
        t0 = time.perf_counter(); O.ref_need_normals(verts, faces); O.ref_need_pointareas(verts, faces)
        t["normals_areas_reference"] = time.perf_counter() - t0
    else:
        t0 = time.perf_counter(); O.mesh_normals(verts, faces); O.pointareas(verts, faces)
        t["normals_areas_port"] = time.perf_counter() - t0
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    xw, yw = float(np.float32(0.5) * (bb[3] - bb[0])), float(np.float32(0.5) * (bb[4] - bb[1]))
    t0 = time.perf_counter(); H, K = O.curvatures(verts, faces, (xw, yw)); t["curvatures_port"] = time.perf_counter() - t0
    if keep is not None:
        keep.update(pnormals=pn, H=H[:V], K=K[:V])
    # KDE: the reference evaluates all V x n_ids pairs; sample = a compact patch of the same leaflet
    ns = int(min(kde_sample_verts, V))
    side = int(round(np.sqrt(V)))
    s_side = int(np.sqrt(ns))
    idx = (np.arange(s_side)[:, None] * side + np.arange(s_side)[None, :]).reshape(-1)
    sub = np.ascontiguousarray(pts[idx]); ns = sub.shape[0]
    sbb = np.concatenate([sub.min(0), sub.max(0)]).astype(np.float32)
    t0 = time.perf_counter()
    if R is not None:
        O.ref_kde(sub, np.zeros(0, np.int32), 2, 20.0, True, sbb)
        kind = "reference"
    else:
        O.kde(sub, np.zeros(0, np.int32), 2, 20.0, True, sbb)
        kind = "port"
    t["kde_sample"] = time.perf_counter() - t0
    pairs_per_s = ns * ns / t["kde_sample"]
    nl = [int((labels == l).sum()) for l in np.unique(labels)]
    pairs_leaflet = len(SIGMAS) * (V * V + sum(V * n for n in nl))
    t_nokde = t["pnormals_port"] + t.get("normals_areas_reference", t.get("normals_areas_port", 0.0)) + t["curvatures_port"]
    t_leaflet = t_nokde + pairs_leaflet / pairs_per_s
    fps = 1.0 / (2.0 * t_leaflet)
    detail = {"seconds": {k: round(v, 4) for k, v in t.items()}, "kde_pairs_per_s": pairs_per_s,
              "kde_pairs_per_leaflet": pairs_leaflet, "extrapolated_frame_seconds": 2.0 * t_leaflet,
              "measured_non_kde_seconds_per_frame": 2.0 * t_nokde,
              "note": "the KDE term (99.9 % of the frame) is EXTRAPOLATED linearly in pairs from the patch; the survey host "
                      "measured 5.3e7 pairs/s/core for the same code (BASELINE.md section 2) -- different CPU"}
    sample = (f"one leaflet (V={V}): point normals [fp64 port of CGAL pca+mst], vertex normals + point areas "
              f"[{'reference C++' if R is not None else 'port'}], curvatures [fp64 port of vtkCurvatures] at full size; "
              f"KDE [{kind}] on a {ns}-vertex patch, type 2, sigma 20, all pairs ({ns * ns:.3g} pairs), extrapolated "
              f"linearly in pairs to the frame's {2 * pairs_leaflet:.3g} pairs; frame = 2 leaflets")
    return fps, kind, (ncores if all_threads else 1), sample, detail


# ---------------------------------------------------------------------------------------------- reference arm
def run_reference_arm(args, rank):
    import numpy as np
    if rank != 0:
        return
    leaflets = make_frame_inputs(args.lipids, 0)[:1]
    vals = []
    for i in range(args.warmup + args.steps):
        # each step = one bounded sample (smaller KDE patch during warm-up)
        fps, kind, cores, sample, detail = cpu_reference_baseline(leaflets[0], 4096 if i < args.warmup else 12288)
        if i >= args.warmup:
            vals.append(fps)
    v = float(np.mean(vals))
    V = leaflets[0]["points"].shape[0]
    cfg = workload_config(args, V)
    cfg["kde"] = "all pairs (the reference's literal sum), EXTRAPOLATED from a 12 288-vertex patch"
    cfg["mesh_producer"] = "host (scipy Delaunay stand-in for CGAL), untimed"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (reference) / f64 (port)", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample,
                             "detail": detail},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "vertices_per_sec": v * 2 * V}
    print(json.dumps(line), flush=True)


def workload_config(args, V):
    return {"workload": f"configs[2]: synthetic periodic undulating bilayer, {2 * V} lipids (2 leaflets x {V} points), "
                        f"1 frame/step: pnormals(k=18) + vertex normals/areas/mean+Gauss curvature + KDE type 2, "
                        f"sigmas {SIGMAS}, labels all + {len(FRACTIONS)} types, get_nlipdis",
            "kde": "all pairs" if args.kde_cutoff <= 0 else f"cut-off {args.kde_cutoff} sigma (cell list)",
            "l2": "flushed between steps (512 MiB device write)",
            "parallelism": f"frames sharded over {args.gpus} GPU(s), 1 frame per rank per step; the two leaflets of a frame on "
                           + ("one stream" if args.single_stream else "two streams"),
            "mesh_producer": "GPU periodic Delaunay + trim (TriMesh.compute_delaunay), untimed"}


# ---------------------------------------------------------------------------------------------- our arm
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    cores, numa = (None, None) if args.no_pin else pin_rank(local_rank, local_world)

    import numpy as np
    import torch
    import torch.distributed as dist
    from memsurfer_b200 import _lib, Membrane, TriMesh
    from memsurfer_b200.frame import LeafletFrame, gather_summaries
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    device = torch.device(f"cuda:{local_rank}")
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    L = _lib.lib()
    Membrane.KDE_CUTOFF_SIGMAS = args.kde_cutoff
    sms = torch.cuda.get_device_properties(device).multi_processor_count

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def gather_floats(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device=device)
        if world == 1:
            return [float(x)]
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    # frame `rank` of the trajectory: every rank owns one frame per step (weak scaling)
    t_setup = time.perf_counter()
    leaflets = [pin_inputs(lf) for lf in make_frame_inputs(args.lipids, rank, device=device)]
    t_setup = time.perf_counter() - t_setup
    V = leaflets[0]["points"].shape[0]
    frames = [LeafletFrame(lf["points"], lf["labels"], lf["bbox"], new_mesh(lf, device), device=device) for lf in leaflets]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=device)

    side_streams = None if args.single_stream else [torch.cuda.Stream(device=device) for _ in frames]
    normal_streams = [torch.cuda.Stream(device=device) for _ in frames] if args.split_streams else None
    use_streams = [True]

    def step_local(cutoff=None):
        # the two leaflets of a frame share nothing (examples/ex_3lipid.py:165-166): with --leaflet-streams each runs on
        # its own stream (forked from and joined back into the timing stream), else one after the other
        recs = []
        main = torch.cuda.current_stream(device)
        for i, fr in enumerate(frames):
            if side_streams is None or not use_streams[0]:
                fr.run(SIGMAS, 18, args.kde_cutoff if cutoff is None else cutoff)
                recs.append(fr.summary())
                continue
            side_streams[i].wait_stream(main)
            with torch.cuda.stream(side_streams[i]):
                if normal_streams is not None:
                    normal_streams[i].wait_stream(side_streams[i])
                    with torch.cuda.stream(normal_streams[i]):
                        fr.run_pnormals(18)
                    fr.run_properties()
                    fr.run_densities(SIGMAS, cutoff_sigmas=args.kde_cutoff if cutoff is None else cutoff)
                    side_streams[i].wait_stream(normal_streams[i])
                else:
                    fr.run(SIGMAS, 18, args.kde_cutoff if cutoff is None else cutoff)
                recs.append(fr.summary())
                recs[-1].record_stream(main)
        if side_streams is not None and use_streams[0]:
            for st_ in side_streams:
                main.wait_stream(st_)
        return torch.stack(recs)

    def step_device():
        rec = step_local()
        return gather_summaries(rec) if world > 1 else rec.unsqueeze(0)

    def timed_local(fn, nwarm, nsteps):
        """This rank only (no collective inside): side measurements of rank 0."""
        for _ in range(nwarm):
            fn()
        total_ms = 0.0
        for _ in range(nsteps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            total_ms += e0.elapsed_time(e1)
        return total_ms

    def timed(fn, nwarm, nsteps, profile=False):
        for _ in range(nwarm):
            fn()
        barrier()
        if profile:
            L.ms_profile_enable(1)
        L.ms_launch_count(1)
        total_ms = 0.0
        for _ in range(nsteps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            total_ms += e0.elapsed_time(e1)
        launches = int(L.ms_launch_count(1))
        prof = _lib.profile_report() if profile else {}
        L.ms_profile_enable(0)
        barrier()
        per_rank = gather_floats(total_ms)
        return max(per_rank), launches, prof, per_rank

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    # nvidia-smi takes a few hundred ms to come up (NVML init stalls kernel submission while it does): keep
    # warming up until its first sample has arrived, so that the timed steps run next to a settled sampler
    # (step_device ends in a collective when world > 1: rank 0 decides, every rank follows)
    t_wait = time.perf_counter()
    while True:
        go = torch.tensor([int(rank == 0 and clocks.proc is not None and not clocks.lines
                               and time.perf_counter() - t_wait < 3.0)], dtype=torch.int32, device=device)
        if world > 1:
            dist.broadcast(go, 0)
        if int(go.item()) == 0:
            break
        step_device()
        torch.cuda.synchronize()
    total_ms, launches, _, per_rank_ms = timed(step_device, args.warmup, args.steps, profile=False)
    clk = clocks.stop() if rank == 0 else None
    ms_per_step = total_ms / args.steps
    value = world * 1e3 / ms_per_step
    # second, shorter pass with an event around every kernel launch: the per-kernel times behind the rooflines
    psteps = max(1, min(args.steps, 5))
    use_streams[0] = False                      # one stream: the per-kernel events must follow each other
    ptotal_ms, _, prof, _ = timed(step_device, 1, psteps, profile=True)
    use_streams[0] = True
    pinfo = [dict(m=fr.info['m'], cands=fr.info['knn_candidates'], rounds=fr.info['mst_rounds']) for fr in frames]

    # ---- end to end through the drop-in API, from page-locked host arrays
    e2e, per_rank_e2e = None, None
    if not args.no_e2e:
        bytes_io = {"h2d": 0, "d2h": 0, "queue_ms": []}

        def step_e2e():
            # one frame in the order of the reference's own driver (examples/ex_3lipid.py:165-175): build BOTH
            # membranes from host arrays (Membrane.compute: point normals + surface; the surface is the producer's
            # output, injected), then properties and compute_densities([top, bottom], ...) for 'all' and every
            # lipid type; finally EVERY result is read back as numpy (downloads inside the timing)
            h2d = d2h = 0
            ms = []
            t_host = time.perf_counter()
            for lf in leaflets:
                m = Membrane(lf["points"], labels=lf["labels"], periodic=True, bbox=lf["bbox"], device=device)
                m.compute_pnormals(18, wait=False)              # queued, like Membrane.compute() does
                m.fit_points_to_box_xy()                        # membrane.py:305 (nothing to move here; the check is host work)
                h2d += m._pset_nbytes
                m.memb_smooth = new_mesh(lf, device)
                ms.append(m)
            for m, lf in zip(ms, leaflets):
                m.compute_properties('smooth')
                h2d += sum(a.nbytes for a in m.memb_smooth._dev['_host'])
                h2d += 4 * lf["labels"].shape[0]                # the per-vertex label index
            for l in ['all'] + list(range(len(FRACTIONS))):
                Membrane.compute_densities(ms, [2], SIGMAS, True, l)
            bytes_io["queue_ms"].append((time.perf_counter() - t_host) * 1e3)  # host time to queue the frame (no waiting in it)
            recs = []
            for m in ms:
                mesh = m.memb_smooth
                d2h += m.pnormals.nbytes
                d2h += sum(mesh.attributes[k].nbytes for k in ('pnormal', 'parea', 'mean_curv', 'gaus_curv'))
                d2h += sum(v.nbytes for v in m.attributes.values())
                recs.append([float(mesh.attributes['parea'].sum()), float(mesh.attributes['mean_curv'].mean())])
            bytes_io["h2d"], bytes_io["d2h"] = h2d, d2h
            return recs

        esteps = max(1, min(args.steps, 10))
        e2e_ms, _, _, per_rank_e2e = timed(step_e2e, 2, esteps)
        e2e_ms /= esteps
        per_rank_e2e = [x / esteps for x in per_rank_e2e]
        per_rank_queue = gather_floats(float(np.median(bytes_io["queue_ms"][-esteps:])))
        e2e = {"value": world * 1e3 / e2e_ms, "unit": "frames/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": bytes_io["h2d"], "d2h_bytes_per_step": bytes_io["d2h"],
               "host_queue_ms_per_rank": per_rank_queue,
               "api": "memsurfer_b200.Membrane / TriMesh (page-locked host numpy in, numpy out)"}

    # ---- side measurements (every rank takes part in the trajectory; the rest is rank 0 only)
    extras = {}
    if not args.no_extras:
        extras["trajectory"] = measure_trajectory(args, rank, world, device, barrier, gather_floats)
    rank_cores = gather_floats(-1 if cores is None else cores[0])
    rank_ncores = gather_floats(0 if cores is None else len(cores))
    rank_numa = gather_floats(-1 if numa is None else numa)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines (rank 0): DFMA probe for the fp64 peak, measured HBM peak from MEASURED_PEAKS.json
    import ctypes
    probe = torch.empty(sms * 8 * 256, dtype=torch.float64, device=device)
    flop = ctypes.c_double(0)
    st = _lib.stream_ptr(device)
    best = 0.0
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); _lib.check(L.ms_dfma_probe(probe.data_ptr(), 200000, ctypes.byref(flop), st), "probe"); e1.record()
        torch.cuda.synchronize()
        best = max(best, flop.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    pairs_allpairs_equiv = sum(fr.pairs_per_run(SIGMAS) for fr in frames)
    both = [fr.evaluated_pairs(SIGMAS, cutoff_sigmas=args.kde_cutoff, split=True) for fr in frames]
    pairs_step, pairs_fp64_step = sum(b[0] for b in both), sum(b[1] for b in both)
    kname = max(("kde_groups", "kde_pairs"), key=lambda k: prof.get(k, {"ms": 0.0})["ms"])
    kp = prof.get(kname, {"launches": 0, "ms": 0.0})
    roof = None
    if kp["launches"]:
        # fp64-pipe instructions per pair, counted in the SASS of the 2-D kernels (profiles/*sass*):
        # k_kde_pairs periodic 7 DADD + 4 DFMA + 2 DSETP + 1 DMUL = 14; k_kde_groups 5 DADD + 4 DFMA + 1 DMUL = 10
        instr = FP64_INSTR_PER_PAIR[kname]
        pairs_per_launch = pairs_step * psteps / kp["launches"]
        fp64_pairs_per_launch = pairs_fp64_step * psteps / kp["launches"]
        avg_ms = kp["ms"] / kp["launches"]
        pps = pairs_per_launch / (avg_ms * 1e-3)
        # evaluated pairs/s against the fp64 peak.  A pair counts as the 10 fp64-pipe instructions the
        # fp64 path spends on it, whichever path evaluated it: `frac` is throughput in fp64-equivalent work,
        # `fp64_pipe_frac` is what really went through the fp64 pipe.
        ach = pps * instr * 2.0 / 1e12
        ach64 = fp64_pairs_per_launch / (avg_ms * 1e-3) * instr * 2.0 / 1e12
        roof = {"kernel": "k_" + kname, "bound": "fp64_fma", "achieved": ach, "peak": best, "unit": "TFLOP/s",
                "frac": ach / best if best > 0 else None, "traffic": traffic.get(kname),
                "definition": "evaluated pairs/s x 10 fp64-pipe instructions per pair x 2 (every DADD/DMUL/DSETP/DFMA holds one "
                              "FMA slot of the fp64 pipe) / DFMA-probe peak.  Sources at least 4.5 sigma from a group's box are "
                              "summed in fp32 (MUFU.EX2, error <= 2e-10 of a density) in the same launch: they count as evaluated "
                              "pairs in `achieved` (fp64-equivalent work), and `fp64_pipe_frac` counts only the pairs that went "
                              "through the fp64 pipe",
                "fp64_pipe_achieved": ach64, "fp64_pipe_frac": ach64 / best if best > 0 else None,
                "fp64_pairs_per_launch": fp64_pairs_per_launch, "fp32_far_field_pairs_per_launch": pairs_per_launch - fp64_pairs_per_launch,
                "peak_source": "ms_dfma_probe in this process (no fp64 figure in MEASURED_PEAKS.json)",
                "fp64_instr_per_pair": instr, "pairs_per_launch": pairs_per_launch, "avg_launch_ms": avg_ms,
                "launches": kp["launches"], "evaluated_pairs_per_s": pps,
                "allpairs_equivalent_pairs_per_s": pairs_allpairs_equiv * psteps / (kp["ms"] * 1e-3),
                "tflops_at_32_flop_per_pair_convention": pps * FLOP_PER_PAIR_2D / 1e12,
                "share_of_step": kp["ms"] / ptotal_ms}
    mp_ms = sum(prof.get(k, {"ms": 0.0})["ms"] for k in MESH_KERNELS)
    mp_n = max([prof.get(k, {"launches": 0})["launches"] for k in MESH_KERNELS] + [0])
    roof_g = None
    if mp_n:
        avg_ms = mp_ms / mp_n
        ach = V * GEOM_BYTES_PER_VERTEX / (avg_ms * 1e-3) / 1e9
        roof_g = {"kernel": "ms_mesh_properties = " + " + ".join("k_" + k for k in MESH_KERNELS if k in prof), "bound": "hbm",
                  "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                  "traffic": traffic.get("mesh_properties"), "peak_source": hbm_src,
                  "bytes_per_vertex": GEOM_BYTES_PER_VERTEX, "avg_ms_per_mesh": avg_ms, "meshes": mp_n,
                  "share_of_step": mp_ms / ptotal_ms,
                  "note": "algorithmic bytes per SURVEY 8(d); the passes are bound by the fp64 pipe (sqrt/div/atan2, "
                          "~1 250 fp64 instructions per face) and L2 gathers, not by HBM (DESIGN.md 4.2)"}

    def hbm_record(names, units, bytes_per_unit, what, extra=None, traffic_key=None):
        ms_ = sum(prof.get(k, {"ms": 0.0})["ms"] for k in names)
        n_ = max([prof.get(k, {"launches": 0})["launches"] for k in names] + [0])
        if not n_:
            return None
        calls = n_ if len(names) == 1 else prof.get(names[-1], {"launches": n_})["launches"]
        avg = ms_ / calls
        ach = units * bytes_per_unit / (avg * 1e-3) / 1e9
        r = {"kernel": " + ".join("k_" + k for k in names if k in prof), "bound": "hbm", "achieved": ach, "peak": hbm_peak,
             "unit": "GB/s", "frac": ach / hbm_peak, "traffic": traffic.get(traffic_key) if traffic_key else None, "peak_source": hbm_src, "bytes_per_unit": bytes_per_unit,
             "units_per_call": units, "avg_ms_per_call": avg, "calls": calls, "share_of_step": ms_ / ptotal_ms, "what": what}
        r.update(extra or {})
        return r

    m_pts = float(np.mean([p["m"] for p in pinfo]))
    knn_names = ("bbox", "make_grid", "cell_count", "cell_scatter", "knn")
    knn_ms = sum(prof.get(k, {"ms": 0.0})["ms"] for k in knn_names)
    roof_knn = hbm_record(knn_names, m_pts, KNN_BYTES_PER_POINT[1],
                          "exact 19-NN of the points incl. ghost images through a uniform cell list (neighbour indices written out)",
                          {"achieved_at_36_bytes_per_point": m_pts * KNN_BYTES_PER_POINT[0] / (knn_ms / max(1, len(frames) * psteps) * 1e-3) / 1e9 if knn_ms else None,
                           "candidate_distances_per_s": sum(p["cands"] for p in pinfo) * psteps / (prof.get("knn", {"ms": 0.0})["ms"] * 1e-3) if prof.get("knn") else None,
                           "candidates_per_query": float(np.mean([p["cands"] / p["m"] for p in pinfo])),
                           "note": "latency / L2-bound in practice (SURVEY 8(d)): report both GB/s and candidate distances/s; traffic = k_knn_thread alone"},
                          traffic_key="knn_thread")
    roof_or = hbm_record(("orient",), m_pts, ORIENT_BYTES_PER_POINT(19),
                         "MST orientation: ONE cooperative kernel per leaflet on HALF the SMs (Boruvka rounds decided on the device); the two leaflets of a frame run side by side on their streams, so avg_ms_per_call is also the wall time per FRAME (DESIGN.md 4.3)",
                         {"boruvka_rounds": [p["rounds"] for p in pinfo],
                          "graph_edges_per_s": None if not prof.get("orient") else 9.7 * m_pts * len(frames) * psteps / (prof["orient"]["ms"] * 1e-3)},
                         traffic_key="orient")

    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, V),
            "vertices_per_sec": value * 2 * V, "gpu_launches": launches, "clocks": clk, "e2e": e2e,
            "roofline": roof, "roofline_geometry": roof_g, "roofline_knn": roof_knn, "roofline_orient": roof_or,
            "per_rank": {"ms_per_step": [x / args.steps for x in per_rank_ms], "e2e_ms_per_step": per_rank_e2e,
                         "first_core": [int(x) for x in rank_cores], "cores": [int(x) for x in rank_ncores],
                         "numa_node": [int(x) for x in rank_numa]},
            "setup_seconds": round(t_setup, 2),
            "kernel_ms_per_step": {k: round(v["ms"] / psteps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}}
    line.update(extras)

    if not args.no_extras:
        # like-for-like with the CPU arm: the reference's literal all-pairs sum on the same frame
        try:
            ap_ms = timed_local(lambda: step_local(0.0), 1, 1)
            line["allpairs"] = {"ms_per_step": ap_ms, "frames_per_s": 1e3 / ap_ms, "kde": "all pairs (--kde-cutoff 0), 1 step, this rank",
                                "pairs_per_step": pairs_allpairs_equiv}
        except Exception as e:
            line["allpairs"] = {"error": repr(e)}
        # type-3 (3-D Gaussian) fields: same frame, same cut-off
        try:
            fr = frames[0]
            def t3():
                fr.run_densities(SIGMAS, types=(3,), cutoff_sigmas=args.kde_cutoff)
            t3_ms = timed_local(t3, 1, 3)
            for k in [k for k in fr.densities if 'type3' in k]:
                del fr.densities[k]
            line["kde_type3"] = {"ms_per_leaflet": t3_ms / 3, "what": "16 type-3 fields of one leaflet (4 sigmas x (all + 3 types)), device-resident"}
        except Exception as e:
            line["kde_type3"] = {"error": repr(e)}
        # the same 16 type-2 fields on a NON-periodic mesh of the same vertices (DistanceKernels.hpp:75-86: plain
        # Euclidean distance, no images; the window is clipped at the data's extent)
        try:
            from memsurfer_b200.frame import LeafletFrame
            lf0 = leaflets[0]
            fmesh = TriMesh(lf0["points"], faces=np.asarray(lf0["tri"]["faces"]), periodic=False, device=device)
            fnp = LeafletFrame(lf0["points"], lf0["labels"], lf0["bbox"], fmesh, device=device)
            def tnp():
                fnp.run_densities(SIGMAS, types=(2,), cutoff_sigmas=args.kde_cutoff)
            np_ms = timed_local(tnp, 1, 3)
            line["kde_nonperiodic"] = {"ms_per_leaflet": np_ms / 3, "what": "16 type-2 fields of one leaflet on a non-periodic mesh, device-resident"}
            del fnp, fmesh
        except Exception as e:
            line["kde_nonperiodic"] = {"error": repr(e)}
        # the mesh producer is NOT part of a frame; reported beside it
        lf = leaflets[0]
        try:
            t_prod = []
            for _ in range(4):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                pm = TriMesh(lf["points"][:, :2], periodic=True, device=device)
                pm.set_bbox(lf["bbox"][0, :2], lf["bbox"][1, :2])
                pm.compute_delaunay()
                t_prod.append((time.perf_counter() - t0) * 1e3)
            line["mesh_producer"] = {"ms_per_leaflet": min(t_prod[1:]), "faces": int(pm.faces.shape[0] + pm._trimmed_faces.shape[0]),
                                     "what": "TriMesh.compute_delaunay: ms_ghosts + ms_knn + ms_delaunay_periodic + trim; not included in value / e2e"}
        except Exception as e:            # a side measurement must never cost the bench line
            line["mesh_producer"] = {"error": repr(e)}

    # ---- parity of this run's own outputs against the oracle (outside every timed region)
    keep = {}
    if world == 1 and not args.no_cpu_baseline:
        fps, kind, ncpu, sample, detail = cpu_reference_baseline(leaflets[0], keep=keep)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": ncpu, "kind": kind, "sample": sample,
                                "detail": detail}
    try:
        line["parity"] = parity_record(frames[0], leaflets[0], step_local, keep)
    except Exception as e:
        line["parity"] = {"error": repr(e)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def parity_record(fr, lf, step_device, keep):
    """Max errors of the frame bench.py just timed (leaflet 0) against the CPU oracle."""
    import numpy as np
    import torch
    from oracle import oracle as O
    step_device()
    torch.cuda.synchronize()
    pts, labels, tri = lf["points"], lf["labels"], lf["tri"]
    V = pts.shape[0]
    verts = np.ascontiguousarray(np.concatenate([pts, tri["duplicated_verts"]]), np.float32)
    faces = np.ascontiguousarray(np.concatenate([tri["faces"], tri["trimmed_faces"]]), np.uint32)
    out = {"against": "oracle/ms_oracle.c (fp64 restatement), leaflet 0 of the timed frame; bars: 1e-9 relative / 1e-12 absolute"}
    out["vertex_normals_max_abs"] = float(np.abs(fr.vn64.cpu().numpy() - O.mesh_normals(verts, faces)[:V]).max())
    out["point_areas_max_abs"] = float(np.abs(fr.pa64.cpu().numpy() - O.pointareas(verts, faces)[:V]).max())
    if "H" in keep:
        H, K = keep["H"], keep["K"]
    else:
        H, K = O.curvatures(verts, faces, fr.half_box)
        H, K = H[:V], K[:V]

    def rel(a, b):
        return float((np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * np.abs(b).max())).max())
    out["mean_curv_max_rel"] = rel(fr.H.cpu().numpy(), H)
    out["gaus_curv_max_rel"] = rel(fr.K.cpu().numpy(), K)
    if "pnormals" in keep:
        out["point_normals_max_abs"] = float(np.abs(fr.pnormals64.cpu().numpy() - keep["pnormals"]).max())
    tg = np.arange(0, V, max(1, V // 256))[:256]
    bb = fr.bboxl
    worst = 0.0
    for s, l in ((10.0, 'all'), (40.0, 1)):
        ids = np.zeros(0, np.int32) if l == 'all' else np.where(labels == l)[0].astype(np.int32)
        got = fr.densities['density_type2_{}_k{:.1f}'.format(l, s)][0].cpu().numpy()
        raw = O.kde_raw_targets(pts, ids, tg, 2, s, bb)
        ratio = got[tg] / raw
        worst = max(worst, float(np.abs(ratio / np.median(ratio) - 1).max()))
        out['density_{}_k{:.0f}_sum_minus_count'.format(l, s)] = float(got.sum() - (ids.shape[0] or V))
    out["density_max_rel_dev_256_targets"] = worst
    return out


def measure_trajectory(args, rank, world, device, barrier, gather_floats):
    """BASELINE configs[3]: frames of a 100k-lipid, 8-type bilayer sharded over the ranks (frame f -> rank f mod W,
    strong scaling: the total number of frames is fixed).  Device-resident inputs; every frame = 2 leaflets through
    LeafletFrame.run + one summary record; the records of all frames are exchanged in ONE all_gather at the end."""
    import numpy as np
    import torch
    from memsurfer_b200 import _lib
    from memsurfer_b200.frame import LeafletFrame, gather_summaries, shard_frames
    try:
        mine = shard_frames(args.trajectory_frames, rank, world)
        ndistinct = min(len(mine), 8)                       # distinct frames kept resident; the loop cycles through them
        sets = []
        for f in mine[:ndistinct]:
            lfs = make_frame_inputs(100000, f, device=device, ntypes=8, fractions=None)
            sets.append([LeafletFrame(lf["points"], lf["labels"], lf["bbox"], new_mesh(lf, device), device=device) for lf in lfs])
        Vt = sets[0][0].nv

        streams = None if args.single_stream else [torch.cuda.Stream(device=device) for _ in range(2)]

        graphs = not args.no_graphs and args.kde_cutoff > 0
        if graphs:                     # one CUDA graph per resident leaflet: a frame is 2 launches instead of 92
            for fs in sets:
                for fr in fs:
                    fr.capture(SIGMAS, 18, args.kde_cutoff)

        def one(fr):
            if graphs:
                return fr.replay().clone()          # the graph writes its summary in place: keep a copy per frame
            fr.run(SIGMAS, 18, args.kde_cutoff)
            return fr.summary()

        def run_all():
            recs = []
            main = torch.cuda.current_stream(device)
            if streams is not None:
                for st_ in streams:
                    st_.wait_stream(main)
            for i in range(len(mine)):
                for li, fr in enumerate(sets[i % ndistinct]):       # leaflet li of every frame on stream li
                    if streams is None:
                        recs.append(one(fr))
                    else:
                        with torch.cuda.stream(streams[li]):
                            recs.append(one(fr))
                            recs[-1].record_stream(main)
            if streams is not None:
                for st_ in streams:
                    main.wait_stream(st_)
            rec = torch.stack(recs)
            e_mid.record()                                  # this rank's frames are queued up to here; then the exchange
            return gather_summaries(rec) if world > 1 else rec

        for fr in sets[0]:
            fr.run(SIGMAS, 18, args.kde_cutoff)             # warm-up (workspaces, function attributes)
        barrier()
        L = _lib.lib()
        L.ms_launch_count(1)
        e0, e1, e_mid = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        t_host = time.perf_counter()
        e0.record(); run_all(); e1.record()
        t_host = (time.perf_counter() - t_host) * 1e3       # host time to QUEUE the rank's frames (launch-bound if ~ ms_total)
        torch.cuda.synchronize()
        launches = int(L.ms_launch_count(1))
        per_rank = gather_floats(e0.elapsed_time(e1))
        per_rank_frames = gather_floats(e0.elapsed_time(e_mid))
        per_rank_host = gather_floats(t_host)
        ms = max(per_rank)
        return {"config": f"configs[3]: {args.trajectory_frames} frames x 100k lipids (2 x {Vt} points), 8 types, sigmas {SIGMAS}, "
                          f"frame f -> rank f mod {world}; {ndistinct} distinct frames resident per rank, cycled",
                "frames": args.trajectory_frames, "frames_per_s": args.trajectory_frames / (ms * 1e-3), "scaling": "strong",
                "ms_total": ms, "ms_per_frame_per_rank": ms / max(1, len(mine)), "per_rank_ms": per_rank,
                "per_rank_frames_ms": per_rank_frames, "per_rank_host_queue_ms": per_rank_host,
                "launches_per_frame": launches / max(1, len(mine)),
                "cuda_graphs": None if not graphs else {"graphs_per_frame": 2, "kernels_per_graph": sets[0][0].graph_launches,
                                                        "what": "LeafletFrame.capture(): one graph per leaflet, replayed per frame"},
                "collective": "one all_gather of the summary records at the end"}
    except Exception as e:
        return {"error": repr(e)}


if __name__ == "__main__":
    main()
